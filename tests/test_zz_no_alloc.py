"""mmimo_welch_psd_noalloc (include/mmimo_b200.h): the allocation-free variant of mmimo_welch_psd whose scratch is a file-scope
`static __device__` array.  No counterpart in the reference.

Called through the C ABI with device pointers (the TEST allocates them with torch; the entry point allocates nothing) and
checked against the existing entry point mmimo_welch_psd on the same signals, on calls in a row - the shared array keeps the
previous call's contents, so the second and third call also prove that the accumulator is cleared.

This file sorts last on purpose (a failure here hides nothing else) and passes both alone in a
fresh process and after the other tests (it creates and destroys its own context on device 0)."""
import ctypes as C

import numpy as np
import pytest

from conftest import rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-5  # the tolerance of the stage tests (tests/test_gpu_dropin_blocks.py): fp32 transforms, fp64 accumulation
MAX_ELEMS = 1 << 19  # MMIMO_WELCH_NOALLOC_MAX_ELEMS


def _signal(seed, n):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal(n) + 1j * rng.standard_normal(n)
    # coloured, with an offset: exercises the detrend and leaves a spectrum that is not flat
    x = np.convolve(x, np.array([1.0, 0.6 - 0.3j, 0.2j]), mode="same") + (0.3 - 0.1j)
    return np.ascontiguousarray(x.astype(np.complex64))


@pytest.fixture()
def env():
    import torch
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import _lib
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    lib = _lib.load()
    h = C.c_void_p()
    _lib.check(lib, lib.mmimo_create(C.byref(h), 0))
    torch.cuda.set_device(0)  # the entry point under test works on the CURRENT device
    yield torch, lib, h
    lib.mmimo_destroy(h)


def _existing(lib, h, x, nperseg, nfft, fs):
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import _lib
    psd = np.empty(nfft, dtype=np.float64)
    freqs = np.empty(nfft, dtype=np.float64)
    fp = lambda a, t: a.ctypes.data_as(C.POINTER(t))
    _lib.check(lib, lib.mmimo_welch_psd(h, x.size, fp(x.view(np.float32), C.c_float), nperseg, nfft, float(fs),
                                        fp(freqs, C.c_double), fp(psd, C.c_double)))
    return psd


def _noalloc(torch, lib, x_dev, n, nperseg, nfft, fs, psd_dev):
    rc = lib.mmimo_welch_psd_noalloc(n, C.c_void_p(x_dev.data_ptr()), nperseg, nfft, float(fs), C.c_void_p(psd_dev.data_ptr()),
                                     C.c_void_p(torch.cuda.current_stream().cuda_stream))
    return rc


def test_noalloc_welch_matches_the_existing_entry_point_on_calls_in_a_row(env):
    torch, lib, h = env
    cases = [(11, 40000, 2048, 4096, 4096.0),   # the (nfft, nperseg) the reference's PSD scripts use most
             (12, 40000, 2048, 4096, 4096.0),   # same shape, other signal: a stale accumulator would show
             (13, 9000, 256, 1024, 30.72e6),    # smaller transform after a larger one: stale rows beyond this call's extent
             (14, 65600, 1024, 4096, 1.0)]      # 127 segments x 4096: one row short of the array's capacity
    # all device buffers are allocated by the TEST before the calls under test
    xs = [_signal(seed, n) for seed, n, _, _, _ in cases]
    x_dev = [torch.from_numpy(x.view(np.float32).copy()).cuda() for x in xs]
    psd_dev = [torch.full((nfft,), -1.0, dtype=torch.float64, device="cuda") for _, _, _, nfft, _ in cases]
    torch.cuda.synchronize()
    free_after_first = None
    for i, (seed, n, nperseg, nfft, fs) in enumerate(cases):
        rc = _noalloc(torch, lib, x_dev[i], n, nperseg, nfft, fs, psd_dev[i])
        assert rc == 0, lib.mmimo_last_error().decode()
        torch.cuda.synchronize()
        free = torch.cuda.mem_get_info()[0]
        if free_after_first is None:
            free_after_first = free      # the first call loads the module (and with it the static array)
        else:
            assert free >= free_after_first, "device memory shrank across an allocation-free call"
    for i, (seed, n, nperseg, nfft, fs) in enumerate(cases):
        got = psd_dev[i].cpu().numpy()
        want = _existing(lib, h, xs[i], nperseg, nfft, fs)
        assert np.all(np.isfinite(got)) and np.all(got >= 0.0)
        assert rel_l2(got, want) < TOL, (i, rel_l2(got, want))
        big = want > want.max() * 1e-5
        assert np.max(np.abs(10 * np.log10(got[big] / want[big]))) < 0.05


def test_noalloc_welch_rejects_sizes_beyond_the_static_array(env):
    torch, lib, h = env
    nfft, nperseg = 4096, 1024
    n = 512 * (MAX_ELEMS // nfft + 2)     # 129 segments > 128 = MAX_ELEMS / nfft
    x_dev = torch.zeros(2 * n, dtype=torch.float32, device="cuda")
    psd_dev = torch.full((nfft,), -1.0, dtype=torch.float64, device="cuda")
    rc = _noalloc(torch, lib, x_dev, n, nperseg, nfft, 1.0, psd_dev)
    assert rc != 0
    msg = lib.mmimo_last_error().decode()
    assert "MMIMO_WELCH_NOALLOC_MAX_ELEMS" in msg and str(MAX_ELEMS) in msg, msg
    torch.cuda.synchronize()
    assert float(psd_dev.min()) == -1.0 and float(psd_dev.max()) == -1.0   # nothing was launched
    # argument errors are reported like the existing entry point's
    assert _noalloc(torch, lib, x_dev, 100, 256, 1024, 1.0, psd_dev) != 0      # shorter than one segment
    assert _noalloc(torch, lib, x_dev, n, 1024, 1000, 1.0, psd_dev) != 0       # nfft not a power of two
    assert lib.mmimo_welch_psd_noalloc(n, None, nperseg, nfft, 1.0, C.c_void_p(psd_dev.data_ptr()), None) != 0
    # one segment fewer fits
    n_ok = 512 * (MAX_ELEMS // nfft + 1)
    assert _noalloc(torch, lib, x_dev, n_ok, nperseg, nfft, 1.0, psd_dev) == 0, lib.mmimo_last_error().decode()
    torch.cuda.synchronize()
    assert float(psd_dev.max()) == 0.0    # an all-zero signal has an all-zero spectrum: the array's old contents are gone
