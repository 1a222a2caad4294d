"""GPU: the drop-in `mp_model.Link` driven exactly like main_mp_miso_mcnc_ber_vs_ebn0.py:101-134 drives the
reference's (one worker, same seeds) must reproduce the reference's shared error / bit counters stored in
tests/golden/link_*.npz - through the fused frame kernel in "numpy" RNG mode (the reference's PCG64 streams are
injected), through the building-block path (MMIMO_LINK_PATH=blocks), and statistically in "philox" mode."""
import ctypes
import multiprocessing as mp
import os

import numpy as np
import pytest

from conftest import load_golden
from oracle import link_oracle as lo
from parity_util import check_frame
from test_gpu_dropin_blocks import M, build_objects, make_channel  # noqa: F401

pytestmark = pytest.mark.gpu

CASES = ["link_k64_los_mcnc", "link_k8_two_path_mcnc_clean", "link_k4_rayleigh_cnc", "link_k8_los_csi", "link_k8_los_rapp",
         "link_k64_los_cnc16", "link_k64_two_path_mcnc8", "link_k64_rayleigh_mcnc8"]


def run_link(M, g, rng_mode="numpy", n_frames=None, bits_max=None, n_err_min=int(1e12), path="fused"):
    n_fft, n_sc = int(g["n_fft"]), int(g["n_sc"])
    mod, dist, tx, rx, arr = build_objects(M, int(g["n_ant"]), int(g["constel"]), n_fft, n_sc, 128 if n_fft >= 1024 else 16,
                                           pa=str(g["pa"]), ibo=0.0)
    ch = make_channel(M, str(g["chan"]), arr, rx)
    nz = M["noise"].Awgn(snr_db=10, seed=1234)
    nb = mod.n_bits_per_ofdm_sym
    n_frames = int(g["n_frames"]) if n_frames is None else n_frames
    csi = float(g["csi_eps"])
    link = M["mp_model"].Link(mod_obj=mod, array_obj=arr, std_rx_obj=rx, chan_obj=ch, noise_obj=nz, rx_loc_var=10.0,
                              n_err_min=n_err_min, bits_sent_max=nb * n_frames if bits_max is None else bits_max,
                              is_mcnc=bool(g["is_mcnc"]), csi_epsylon=None if csi < 0 else csi)
    link.rng_mode = rng_mode
    link.frame_log = []   # per-frame decisions for the margin rule (tests/parity_util.py)
    link.update_distortion(ibo_val_db=float(g["ibo_db"]))
    link.set_snr(snr_db_val=float(g["snr_db"]))
    iters = np.asarray(g["iters"])
    n_res = len(iters) + int(g["incl_clean"])
    e = mp.Array(ctypes.c_double, n_res, lock=True)
    b = mp.Array(ctypes.c_double, n_res, lock=True)
    old = os.environ.get("MMIMO_LINK_PATH")
    os.environ["MMIMO_LINK_PATH"] = path
    try:
        link.simulate(bool(g["incl_clean"]), True, iters, [int(s) for s in g["seeds"]], e, b)
    finally:
        if old is None:
            del os.environ["MMIMO_LINK_PATH"]
        else:
            os.environ["MMIMO_LINK_PATH"] = old
    return np.array(list(e)), np.array(list(b)), link


def oracle_frames(g):
    """the same worker through the CPU oracle (bit-exact with the reference, tests/test_oracle_golden.py): per-frame decisions,
    slicer inputs and error counts"""
    pa = {"softlim": lo.PA_SOFT_LIMITER, "rapp": lo.PA_RAPP, "toi": lo.PA_TOI}[str(g["pa"])]
    cfg = lo.LinkConfig(constel_size=int(g["constel"]), n_fft=int(g["n_fft"]), n_sub_carr=int(g["n_sc"]),
                        n_ant=int(g["n_ant"]), pa_kind=pa, ibo_db=float(g["ibo_db"]), channel=str(g["chan"]))
    csi = float(g["csi_eps"])
    errs, bits, frames = lo.link_simulate(cfg, float(g["snr_db"]), int(np.max(g["iters"])), list(g["seeds"]), int(g["n_frames"]),
                                          is_mcnc=bool(g["is_mcnc"]), incl_clean_run=bool(g["incl_clean"]),
                                          csi_eps=None if csi < 0 else csi, return_frames=True)
    return cfg, frames


def check_against_reference(name, g, err, bits, link, blocks=False):
    """Counters: the bit counters are exact; every frame's decisions and error counts equal the reference's under the margin
    rule; when no frame of the case hit a sub-margin decision the shared error counters equal the reference's exactly."""
    assert np.array_equal(bits, g["bits"])
    cfg, frames = oracle_frames(g)
    iters = np.asarray(g["iters"])
    assert len(link.frame_log) == len(frames)
    n_div = 0
    bps = cfg.n_bits_per_symbol
    for fr, rec in zip(frames, link.frame_log):
        if blocks:  # building-block path: the receiver's bit vectors, ascending iteration order
            its = np.asarray(rec["iters"])
            det = np.stack([lo.bits_to_indices(b, bps) for b in rec["bits"]])
            e = np.array([int(np.sum(np.asarray(b) != rec["tx_bits"])) for b in rec["bits"]])
        else:
            its = np.arange(rec["n_iter"] + 1)
            det, e = rec["det"][its], rec["err"][its]
        corr = np.stack(fr["corr"])[its]
        n_ok = check_frame("drop-in Link %s%s" % (name, " (blocks)" if blocks else ""), det, fr["det"][its], corr,
                           cfg.constel_size, e, fr["err"][its])
        n_div += n_ok < len(its)
    if n_div == 0:
        assert np.array_equal(err, g["err"]), (err, g["err"])
    assert n_div <= max(1, len(frames) // 16)


@pytest.mark.parametrize("name", CASES)
def test_link_fused_numpy_streams_match_reference(M, name):
    g = load_golden(name)
    err, bits, link = run_link(M, g)
    assert link._fused_ok()
    check_against_reference(name, g, err, bits, link)


@pytest.mark.parametrize("name", ["link_k8_two_path_mcnc_clean", "link_k4_rayleigh_cnc", "link_k16_los_small"])
def test_link_block_path_matches_reference(M, name):
    """The same loop composed from the drop-in building blocks (also the only path for n_fft != 4096)."""
    g = load_golden(name)
    err, bits, link = run_link(M, g, path="blocks")
    check_against_reference(name, g, err, bits, link, blocks=True)


def test_link_stop_rule_and_stream_rewind(M):
    """bits_sent_max reached inside a batch: the counters stop exactly there and the generators are left where the
    reference leaves them (here: a second simulate with the same seeds must give identical numbers)."""
    g = load_golden("link_k8_los_rapp")
    nb = 6 * 2048
    err1, bits1, link = run_link(M, g, n_frames=5)
    assert np.all(bits1 == 5 * nb)
    # error threshold: slot 0 (most errors) stops first, the others keep counting
    thr = int(err1[0] * 0.55)
    err2, bits2, _ = run_link(M, g, n_frames=5, n_err_min=thr)
    assert bits2[0] < 5 * nb and err2[0] >= thr and err2[0] - thr < nb
    assert np.all(bits2[1:] >= bits2[0])


def test_link_philox_mode_statistics(M):
    """device RNG through the drop-in Link: BER at Eb/N0 15 dB inside the confidence interval of the reference's shipped
    curve (tests/ci_util.py; frame-level variance from this very run's per-frame error counts)"""
    from ci_util import assert_within_ci, dispersion, ref_bits
    g = load_golden("link_k64_los_mcnc")
    err, bits, link = run_link(M, g, rng_mode="philox", n_frames=1024)
    ref = load_golden("ref_curves")["mcnc_los"]
    col = int(np.argmin(np.abs(ref[0] - 15.0)))
    p_ref = ref[2:8, col]                                   # iterations 0..5
    phi = dispersion(np.stack([r["err"] for r in link.frame_log]))
    assert_within_ci("drop-in Link philox K=64 LOS MCNC Eb/N0 15 dB", err / bits, bits, p_ref, ref_bits(p_ref), phi, min_points=6)


def _runner_pattern(M, n_workers, rng_mode):
    """The launcher of main_mp_miso_mcnc_ber_vs_ebn0.py:121-141: shared counters, one mp.Process per worker running
    link.simulate with its own seeds, join, BER = errors / bits."""
    g = load_golden("link_k8_los_rapp")
    mod, dist, tx, rx, arr = build_objects(M, 8, pa="softlim", ibo=0.0)
    ch = make_channel(M, "los", arr, rx)
    link = M["mp_model"].Link(mod_obj=mod, array_obj=arr, std_rx_obj=rx, chan_obj=ch, noise_obj=M["noise"].Awgn(snr_db=10, seed=1234),
                              rx_loc_var=10.0, n_err_min=int(1e9), bits_sent_max=12288 * 40, is_mcnc=True)
    link.rng_mode = rng_mode
    link.update_distortion(ibo_val_db=0.0)
    link.set_snr(snr_db_val=float(M["utilities"].ebn0_to_snr(15.0, 2048, 2048, 64)))
    iters = np.array([0, 1, 2, 3])
    e = mp.Array(ctypes.c_double, len(iters) + 1, lock=True)
    b = mp.Array(ctypes.c_double, len(iters) + 1, lock=True)
    seeds = np.random.default_rng(2137).integers(0, 2 ** 62, size=(n_workers, 3))
    procs = [mp.Process(target=link.simulate, args=(True, True, iters, seeds[i], e, b)) for i in range(n_workers)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0
    return np.array(list(e)), np.array(list(b))


@pytest.mark.parametrize("rng_mode", ["numpy", "philox"])
def test_multiprocessing_runner_pattern(M, rng_mode):
    """Two worker PROCESSES (spawned: CUDA contexts do not survive fork) sharing the counters like the reference's runners."""
    assert mp.get_start_method() == "spawn"  # set by importing the drop-in mp_model
    err, bits = _runner_pattern(M, 2, rng_mode)
    # every slot stops at bits_sent_max; two workers may overshoot by at most one frame each (like the reference)
    assert np.all(bits >= 12288 * 40) and np.all(bits <= 12288 * 42)
    ber = err / bits
    assert ber[0] < 5e-3                                  # no distortion
    assert 0.05 < ber[1] < 0.14 and np.all(np.diff(ber[1:]) < 0)  # standard RX, then improving with MCNC iterations


def test_fixed_random_paths_channel_through_the_fused_link(M):
    """A channel object the fused path has no generator for (MisoRandomPathsFd, BASELINE config 3) with reroll_chan=False: the
    drop-in Link injects the object's channel_mat_fd into the fused kernels; decisions and counters against the reference's
    run of the same worker (tests/golden/link_k8_random_paths_fixed.npz) under the margin rule."""
    from test_oracle_golden import random_paths_channel
    g = load_golden("link_k8_random_paths_fixed")
    mod, dist, tx, rx, arr = build_objects(M, 8, pa="softlim", ibo=0.0)
    ch = M["channel"].MisoRandomPathsFd(tx_transceivers=arr.array_elements, rx_transceiver=rx, seed=5, n_paths=8,
                                        max_delay_spread=1e-6)
    link = M["mp_model"].Link(mod_obj=mod, array_obj=arr, std_rx_obj=rx, chan_obj=ch, noise_obj=M["noise"].Awgn(snr_db=10, seed=1234),
                              rx_loc_var=10.0, n_err_min=int(1e12), bits_sent_max=12288 * int(g["n_frames"]), is_mcnc=True)
    link.frame_log = []
    link.update_distortion(ibo_val_db=float(g["ibo_db"]))
    link.set_snr(snr_db_val=float(g["snr_db"]))
    assert link._fused_ok(reroll_chan=False) and not link._fused_ok(reroll_chan=True)
    iters = np.asarray(g["iters"])
    e = mp.Array(ctypes.c_double, len(iters), lock=True)
    b = mp.Array(ctypes.c_double, len(iters), lock=True)
    link.simulate(False, False, iters, [int(s) for s in g["seeds"]], e, b)
    assert np.array_equal(np.array(list(b)), g["bits"])
    cfg = lo.LinkConfig(n_ant=8, ibo_db=float(g["ibo_db"]))
    errs, bits, frames = lo.link_simulate(cfg, float(g["snr_db"]), int(iters.max()), list(g["seeds"]), int(g["n_frames"]),
                                          reroll_chan=False, h_fixed=random_paths_channel(cfg), return_frames=True)
    n_div = 0
    for fr, rec in zip(frames, link.frame_log):
        its = np.arange(rec["n_iter"] + 1)
        n_ok = check_frame("drop-in Link fixed random-paths channel", rec["det"][its], fr["det"][its], np.stack(fr["corr"])[its], 64,
                           rec["err"][its], fr["err"][its])
        n_div += n_ok < len(its)
    if n_div == 0:
        assert np.array_equal(np.array(list(e)), g["err"])
