"""Index algebra of csrc/fft4096.cuh and the native slot order of csrc/link_common.cuh, restated in numpy and checked
against numpy's FFT: the three radix-16 passes with the two shared-memory exchanges compute a 4096-point DFT, and the
pruned variants touch exactly the bins the OFDM grid uses.  CPU only (documents the kernel's layout contracts)."""
import numpy as np

N = 4096


def fft_three_pass(x, sign):
    """n = n0 + 16 n1 + 256 n2, k = k0 + 16 k1 + 256 k2 (fft4096.cuh header comment)."""
    w = lambda m, e: np.exp(sign * 2j * np.pi * e / m)
    t = np.arange(256)
    # pass 1: thread t = n0 + 16 n1 holds x[t + 256 j]; DFT16 over j -> k0; times W4096^(t k0)
    a = x.reshape(16, 256)                                   # [j][t]
    d16 = w(16, np.outer(np.arange(16), np.arange(16)))      # [k][j]
    p1 = (d16 @ a) * w(N, np.outer(np.arange(16), t))        # [k0][t]
    # exchange 1: thread t' = n0 + 16 k0 holds n1 = 0..15
    p1 = p1.reshape(16, 16, 16)                              # [k0][n1][n0]
    b = np.transpose(p1, (1, 0, 2))                          # [n1][k0][n0]
    # pass 2: DFT16 over n1 -> k1; times W256^(n0 k1)
    p2 = np.einsum("kn,nab->kab", d16, b) * w(256, np.arange(16)[:, None, None] * np.arange(16)[None, None, :])  # [k1][k0][n0]
    # exchange 2: thread t'' = k0 + 16 k1 holds n0 = 0..15; pass 3: DFT16 over n0 -> k2; bin = t'' + 256 k2
    c = np.transpose(p2, (2, 0, 1))                          # [n0][k1][k0]
    p3 = np.einsum("kn,nab->kab", d16, c)                    # [k2][k1][k0]
    return p3.reshape(N)


def test_three_pass_decomposition_is_a_dft():
    rng = np.random.default_rng(0)
    x = rng.standard_normal(N) + 1j * rng.standard_normal(N)
    assert np.allclose(fft_three_pass(x, -1), np.fft.fft(x), atol=1e-9)
    assert np.allclose(fft_three_pass(x, +1), np.fft.ifft(x) * N, atol=1e-9)


def slot_to_bin(p):
    if p == 0:
        return 1024
    jp, t = p >> 8, p & 255
    return t + 256 * (jp if jp < 4 else jp + 8)


def bin_to_sc(b):
    return b - 3072 if b >= 3072 else 1023 + b


def test_native_slot_order_is_a_bijection_onto_the_data_sub_carriers():
    bins = [slot_to_bin(p) for p in range(2048)]
    data_bins = list(range(1, 1025)) + list(range(3072, 4096))
    assert sorted(bins) == sorted(data_bins)
    sc = [bin_to_sc(b) for b in bins]
    assert sorted(sc) == list(range(2048))
    # reference data order: first half -> bins 3072.., second half -> bins 1.. (modulation.py:266-267)
    assert bin_to_sc(3072) == 0 and bin_to_sc(4095) == 1023 and bin_to_sc(1) == 1024 and bin_to_sc(1024) == 2047
    # thread t owns FFT registers j = 0..3 and 12..15 (bins t + 256 j): exactly what PRUNE_IN / PRUNE_OUT keep,
    # except bin 0 (DC, unused) whose slot carries bin 1024 (register 4 of thread 0)
    for p in range(1, 2048):
        jp, t = p >> 8, p & 255
        j = jp if jp < 4 else jp + 8
        assert slot_to_bin(p) == t + 256 * j and (j < 4 or j >= 12)
