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


# ---- two-pass radix-64 form (csrc/fft4096_r64.cuh) ---------------------------------------------------------------
def _rev64(p):
    return (p >> 4) | (((p >> 2) & 3) << 2) | ((p & 3) << 4)


def _radix64_dif(x, sign):
    """the in-register 64-point DFT of fft4096_r64.cuh: three radix-4 stages, decimation in frequency, natural order in,
    base-4 digit-reversed order out (X[k] at position rev64(k))"""
    x = np.array(x, dtype=complex)
    w = lambda m: np.exp(sign * 2j * np.pi * m / 64)

    def bf(a0, a1, a2, a3):
        return [a0 + a1 + a2 + a3, (a0 - a2) + sign * 1j * (a1 - a3), (a0 + a2) - (a1 + a3), (a0 - a2) - sign * 1j * (a1 - a3)]

    for q in range(16):
        o = bf(x[q], x[q + 16], x[q + 32], x[q + 48])
        for r in range(4):
            x[q + 16 * r] = o[r] * w(q * r)
    for r in range(4):
        for q in range(4):
            o = bf(*[x[16 * r + q + 4 * i] for i in range(4)])
            for rp in range(4):
                x[16 * r + q + 4 * rp] = o[rp] * w(4 * q * rp)
    for b in range(16):
        x[4 * b:4 * b + 4] = bf(*x[4 * b:4 * b + 4])
    return x


def test_radix64_two_pass_decomposition():
    rng = np.random.default_rng(5)
    v = rng.standard_normal(64) + 1j * rng.standard_normal(64)
    for sign, ref in ((-1, np.fft.fft(v)), (+1, np.fft.ifft(v) * 64)):
        o = _radix64_dif(v, sign)
        assert np.allclose([o[_rev64(k)] for k in range(64)], ref)
    assert sorted(_rev64(p) for p in range(64)) == list(range(64)) and all(_rev64(_rev64(p)) == p for p in range(64))
    x = rng.standard_normal(4096) + 1j * rng.standard_normal(4096)
    sign = -1
    buf = np.zeros((64, 64), complex)                      # [k0][n0], stored at column n0 ^ k0 on the device
    for n0 in range(64):                                   # pass 1: lane n0 holds x[n0 + 64 n1]
        o = _radix64_dif(x[n0 + 64 * np.arange(64)], sign)
        for p in range(64):
            k0 = _rev64(p)
            buf[k0, n0] = o[p] * np.exp(sign * 2j * np.pi * n0 * k0 / 4096)
    out = np.zeros(4096, complex)
    for k0 in range(64):                                   # pass 2: lane k0 holds n0 = 0..63, output bin k0 + 64 k1
        o = _radix64_dif(buf[k0], sign)
        for k1 in range(64):
            out[k0 + 64 * k1] = o[_rev64(k1)]
    assert np.allclose(out, np.fft.fft(x))
    # the XOR swizzle of the exchange is a bijection per row and conflict-free per half-warp (16 lanes, 8-byte words)
    for k0 in range(64):
        assert sorted(n0 ^ k0 for n0 in range(64)) == list(range(64))
    for n0 in range(64):
        for h in range(4):
            banks = [((n0 ^ lane) % 16) for lane in range(16 * h, 16 * h + 16)]
            assert len(set(banks)) == 16
    # pruning sets: data bins of lane n0 are n0 + 64 j, j in 0..15 and 48..63, plus bin 1024 = 0 + 64 * 16 for lane 0
    data = [b for b in range(4096) if 1 <= b <= 1024 or b >= 3072]
    js = {(b % 64, b // 64) for b in data}
    assert all(j < 16 or j >= 48 or (n0 == 0 and j == 16) for n0, j in js)
