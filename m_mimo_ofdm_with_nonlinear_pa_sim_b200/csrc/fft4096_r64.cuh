// 4096-point complex FFT / IFFT as TWO radix-64 passes: 64 threads (two warps) per transform, 64 points per
// thread, ONE shared-memory exchange per transform (the three-pass radix-16 form in fft4096.cuh needs two).
//
// Replaces the reference's torch.fft.ifft / torch.fft.fft (norm="ortho") call sites modulation.py:270,290,
// utilities.py:329,339, corrector.py:93,98.  The 1/sqrt(N) factors are NOT applied here (callers fold 1/64 into
// the precoder and the AGC).  Arithmetic: packed fp32 (FADD2 / FMUL2 / FFMA2), see fft4096.cuh.
//
// Index algebra (validated against numpy in tests/test_fft_index_math.py):
//   n = n0 + 64 n1,  k = k0 + 64 k1                       (n0, n1, k0, k1 in 0..63)
//   pass 1: lane n0 of the group holds x[n0 + 64 n1], radix-64 over n1 -> k0, times W4096^(n0 k0)
//   exchange: buf[k0][n0 ^ (k0 & 63)]  (XOR swizzle: both sides conflict-free for 64-bit accesses)
//   pass 2: lane k0 holds n0 = 0..63, radix-64 over n0 -> k1, output bin k0 + 64 k1
// The in-register radix-64 is three radix-4 stages, decimation in frequency: natural order in, base-4
// digit-reversed order out (position p = 16 r + 4 r' + r''  <->  index r + 4 r' + 16 r''); the reversal is a
// compile-time renaming of registers (rev64).
#pragma once
#include "fft4096.cuh"

namespace mmimo {

constexpr int R64_GROUP = 64;  // threads per transform

__host__ __device__ constexpr int rev64(int p) { return (p >> 4) | (((p >> 2) & 3) << 2) | ((p & 3) << 4); }

namespace r64 {
constexpr float QC[17] = {1.f,
                          0.99518472667219693f, 0.98078528040323043f, 0.95694033573220882f, 0.92387953251128674f,
                          0.88192126434835505f, 0.83146961230254524f, 0.77301045336273699f, 0.70710678118654757f,
                          0.63439328416364549f, 0.55557023301960229f, 0.47139673682599781f, 0.38268343236508984f,
                          0.29028467725446233f, 0.19509032201612833f, 0.09801714032956077f, 0.f};
__host__ __device__ constexpr float cos64(int m) {
    m &= 63;
    return m <= 16 ? QC[m] : m <= 32 ? -QC[32 - m] : m <= 48 ? -QC[m - 32] : QC[64 - m];
}
__host__ __device__ constexpr float sin64(int m) { return cos64(m - 16); }
}  // namespace r64

// multiply by W64^M, W64 = exp(S * 2 pi i / 64), M compile-time
template <int S, int M>
__device__ __forceinline__ pf2 pmulw64(pf2 a) {
    constexpr int m = M & 63;
    if (m == 0) return a;
    if (m == 16) return pmulj<S>(a);
    if (m == 32) return pmul(a, pbc(-1.f));
    if (m == 48) return pmulj<-S>(a);
    constexpr float c = r64::cos64(m), s = (S > 0 ? 1.f : -1.f) * r64::sin64(m);
    return pcmul(a, c, s);
}

// radix-4 butterfly on (a0,a1,a2,a3) -> outputs r = 0..3:  y_r = sum_m a_m (S j)^(m r)
template <int S>
__device__ __forceinline__ void bf4(pf2& a0, pf2& a1, pf2& a2, pf2& a3) {
    const pf2 s02 = padd(a0, a2), d02 = psub(a0, a2);
    const pf2 s13 = padd(a1, a3), d13 = psub(a1, a3);
    a0 = padd(s02, s13);
    a1 = paddj<S>(d02, d13);
    a2 = psub(s02, s13);
    a3 = paddj<-S>(d02, d13);
}

// In-register 64-point DFT (sign S), natural order in, digit-reversed out: X[k] ends up in x[rev64(k)].
//   PRUNE_IN : x[17..47] are known to be zero (values ignored); x[16] is honoured (the stray bin, see below)
//   PRUNE_OUT: only k in 0..16 and 48..63 are produced
template <int S, bool PRUNE_IN>
__device__ __forceinline__ void radix64_stage_a(pf2 (&x)[64]) {
    // stage A: butterflies over m (stride 16), then W64^(q r)
#pragma unroll
    for (int q = 0; q < 16; ++q) {
        if (PRUNE_IN) {
            const pf2 a0 = x[q], a3 = x[q + 48];
            if (q == 0) {
                const pf2 a1 = x[16];
                x[0] = padd(padd(a0, a3), a1);
                x[16] = paddj<S>(paddj<-S>(a0, a3), a1);
                x[32] = psub(psub(a0, a3), a1);
                x[48] = paddj<-S>(paddj<S>(a0, a3), a1);
            } else {
                x[q] = padd(a0, a3);
                x[q + 16] = paddj<-S>(a0, a3);
                x[q + 32] = psub(a0, a3);
                x[q + 48] = paddj<S>(a0, a3);
            }
        } else {
            bf4<S>(x[q], x[q + 16], x[q + 32], x[q + 48]);
        }
    }
}
template <int S, bool PRUNE_OUT>
__device__ __forceinline__ void radix64_stage_c(pf2 (&x)[64]) {
    // stage C: sixteen radix-4 butterflies on adjacent quadruples; output r'' is the HIGH digit of k
#pragma unroll
    for (int b = 0; b < 16; ++b) {
        if (PRUNE_OUT && b != 0) {  // only r'' = 0 and 3 (k < 16 or k >= 48); block 0 also yields k = 16
            const pf2 a0 = x[4 * b], a1 = x[4 * b + 1], a2 = x[4 * b + 2], a3 = x[4 * b + 3];
            x[4 * b] = padd(padd(a0, a2), padd(a1, a3));
            x[4 * b + 3] = paddj<-S>(psub(a0, a2), psub(a1, a3));
        } else {
            bf4<S>(x[4 * b], x[4 * b + 1], x[4 * b + 2], x[4 * b + 3]);
        }
    }
}
template <int S>
__device__ __forceinline__ void radix64_stage_b(pf2 (&x)[64]) {
#define MMIMO_TWA(q)                                   \
    x[q + 16] = pmulw64<S, (q) * 1>(x[q + 16]);        \
    x[q + 32] = pmulw64<S, (q) * 2>(x[q + 32]);        \
    x[q + 48] = pmulw64<S, (q) * 3>(x[q + 48]);
    MMIMO_TWA(1) MMIMO_TWA(2) MMIMO_TWA(3) MMIMO_TWA(4) MMIMO_TWA(5) MMIMO_TWA(6) MMIMO_TWA(7) MMIMO_TWA(8)
    MMIMO_TWA(9) MMIMO_TWA(10) MMIMO_TWA(11) MMIMO_TWA(12) MMIMO_TWA(13) MMIMO_TWA(14) MMIMO_TWA(15)
#undef MMIMO_TWA
    // stage B: four 16-point blocks, butterflies over stride 4, then W16^(q' r')
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
        for (int q = 0; q < 4; ++q) bf4<S>(x[16 * r + q], x[16 * r + q + 4], x[16 * r + q + 8], x[16 * r + q + 12]);
        x[16 * r + 1 + 4] = pmulw64<S, 4>(x[16 * r + 1 + 4]);
        x[16 * r + 2 + 4] = pmulw64<S, 8>(x[16 * r + 2 + 4]);
        x[16 * r + 3 + 4] = pmulw64<S, 12>(x[16 * r + 3 + 4]);
        x[16 * r + 1 + 8] = pmulw64<S, 8>(x[16 * r + 1 + 8]);
        x[16 * r + 2 + 8] = pmulw64<S, 16>(x[16 * r + 2 + 8]);
        x[16 * r + 3 + 8] = pmulw64<S, 24>(x[16 * r + 3 + 8]);
        x[16 * r + 1 + 12] = pmulw64<S, 12>(x[16 * r + 1 + 12]);
        x[16 * r + 2 + 12] = pmulw64<S, 24>(x[16 * r + 2 + 12]);
        x[16 * r + 3 + 12] = pmulw64<S, 36>(x[16 * r + 3 + 12]);
    }
}
template <int S, bool PRUNE_IN, bool PRUNE_OUT>
__device__ __forceinline__ void radix64(pf2 (&x)[64]) {
    radix64_stage_a<S, PRUNE_IN>(x);
    radix64_stage_b<S>(x);
    radix64_stage_c<S, PRUNE_OUT>(x);
}
// same with the pruning decided at run time (uniform branches around the two small pruned stages): lets ONE
// unrolled transform body serve a pruned-input and a pruned-output transform
template <int S>
__device__ __forceinline__ void radix64_rt(pf2 (&x)[64], const bool prune_in, const bool prune_out) {
    if (prune_in) radix64_stage_a<S, true>(x); else radix64_stage_a<S, false>(x);
    radix64_stage_b<S>(x);
    if (prune_out) radix64_stage_c<S, true>(x); else radix64_stage_c<S, false>(x);
}

// Named barrier of one 64-thread group (ids 1..15; 0 is __syncthreads)
__device__ __forceinline__ void group_bar(const int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }

struct SyncGroup {
    static constexpr bool kExchange = true;
    int id;
    __device__ __forceinline__ void x_sync() const { group_bar(id); }
};

// Inter-pass twiddles W4096^(n0 k0), k0 = a + 8 b, as products of 7 + 7 per-thread register values
// A[a] = W4096^(n0 a), B[b] = W512^(n0 b) (forward sign): no shared-memory traffic, 49 extra complex multiplies
// per transform.
struct TwReg64 {
    float2 A[7], B[7];
};
__device__ __forceinline__ void load_tw64(TwReg64& tw, const float2* __restrict__ tw4096, const int lane) {
#pragma unroll
    for (int a = 1; a < 8; ++a) {
        tw.A[a - 1] = tw4096[(lane * a) & 4095];
        tw.B[a - 1] = tw4096[(lane * a * 8) & 4095];
    }
}
template <int S>
__device__ __forceinline__ pf2 tw_apply(pf2 v, const TwReg64& tw, const int k0, const int /*lane*/) {
    const int a = k0 & 7, b = k0 >> 3;
    if (a) v = (S > 0) ? pcmulc(v, tw.A[a - 1].x, tw.A[a - 1].y) : pcmul(v, tw.A[a - 1].x, tw.A[a - 1].y);
    if (b) v = (S > 0) ? pcmulc(v, tw.B[b - 1].x, tw.B[b - 1].y) : pcmul(v, tw.B[b - 1].x, tw.B[b - 1].y);
    return v;
}
// table W4096^(n0 k0) at [k0 * 64 + n0] (forward sign) in shared memory
struct TwSmem64 {
    const float2* t;
};
template <int S>
__device__ __forceinline__ pf2 tw_apply(pf2 v, const TwSmem64& tw, const int k0, const int lane) {
    if (k0 == 0) return v;
    const float2 w = tw.t[k0 * 64 + lane];
    return (S > 0) ? pcmulc(v, w.x, w.y) : pcmul(v, w.x, w.y);
}

// One full transform by a 64-thread group.  x in: x[n1] = element n0 + 64 n1 (lane n0 = threadIdx & 63).
// x out: X[bin k0 + 64 k1] in x[rev64(k1)] of lane k0.  buf: the group's private 4096-entry exchange buffer.
// Two group barriers: one between the stores and the loads of the exchange, one after the loads (the buffer
// is reused by the next transform).
template <int S, bool PRUNE_IN, bool PRUNE_OUT, class TW, class SYNC = SyncGroup>
__device__ __forceinline__ void fft4096_r64(pf2 (&x)[64], pf2* __restrict__ buf, const TW& tw, const int lane,
                                            const SYNC sy) {
    radix64<S, PRUNE_IN, false>(x);
    // X1[k0] sits in x[rev64(k0)]: twiddle and store row k0
#pragma unroll
    for (int p = 0; p < 64; ++p) {
        const int k0 = rev64(p);
        const pf2 v = tw_apply<S>(x[p], tw, k0, lane);
        if (SYNC::kExchange) buf[k0 * 64 + (lane ^ k0)] = v;
        else x[p] = v;
    }
    sy.x_sync();
    if (SYNC::kExchange) {
#pragma unroll
        for (int n0 = 0; n0 < 64; ++n0) x[n0] = buf[lane * 64 + (n0 ^ lane)];
    }
    sy.x_sync();
    radix64<S, false, PRUNE_OUT>(x);
}

// Variant with the loads software-pipelined by hand: the inter-pass twiddles are fetched DEPTH elements ahead of
// the multiply that uses them, and the exchange is read back in the order stage A of the second pass consumes
// it (q, q+16, q+32, q+48), so the first butterflies start after four loads instead of forty-nine.
// `after_exchange` runs once the group has finished reading the exchange buffer (the buffer is free from there
// until the next transform's stores), before the second radix-64 pass.
template <int S, bool PRUNE_IN, bool PRUNE_OUT, int DEPTH = 8, class SYNC = SyncGroup, class HOOK = NoHook>
__device__ __forceinline__ void fft4096_r64_pipelined(pf2 (&x)[64], pf2* __restrict__ buf, const TwSmem64& tw,
                                                      const int lane, const SYNC sy,
                                                      const HOOK after_exchange = HOOK()) {
    radix64<S, PRUNE_IN, false>(x);
    float2 w[DEPTH];
#pragma unroll
    for (int p = 0; p < DEPTH; ++p) w[p] = tw.t[rev64(p) * 64 + lane];
#pragma unroll
    for (int p = 0; p < 64; ++p) {
        const int k0 = rev64(p);
        const float2 wp = w[p % DEPTH];
        if (p + DEPTH < 64) w[p % DEPTH] = tw.t[rev64(p + DEPTH) * 64 + lane];
        const pf2 v = k0 == 0 ? x[p] : (S > 0) ? pcmulc(x[p], wp.x, wp.y) : pcmul(x[p], wp.x, wp.y);
        buf[k0 * 64 + (lane ^ k0)] = v;
    }
    sy.x_sync();
#pragma unroll
    for (int q = 0; q < 16; ++q)
#pragma unroll
        for (int m = 0; m < 4; ++m) x[q + 16 * m] = buf[lane * 64 + ((q + 16 * m) ^ lane)];
    sy.x_sync();
    after_exchange();
    radix64<S, false, PRUNE_OUT>(x);
}

// Run-time pruned form of fft4096_r64_pipelined (see radix64_rt): the first pass prunes its input when
// prune_in, the second pass prunes its output when prune_out.
// `after_exchange` runs once the group has finished reading the exchange buffer (the buffer is free from there
// until the next transform's stores), before the second radix-64 pass.
template <int S, int DEPTH = 8, class SYNC = SyncGroup, class HOOK = NoHook>
__device__ __forceinline__ void fft4096_r64_pipelined_rt(pf2 (&x)[64], pf2* __restrict__ buf, const TwSmem64& tw,
                                                         const int lane, const SYNC sy, const bool prune_in,
                                                         const bool prune_out, const HOOK after_exchange = HOOK()) {
    radix64_rt<S>(x, prune_in, false);
    float2 w[DEPTH];
#pragma unroll
    for (int p = 0; p < DEPTH; ++p) w[p] = tw.t[rev64(p) * 64 + lane];
#pragma unroll
    for (int p = 0; p < 64; ++p) {
        const int k0 = rev64(p);
        const float2 wp = w[p % DEPTH];
        if (p + DEPTH < 64) w[p % DEPTH] = tw.t[rev64(p + DEPTH) * 64 + lane];
        const pf2 v = k0 == 0 ? x[p] : (S > 0) ? pcmulc(x[p], wp.x, wp.y) : pcmul(x[p], wp.x, wp.y);
        buf[k0 * 64 + (lane ^ k0)] = v;
    }
    sy.x_sync();
#pragma unroll
    for (int q = 0; q < 16; ++q)
#pragma unroll
        for (int m = 0; m < 4; ++m) x[q + 16 * m] = buf[lane * 64 + ((q + 16 * m) ^ lane)];
    sy.x_sync();
    after_exchange();
    radix64_rt<S>(x, false, prune_out);
}

}  // namespace mmimo
