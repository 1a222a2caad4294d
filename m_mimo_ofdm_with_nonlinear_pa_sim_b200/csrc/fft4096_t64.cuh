// 4096-point forward complex FFT as TWO radix-64 passes (64 threads per transform, 64 points per thread, ONE
// shared-memory exchange), every twiddle factor in SCALED-TANGENT form so that it costs ONE packed FMA instead of a
// two-instruction complex multiply and every butterfly addition is a packed FMA with a real scale:
//
//     w = j^p * f * (1 + j t)          p in {0,1} chosen at compile time so that |t| stays small
//     w a = j^p f a',   a' = a + t (j a)                       <- one FFMA2 ("tangent step")
//     a0 +- w2 a2 = a0 +- f2 (j^p2 a2')                        <- one FFMA2 each instead of FADD2
//
// Replaces the reference's torch.fft.ifft / torch.fft.fft (norm="ortho") call sites modulation.py:270,290,
// utilities.py:329,339, corrector.py:93,98 inside the fused frame kernel (the 1/sqrt(N) factors are folded into
// the precoder and the AGC by the callers, as in fft4096.cuh).  Forward sign only: the frame kernel obtains the
// inverse transform by index reversal, IFFT(x)[n] = FFT(x)[-n] / N (frame_kernel_t64.cuh).
//
// Decomposition (W = exp(-2 pi i / 4096); validated against a double-precision DFT in tests/test_fft_t64_host.py,
// which compiles this header for the HOST and runs the 64 lanes one after the other):
//   n = n0 + 64 n1,  k = k0 + 64 k1
//   pass 1: lane n0 holds x[n0 + 64 n1] in register n1; 64-point DFT over n1 -> X1[n0][k0] (no twiddle applied)
//   exchange: buf[k0][n0 ^ k0]
//   pass 2: lane k0 holds X1[n0][k0] in register n0; computes  Y[k1] = sum_n0 X1[n0][k0] W^(n0 k0) W64^(n0 k1),
//           i.e. the inter-pass twiddle W^(n0 k0) is MERGED into the butterflies' own twiddles (below).
// One 64-point pass with base w = W^e (pass 1: e = 0; pass 2: e = k0), decimation in frequency with the twiddles
// placed on the butterfly INPUTS (the lead input of every butterfly is never twiddled):
//   level A: n0 = q + 16 m        A[q][r]   = sum_m x[q+16m]      W^(16 m e)            (-j)^(m r)
//   level B: q  = q' + 4 m'       B[q'][r'] = sum_m' A[q'+4m'][r] W^(4 m' (e + 64 r))   (-j)^(m' r')
//   level C:                      Y[r+4r'+16r''] = sum_q' B[q'][r'] W^(q' (e + 64 r + 256 r')) (-j)^(q' r'')
// Output index k1 = r + 4 r' + 16 r'' ends up in register 16 r + 4 r' + r'' (rev64, a compile-time renaming).
// Pass 1 uses compile-time constants; pass 2 reads per-lane parameters (t1,t2,t3,f2,rho,f1 per butterfly class:
// 1 + 4 + 16 classes = 63 twiddles = 126 floats per thread and transform, the same shared-memory traffic as the
// 63 complex inter-pass twiddles of fft4096_r64.cuh) from a 31.5 KB table built on the host in double precision
// (t64::build_table).  Instruction count per transform and thread: 2 x 384 butterfly FMAs + 76 + 63 tangent steps
// = 907 packed FMA-pipe instructions (fft4096_r64.cuh: 1208; nominal 5 N log2 N / 4 flop per FFMA2 = 960).
#pragma once
#include <math.h>
#include <stdint.h>

#include <utility>

#include "fft4096.cuh"

#if defined(__CUDACC__)
#define T64_HD __host__ __device__ __forceinline__
#else
#define T64_HD inline
#endif

namespace mmimo {
namespace t64 {

__host__ __device__ constexpr int rev64(int p) { return (p >> 4) | (((p >> 2) & 3) << 2) | ((p & 3) << 4); }

// ---- packed primitives: PTX on the device, plain float arithmetic on the host (validation build) -----------
T64_HD pf2 qpk(float lo, float hi) {
#if defined(__CUDA_ARCH__)
    return pk(lo, hi);
#else
    union { float f[2]; pf2 u; } c;
    c.f[0] = lo; c.f[1] = hi;
    return c.u;
#endif
}
T64_HD float2 qunpk(pf2 v) {
#if defined(__CUDA_ARCH__)
    return unpk(v);
#else
    union { float f[2]; pf2 u; } c;
    c.u = v;
    return make_float2(c.f[0], c.f[1]);
#endif
}
T64_HD pf2 qadd(pf2 a, pf2 b) {
#if defined(__CUDA_ARCH__)
    return padd(a, b);
#else
    const float2 x = qunpk(a), y = qunpk(b);
    return qpk(x.x + y.x, x.y + y.y);
#endif
}
T64_HD pf2 qsub(pf2 a, pf2 b) {
#if defined(__CUDA_ARCH__)
    return psub(a, b);
#else
    const float2 x = qunpk(a), y = qunpk(b);
    return qpk(x.x - y.x, x.y - y.y);
#endif
}
// c + f b      (f real)
T64_HD pf2 qfma_s(pf2 b, float f, pf2 c) {
#if defined(__CUDA_ARCH__)
    return pfma(b, pbc(f), c);
#else
    const float2 x = qunpk(b), z = qunpk(c);
    return qpk(fmaf(x.x, f, z.x), fmaf(x.y, f, z.y));
#endif
}
// c + f (j b)  = (c.x - f b.y, c.y + f b.x)
T64_HD pf2 qfma_js(pf2 b, float f, pf2 c) {
#if defined(__CUDA_ARCH__)
    return pfma(pswap(b), pk(-f, f), c);
#else
    const float2 x = qunpk(b), z = qunpk(c);
    return qpk(fmaf(x.y, -f, z.x), fmaf(x.x, f, z.y));
#endif
}
// c + f (j^P b),  P in {-1, 0, 1}
template <int P>
T64_HD pf2 qfma_jp(pf2 b, float f, pf2 c) {
    if (P == 0) return qfma_s(b, f, c);
    if (P == 1) return qfma_js(b, f, c);
    return qfma_js(b, -f, c);
}

// ---- twiddles in scaled-tangent form ------------------------------------------------------------------------
// angle unit: 2 pi / 4096, forward sign: W^u = cos(2 pi u / 4096) - j sin(2 pi u / 4096)
constexpr double C64[17] = {1.0,
                            0.99518472667219688624, 0.98078528040323044913, 0.95694033573220886494,
                            0.92387953251128675613, 0.88192126434835502971, 0.83146961230254523708,
                            0.77301045336273696081, 0.70710678118654752440, 0.63439328416364549822,
                            0.55557023301960222474, 0.47139673682599764856, 0.38268343236508977173,
                            0.29028467725446236764, 0.19509032201612826785, 0.09801714032956060199, 0.0};
__host__ __device__ constexpr double cos64(int m) {
    m = ((m % 64) + 64) % 64;
    return m <= 16 ? C64[m] : m <= 32 ? -C64[32 - m] : m <= 48 ? -C64[m - 32] : C64[64 - m];
}
__host__ __device__ constexpr double sin64(int m) { return cos64(m - 16); }

// form of a twiddle whose angle (in units) lies around `twice_mid / 2`: p = 0 (cosine form) when the angle is nearer
// to 0 or pi than to +-pi/2, p = 1 (sine form, w = j f (1 + j t)) otherwise
__host__ __device__ constexpr int form_of(int twice_mid) {
    const int x = ((twice_mid % 4096) + 4096) % 4096;  // twice the angle, mod 2 pi
    return (x > 1024 && x < 3072) ? 1 : 0;
}

// compile-time twiddle triple: inputs 1,2,3 of a butterfly carry W^(U1), W^(U2), W^(U3); U multiples of 64
template <int U1, int U2, int U3>
struct CTw {
    static_assert(U1 % 64 == 0 && U2 % 64 == 0 && U3 % 64 == 0, "pass-1 twiddles are powers of W64");
    static constexpr bool kConst = true;
    static constexpr int p1 = form_of(2 * U1), p2 = form_of(2 * U2), p3 = form_of(2 * U3);
    static constexpr double re(int u) { return cos64(u / 64); }
    static constexpr double im(int u) { return -sin64(u / 64); }
    static constexpr double fd(int u, int p) { return p ? im(u) : re(u); }
    static constexpr double td(int u, int p) { return p ? -re(u) / im(u) : im(u) / re(u); }
    static constexpr float t1 = (float)td(U1, p1), t2 = (float)td(U2, p2), t3 = (float)td(U3, p3);
    static constexpr float f1 = (float)fd(U1, p1), f2 = (float)fd(U2, p2);
    static constexpr float rho = (float)(fd(U3, p3) / fd(U1, p1));
};
// run-time twiddle triple (pass 2): forms fixed at compile time, values from the per-lane table
template <int P1, int P2, int P3>
struct RTw {
    static constexpr bool kConst = false;
    static constexpr int p1 = P1, p2 = P2, p3 = P3;
    float t1, t2, t3, f2, rho, f1;
};

// Radix-4 butterfly, forward sign, twiddles on inputs 1..3:  y_r = sum_m (w_m a_m) (-j)^(m r),  w_0 = 1.
// Outputs in place (a_r <- y_r).  11 packed FMAs (8 when the twiddles are trivial).
// OUT: bit r set <=> output y_r is needed (pruned last level of the frame kernel's second transform).
template <class TW, int OUT = 15>
T64_HD void bf4_tw(pf2& a0, pf2& a1, pf2& a2, pf2& a3, const TW& tw) {
    bool skip1 = false, skip2 = false, skip3 = false;
    if (TW::kConst) {
        skip1 = tw.t1 == 0.f;
        skip2 = tw.t2 == 0.f;
        skip3 = tw.t3 == 0.f;
    }
    if (!skip1) a1 = qfma_js(a1, tw.t1, a1);
    if (!skip2) a2 = qfma_js(a2, tw.t2, a2);
    if (!skip3) a3 = qfma_js(a3, tw.t3, a3);
    pf2 s02, d02, s13, d13;
    if (TW::kConst && TW::p2 == 0 && tw.f2 == 1.f) {
        s02 = qadd(a0, a2);
        d02 = qsub(a0, a2);
    } else {
        s02 = qfma_jp<TW::p2>(a2, tw.f2, a0);
        d02 = qfma_jp<TW::p2>(a2, -tw.f2, a0);
    }
    constexpr int dp = TW::p3 - TW::p1;
    if (TW::kConst && dp == 0 && tw.rho == 1.f) {
        s13 = qadd(a1, a3);
        d13 = qsub(a1, a3);
    } else {
        s13 = qfma_jp<dp>(a3, tw.rho, a1);
        d13 = qfma_jp<dp>(a3, -tw.rho, a1);
    }
    if (TW::kConst && TW::p1 == 0 && tw.f1 == 1.f) {
        if (OUT & 1) a0 = qadd(s02, s13);
        if (OUT & 4) a2 = qsub(s02, s13);
    } else {
        if (OUT & 1) a0 = qfma_jp<TW::p1>(s13, tw.f1, s02);
        if (OUT & 4) a2 = qfma_jp<TW::p1>(s13, -tw.f1, s02);
    }
    // y1 = d02 + (-j) j^p1 f1 d13',  y3 = d02 - (-j) j^p1 f1 d13'
    if (OUT & 2) a1 = qfma_jp<TW::p1 - 1>(d13, tw.f1, d02);
    if (OUT & 8) a3 = qfma_jp<TW::p1 - 1>(d13, -tw.f1, d02);
}

// ---- per-lane table of pass 2 (host-built, double precision) ------------------------------------------------
// level C: class b = r + 4 r' (0..15), e = lane + 64 b, twiddles W^(q' e), q' = 1..3
// level B: class r (0..3), e = lane + 64 r, twiddles W^(4 m' e)
// level A: e = lane, twiddles W^(16 m e)
// forms decided by the middle of the lane range (lane = 31.5)
__host__ __device__ constexpr int form_c(int b, int q) { return form_of(q * (128 * b + 63)); }
__host__ __device__ constexpr int form_b(int r, int m) { return form_of(4 * m * (128 * r + 63)); }
// Table layout, float2 units: three planes per level, plane j of an entry holds (t1,t2), (t3,f2), (rho,f1)
constexpr int TAB_C = 0;               // [3][1024]
constexpr int TAB_B = 3 * 1024;        // [3][256]
constexpr int TAB_A = TAB_B + 3 * 256; // [3][64]
constexpr int TAB_FLOAT2 = TAB_A + 3 * 64;  // 4032
static_assert(TAB_FLOAT2 <= 4096, "table fits the 32 KB twiddle region of the frame kernels");

// host side: scaled-tangent parameters of W^u in form p; |f| is kept >= 2^-40 (exact hits of +-j in the cosine form: the
// representation error 2^-40 is far below fp32 resolution)
inline void tw_params(long long u, int p, double& f, double& t) {
    const double th = -2.0 * M_PI * (double)(((u % 4096) + 4096) % 4096) / 4096.0;
    const long long um = ((u % 4096) + 4096) % 4096;
    double c = cos(th), s = sin(th);
    if (um % 1024 == 0) {  // exact quarter turns
        const int qd = (int)(um / 1024);
        c = qd == 0 ? 1.0 : qd == 2 ? -1.0 : 0.0;
        s = qd == 1 ? -1.0 : qd == 3 ? 1.0 : 0.0;
    }
    const double tiny = ldexp(1.0, -40);
    if (p == 0) {
        f = c;
        if (fabs(f) < tiny) f = tiny;
        t = s / f;
    } else {
        f = s;
        if (fabs(f) < tiny) f = tiny;
        t = -c / f;
    }
}
inline void tw_entry(long long u1, long long u2, long long u3, int p1, int p2, int p3, float (&out)[6]) {
    double f1, f2, f3, t1, t2, t3;
    tw_params(u1, p1, f1, t1);
    tw_params(u2, p2, f2, t2);
    tw_params(u3, p3, f3, t3);
    out[0] = (float)t1; out[1] = (float)t2; out[2] = (float)t3;
    out[3] = (float)f2; out[4] = (float)(f3 / f1); out[5] = (float)f1;
}
// fills TAB_FLOAT2 float2 entries
inline void build_table(float2* tab) {
    float e6[6];
    for (int b = 0; b < 16; ++b)
        for (int lane = 0; lane < 64; ++lane) {
            const long long e = lane + 64 * b;
            tw_entry(e, 2 * e, 3 * e, form_c(b, 1), form_c(b, 2), form_c(b, 3), e6);
            for (int j = 0; j < 3; ++j) tab[TAB_C + j * 1024 + e] = make_float2(e6[2 * j], e6[2 * j + 1]);
        }
    for (int r = 0; r < 4; ++r)
        for (int lane = 0; lane < 64; ++lane) {
            const long long e = lane + 64 * r;
            tw_entry(4 * e, 8 * e, 12 * e, form_b(r, 1), form_b(r, 2), form_b(r, 3), e6);
            for (int j = 0; j < 3; ++j) tab[TAB_B + j * 256 + e] = make_float2(e6[2 * j], e6[2 * j + 1]);
        }
    for (int lane = 0; lane < 64; ++lane) {
        tw_entry(16 * lane, 32 * lane, 48 * lane, 0, 0, 0, e6);
        for (int j = 0; j < 3; ++j) tab[TAB_A + j * 64 + lane] = make_float2(e6[2 * j], e6[2 * j + 1]);
    }
}

struct Raw3 {  // one table entry as loaded (three 64-bit loads)
    float2 a, b, c;
};
T64_HD Raw3 load_raw(const float2* __restrict__ plane0, const int stride, const int idx) {
    Raw3 r;
    r.a = plane0[idx];
    r.b = plane0[stride + idx];
    r.c = plane0[2 * stride + idx];
    return r;
}
template <int P1, int P2, int P3>
T64_HD RTw<P1, P2, P3> as_tw(const Raw3& r) {
    RTw<P1, P2, P3> tw;
    tw.t1 = r.a.x; tw.t2 = r.a.y; tw.t3 = r.b.x; tw.f2 = r.b.y; tw.rho = r.c.x; tw.f1 = r.c.y;
    return tw;
}

// ---- pass 1: compile-time twiddles (e = 0) ---------------------------------------------------------------------
template <int R>
T64_HD void p1_level_b(pf2 (&x)[64]) {
    const CTw<256 * R, 512 * R, 768 * R> tw{};
#pragma unroll
    for (int q = 0; q < 4; ++q) bf4_tw(x[16 * R + q], x[16 * R + q + 4], x[16 * R + q + 8], x[16 * R + q + 12], tw);
}
template <int B>  // B = r + 4 r'
T64_HD void p1_level_c(pf2 (&x)[64]) {
    const CTw<64 * B, 128 * B, 192 * B> tw{};
    constexpr int base = 16 * (B & 3) + 4 * (B >> 2);
    bf4_tw(x[base], x[base + 1], x[base + 2], x[base + 3], tw);
}
template <int... B>
T64_HD void p1_level_c_all(pf2 (&x)[64], std::integer_sequence<int, B...>) {
    (p1_level_c<B>(x), ...);
}
T64_HD void p1_level_a(pf2 (&x)[64]) {
    const CTw<0, 0, 0> tw0{};
#pragma unroll
    for (int q = 0; q < 16; ++q) bf4_tw(x[q], x[q + 16], x[q + 32], x[q + 48], tw0);
}
// level A when registers 17..47 are known to be zero (guard band of the frame kernel's first transform; register 16
// may be non-zero: lane 0 carries the stray bin there, frame_kernel_t64.cuh): y_r = a0 + j^r a3, 4 instead of 8
// instructions per butterfly.  The values of registers 17..47 on entry are ignored.
T64_HD void p1_level_a_pruned(pf2 (&x)[64]) {
    const CTw<0, 0, 0> tw0{};
    x[32] = qpk(0.f, 0.f);
    bf4_tw(x[0], x[16], x[32], x[48], tw0);
#pragma unroll
    for (int q = 1; q < 16; ++q) {
        const pf2 a0 = x[q], a3 = x[q + 48];
        x[q] = qadd(a0, a3);
        x[q + 16] = qfma_js(a3, 1.f, a0);
        x[q + 32] = qsub(a0, a3);
        x[q + 48] = qfma_js(a3, -1.f, a0);
    }
}
T64_HD void p1_levels_bc(pf2 (&x)[64]) {
    p1_level_b<0>(x);
    p1_level_b<1>(x);
    p1_level_b<2>(x);
    p1_level_b<3>(x);
    p1_level_c_all(x, std::make_integer_sequence<int, 16>());
}
T64_HD void radix64_const(pf2 (&x)[64]) {
    p1_level_a(x);
    p1_levels_bc(x);
}

// ---- pass 2: run-time twiddles from the per-lane table ------------------------------------------------------------
template <int R>
T64_HD void p2_level_b(pf2 (&x)[64], const Raw3& raw) {
    const auto tw = as_tw<form_b(R, 1), form_b(R, 2), form_b(R, 3)>(raw);
#pragma unroll
    for (int q = 0; q < 4; ++q) bf4_tw(x[16 * R + q], x[16 * R + q + 4], x[16 * R + q + 8], x[16 * R + q + 12], tw);
}
// PRUNE_OUT: only outputs r'' = 0 and 3 (k1 = 0..15, 48..63: the data bins) are produced, plus r'' = 1 of class 0
// (k1 = 16: bin 1024 of lane 0); the other registers hold garbage afterwards.
template <int B, bool PRUNE_OUT = false>
T64_HD void p2_level_c(pf2 (&x)[64], const Raw3& raw) {
    const auto tw = as_tw<form_c(B, 1), form_c(B, 2), form_c(B, 3)>(raw);
    constexpr int base = 16 * (B & 3) + 4 * (B >> 2);
    using TW = RTw<form_c(B, 1), form_c(B, 2), form_c(B, 3)>;
    bf4_tw<TW, PRUNE_OUT ? (B == 0 ? 11 : 9) : 15>(x[base], x[base + 1], x[base + 2], x[base + 3], tw);
}
// level C with the table entries fetched DEPTH classes ahead of the butterfly that uses them
template <int DEPTH, int B, bool PRUNE_OUT = false>
struct P2LevelC {
    static T64_HD void run(pf2 (&x)[64], Raw3 (&ring)[DEPTH], const float2* __restrict__ tab, const int lane) {
        const Raw3 cur = ring[B % DEPTH];
        if (B + DEPTH < 16) ring[B % DEPTH] = load_raw(tab + TAB_C, 1024, lane + 64 * (B + DEPTH));
        p2_level_c<B, PRUNE_OUT>(x, cur);
        P2LevelC<DEPTH, B + 1, PRUNE_OUT>::run(x, ring, tab, lane);
    }
};
template <int DEPTH, bool PRUNE_OUT>
struct P2LevelC<DEPTH, 16, PRUNE_OUT> {
    static T64_HD void run(pf2 (&)[64], Raw3 (&)[DEPTH], const float2* __restrict__, const int) {}
};

struct TabA {  // level-A entry of this lane, loaded once per kernel (the same for every transform)
    Raw3 raw;
};
T64_HD TabA load_tab_a(const float2* __restrict__ tab, const int lane) { return TabA{load_raw(tab + TAB_A, 64, lane)}; }

// level A of pass 2 (the table entries of level B are fetched first and returned in rb)
T64_HD void p2_level_a(pf2 (&x)[64], Raw3 (&rb)[4], const float2* __restrict__ tab, const TabA& ta, const int lane) {
#pragma unroll
    for (int r = 0; r < 4; ++r) rb[r] = load_raw(tab + TAB_B, 256, lane + 64 * r);
    const auto tw = as_tw<0, 0, 0>(ta.raw);
#pragma unroll
    for (int q = 0; q < 16; ++q) bf4_tw(x[q], x[q + 16], x[q + 32], x[q + 48], tw);
}
// level B of pass 2; `ring` comes back holding the first DEPTH level-C table entries
template <int DEPTH>
T64_HD void p2_level_b_all(pf2 (&x)[64], const Raw3 (&rb)[4], Raw3 (&ring)[DEPTH], const float2* __restrict__ tab,
                           const int lane) {
#pragma unroll
    for (int b = 0; b < DEPTH; ++b) ring[b] = load_raw(tab + TAB_C, 1024, lane + 64 * b);
    p2_level_b<0>(x, rb[0]);
    p2_level_b<1>(x, rb[1]);
    p2_level_b<2>(x, rb[2]);
    p2_level_b<3>(x, rb[3]);
}
template <int DEPTH>
T64_HD void p2_levels_ab(pf2 (&x)[64], Raw3 (&ring)[DEPTH], const float2* __restrict__ tab, const TabA& ta,
                         const int lane) {
    Raw3 rb[4];
    p2_level_a(x, rb, tab, ta, lane);
    p2_level_b_all<DEPTH>(x, rb, ring, tab, lane);
}
template <int DEPTH, bool PRUNE_OUT>
T64_HD void p2_level_c_all(pf2 (&x)[64], Raw3 (&ring)[DEPTH], const float2* __restrict__ tab, const int lane) {
    P2LevelC<DEPTH, 0, PRUNE_OUT>::run(x, ring, tab, lane);
}
template <int DEPTH>
T64_HD void radix64_table(pf2 (&x)[64], const float2* __restrict__ tab, const TabA& ta, const int lane) {
    Raw3 ring[DEPTH];
    p2_levels_ab<DEPTH>(x, ring, tab, ta, lane);
    p2_level_c_all<DEPTH, false>(x, ring, tab, lane);
}

// ---- the two halves of a transform (the device wrapper puts the group barriers between them) -------------------
// Exchange layout: element (k0, n0) of the 64 x 64 intermediate lives at buf[k0 * T64_ROW + n0] with T64_ROW = 65
// (one float2 of padding per row).  Stores (lane = n0, 64 consecutive 8-byte words per k0) and loads (lane = k0,
// stride 65 words: 16 lanes of a half-warp hit 16 different bank pairs) are both conflict-free, and every address
// is `lane-dependent base + compile-time offset`: no per-access integer instructions (the XOR swizzle this replaces
// cost a LOP3 and an IADD3 per access, ~10 % of the issue slots of a transform).
#ifndef T64_ROW
#define T64_ROW 65
#endif
constexpr int XROW = T64_ROW;
constexpr int XBUF = 64 * XROW;  // float2 elements of one group's exchange buffer
T64_HD int xidx(const int k0, const int n0) { return XROW == 64 ? k0 * 64 + (n0 ^ k0) : k0 * XROW + n0; }
// pass 1 + exchange stores.  x in: x[n1] = element lane + 64 n1.
T64_HD void exchange_store(const pf2 (&x)[64], pf2* __restrict__ buf, const int lane) {
#pragma unroll
    for (int p = 0; p < 64; ++p) buf[xidx(rev64(p), lane)] = x[p];
}
T64_HD void pass1_store(pf2 (&x)[64], pf2* __restrict__ buf, const int lane) {
    radix64_const(x);
    exchange_store(x, buf, lane);
}
// exchange loads (in the order level A consumes them)
T64_HD void pass2_load(pf2 (&x)[64], const pf2* __restrict__ buf, const int lane) {
#pragma unroll
    for (int q = 0; q < 16; ++q)
#pragma unroll
        for (int m = 0; m < 4; ++m) x[q + 16 * m] = buf[xidx(lane, q + 16 * m)];
}
// pass 2.  x out: X[bin lane + 64 k1] in x[rev64(k1)].
template <int DEPTH>
T64_HD void pass2_compute(pf2 (&x)[64], const float2* __restrict__ tab, const TabA& ta, const int lane) {
    radix64_table<DEPTH>(x, tab, ta, lane);
}

}  // namespace t64

#if defined(__CUDACC__)
// One forward transform by a 64-thread group.  `after_exchange` runs once the group has finished reading the
// exchange buffer (it is free from there until the next transform's stores).
template <int DEPTH, class SYNC, class HOOK>
__device__ __forceinline__ void fft4096_t64(pf2 (&x)[64], pf2* __restrict__ buf, const float2* __restrict__ tab,
                                            const t64::TabA& ta, const int lane, const SYNC sy,
                                            const HOOK after_exchange) {
    t64::pass1_store(x, buf, lane);
    sy.x_sync();
    t64::pass2_load(x, buf, lane);
    sy.x_sync();
    after_exchange();
    t64::pass2_compute<DEPTH>(x, tab, ta, lane);
}
#endif

}  // namespace mmimo
