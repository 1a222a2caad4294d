// 4096-point complex FFT / IFFT for one CTA of 256 threads, 16 points per thread,
// three radix-16 passes with two shared-memory exchanges (sm_100a, fp32).
//
// Replaces the reference's torch.fft.ifft / torch.fft.fft (norm="ortho") call sites
// modulation.py:270,290, utilities.py:329,339, corrector.py:93,98.  The 1/sqrt(N)
// factors are NOT applied here: callers fold 1/64 into the precoder and the AGC.
//
// All complex arithmetic uses the sm_100 packed-fp32 instructions (PTX add/sub/mul/fma.f32x2 ->
// SASS FADD2 / FMUL2 / FFMA2): one complex value = one 64-bit register pair (lo = re, hi = im), one
// complex add = ONE issue slot.  ptxas folds the pair swizzles written below as pk(hi, lo) / pk(s, s) /
// pk(-s, s) into operand modifiers (.LO_HI, .F32 broadcast, .NP per-lane sign), so a complex
// multiply is FMUL2 + FFMA2 with no extra moves.  Measured on B200 (profiles/r01_fp32x2_microbench.txt):
// the packed forms run at the full 128 fp32 lanes/clk/SM with half the issue slots of the scalar forms,
// which is what moves this kernel from issue-bound to FMA-pipe-bound.
//
// Index algebra (validated against numpy in tests/test_fft_index_math.py):
//   n = n0 + 16 n1 + 256 n2,  k = k0 + 16 k1 + 256 k2
//   pass 1: thread t = n0+16n1 holds x[t+256 j], radix-16 over j -> k0, times W4096^(t k0)
//   pass 2: thread t = n0+16k0 holds n1=0..15,   radix-16 -> k1, times W256^(n0 k1)
//   pass 3: thread t = k0+16k1 holds n0=0..15,   radix-16 -> k2, output bin t+256 k2
// Both exchanges are conflict-free for 64-bit accesses (half-warp = 16 consecutive 8-byte words;
// exchange 2 uses an XOR swizzle n0^k0).
#pragma once
#include <cuda_runtime.h>

namespace mmimo {

constexpr int FFT_N = 4096;
constexpr int FFT_THREADS = 256;

// ---- scalar complex helpers (prep / stage kernels) -------------------------------------------
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 cmulc(float2 a, float2 b) {  // a * conj(b)
    return make_float2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y);
}
__device__ __forceinline__ float2 cfma(float2 a, float2 b, float2 c) {  // a*b + c
    return make_float2(fmaf(a.x, b.x, fmaf(-a.y, b.y, c.x)), fmaf(a.x, b.y, fmaf(a.y, b.x, c.y)));
}

// ---- packed complex (f32x2) ------------------------------------------------------------------
typedef unsigned long long pf2;  // lo 32 bits = re, hi 32 bits = im

__device__ __forceinline__ pf2 pk(float lo, float hi) {
    pf2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float2 unpk(pf2 v) {
    float2 r;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
    return r;
}
__device__ __forceinline__ pf2 pk2(float2 v) { return pk(v.x, v.y); }
__device__ __forceinline__ pf2 pswap(pf2 v) {
    const float2 a = unpk(v);
    return pk(a.y, a.x);
}
__device__ __forceinline__ pf2 pbc(float s) { return pk(s, s); }
__device__ __forceinline__ pf2 padd(pf2 a, pf2 b) {
    pf2 r;
    asm("add.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ pf2 psub(pf2 a, pf2 b) {
    pf2 r;
    asm("sub.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ pf2 pmul(pf2 a, pf2 b) {
    pf2 r;
    asm("mul.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ pf2 pfma(pf2 a, pf2 b, pf2 c) {
    pf2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
// a * (c + j s)   = (a.x c - a.y s, a.y c + a.x s)
__device__ __forceinline__ pf2 pcmul(pf2 a, float c, float s) { return pfma(pswap(a), pk(-s, s), pmul(a, pbc(c))); }
// a * (c - j s)
__device__ __forceinline__ pf2 pcmulc(pf2 a, float c, float s) { return pfma(pswap(a), pk(s, -s), pmul(a, pbc(c))); }
// acc + a * b (full complex)
__device__ __forceinline__ pf2 pcfma(pf2 a, pf2 b, pf2 acc) {
    const float2 bb = unpk(b);
    return pfma(pswap(a), pk(-bb.y, bb.y), pfma(a, pbc(bb.x), acc));
}
// c + S j a   (S = +1 inverse transform, -1 forward)
template <int S>
__device__ __forceinline__ pf2 paddj(pf2 c, pf2 a) {
    return pfma(pswap(a), S > 0 ? pk(-1.f, 1.f) : pk(1.f, -1.f), c);
}
template <int S>
__device__ __forceinline__ pf2 pmulj(pf2 a) {
    return pmul(pswap(a), S > 0 ? pk(-1.f, 1.f) : pk(1.f, -1.f));
}

// multiply by W16^M with W16 = exp(S * 2 pi i / 16), M compile-time
template <int S, int M>
__device__ __forceinline__ pf2 pmulw16(pf2 a) {
    constexpr int m = M & 15;
    if (m == 0) return a;
    if (m == 4) return pmulj<S>(a);
    if (m == 8) return pmul(a, pbc(-1.f));
    if (m == 12) return pmulj<-S>(a);
    constexpr float C1 = 0.92387953251128675613f, S1 = 0.38268343236508977173f, H = 0.70710678118654752440f;
    constexpr float cs[16] = {1.f, C1, H, S1, 0.f, -S1, -H, -C1, -1.f, -C1, -H, -S1, 0.f, S1, H, C1};
    constexpr float sn[16] = {0.f, S1, H, C1, 1.f, C1, H, S1, 0.f, -S1, -H, -C1, -1.f, -C1, -H, -S1};
    return pcmul(a, cs[m], S > 0 ? sn[m] : -sn[m]);
}

// In-register 16-point DFT, natural order in and out: x[k] = sum_n x[n] W16^(n k).
//   PRUNE_IN : x[4..11] are known to be zero (values ignored)
//   PRUNE_OUT: only x[0..3] and x[12..15] are produced (others undefined)
template <int S, bool PRUNE_IN, bool PRUNE_OUT>
__device__ __forceinline__ void radix16(pf2 (&x)[16]) {
    pf2 y[16];  // y[b + 4c]
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        const pf2 a0 = x[b], a3 = x[b + 12];
        if (PRUNE_IN) {
            y[b + 0] = padd(a0, a3);
            y[b + 4] = paddj<-S>(a0, a3);
            y[b + 8] = psub(a0, a3);
            y[b + 12] = paddj<S>(a0, a3);
        } else {
            const pf2 a1 = x[b + 4], a2 = x[b + 8];
            const pf2 s02 = padd(a0, a2), d02 = psub(a0, a2);
            const pf2 s13 = padd(a1, a3), d13 = psub(a1, a3);
            y[b + 0] = padd(s02, s13);
            y[b + 4] = paddj<S>(d02, d13);
            y[b + 8] = psub(s02, s13);
            y[b + 12] = paddj<-S>(d02, d13);
        }
    }
    // twiddle W16^(b c)
    y[1 + 4] = pmulw16<S, 1>(y[1 + 4]);
    y[2 + 4] = pmulw16<S, 2>(y[2 + 4]);
    y[3 + 4] = pmulw16<S, 3>(y[3 + 4]);
    y[1 + 8] = pmulw16<S, 2>(y[1 + 8]);
    y[3 + 8] = pmulw16<S, 6>(y[3 + 8]);
    y[1 + 12] = pmulw16<S, 3>(y[1 + 12]);
    y[2 + 12] = pmulw16<S, 6>(y[2 + 12]);
    y[3 + 12] = pmulw16<S, 9>(y[3 + 12]);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const pf2 a0 = y[4 * c], a1 = y[4 * c + 1], a2 = y[4 * c + 2], a3 = y[4 * c + 3];
        pf2 s02, d02;
        if (c == 2) {  // y[2+8] carries W16^4 = S j: folded into the butterfly instead of a separate multiply
            s02 = paddj<S>(a0, a2);
            d02 = paddj<-S>(a0, a2);
        } else {
            s02 = padd(a0, a2);
            d02 = psub(a0, a2);
        }
        const pf2 s13 = padd(a1, a3), d13 = psub(a1, a3);
        x[c] = padd(s02, s13);            // d = 0
        x[c + 12] = paddj<-S>(d02, d13);  // d = 3
        if (!PRUNE_OUT) {
            x[c + 4] = paddj<S>(d02, d13);  // d = 1
            x[c + 8] = psub(s02, s13);      // d = 2
        }
    }
}

// Per-thread twiddles W4096^(t k), k = 1..15 (forward sign, exp(-2 pi i t k / 4096)); the inverse
// transform uses the conjugates.  Kept in registers for the whole frame.
struct Tw1 {
    float2 w[15];
    __device__ __forceinline__ float2 at(int k) const { return w[k - 1]; }
};

// Compact variant: only W^t, W^2t, W^4t, W^8t live in registers (8 instead of 30); the other eleven twiddles are
// products formed on the fly (11 complex multiplies per use).  Trades FMA work for registers / occupancy.
struct Tw1Compact {
    float2 w1, w2, w4, w8;
    __device__ __forceinline__ float2 at(int k) const {
        float2 r = make_float2(1.f, 0.f);
        bool first = true;
        if (k & 1) { r = w1; first = false; }
        if (k & 2) { r = first ? w2 : cmul(r, w2); first = false; }
        if (k & 4) { r = first ? w4 : cmul(r, w4); first = false; }
        if (k & 8) { r = first ? w8 : cmul(r, w8); first = false; }
        return r;
    }
};
__device__ __forceinline__ void load_tw1(Tw1Compact& tw, const float2* __restrict__ tw4096, int t) {
    tw.w1 = tw4096[t & (4096 - 1)];
    tw.w2 = tw4096[(2 * t) & (4096 - 1)];
    tw.w4 = tw4096[(4 * t) & (4096 - 1)];
    tw.w8 = tw4096[(8 * t) & (4096 - 1)];
}

__device__ __forceinline__ void load_tw1(Tw1& tw, const float2* __restrict__ tw4096, int t) {
#pragma unroll
    for (int k = 1; k < 16; ++k) tw.w[k - 1] = tw4096[(t * k) & (FFT_N - 1)];
}

// Synchronisation policy of the two shared-memory exchanges of a transform (a template hook: the microbenchmark
// tools/microbench/fft_phase_bench.cu instantiates the transform without barriers / without data movement to
// measure the FMA-pipe and LSU-pipe demands separately).  SyncCta: the CTA of 256 threads runs one transform.
struct SyncCta {
    static constexpr bool kExchange = true;
    static constexpr bool kFakeTw2 = false;  // timing experiments only: pass-2 twiddles taken from registers
    __device__ __forceinline__ void x_begin() const {}
    __device__ __forceinline__ void x_sync() const { __syncthreads(); }
    __device__ __forceinline__ void x_end() const {}
};

// One full transform.  bufA/bufB: two 4096-entry shared buffers; tw2: shared table
// W256^(n0 k1) at [k1*16 + n0] (forward sign).  x in: element j <-> index t + 256 j.
// x out: element k2 <-> bin t + 256 k2.  Exactly two __syncthreads; safe to call back to back
// (A/B alternate, see DESIGN.md "exchange buffers").
struct NoHook {
    __device__ __forceinline__ void operator()() const {}
};

// `before_pass3` runs after the loads of the second exchange have been issued and before the last radix-16 pass:
// the frame kernel issues the next antenna's global loads there, so that their latency hides behind that pass.
template <int S, bool PRUNE_IN, bool PRUNE_OUT, class SYNC = SyncCta, class HOOK = NoHook, class TW = Tw1>
__device__ __forceinline__ void fft4096(pf2 (&x)[16], pf2* __restrict__ bufA, pf2* __restrict__ bufB,
                                        const float2* __restrict__ tw2, const TW& tw1, const int t,
                                        const SYNC sy = SYNC(), const HOOK before_pass3 = HOOK()) {
    // pass 1
    if (PRUNE_IN) {
        if ((t >> 5) == 0) radix16<S, false, false>(x);  // warp 0 holds the stray bin 1024 in x[4] of thread 0
        else radix16<S, true, false>(x);
    } else {
        radix16<S, false, false>(x);
    }
#pragma unroll
    for (int k = 1; k < 16; ++k) {
        const float2 w = tw1.at(k);
        x[k] = (S > 0) ? pcmulc(x[k], w.x, w.y) : pcmul(x[k], w.x, w.y);
    }
    sy.x_begin();
    if (SYNC::kExchange) {
#pragma unroll
        for (int k = 0; k < 16; ++k) bufA[k * 256 + t] = x[k];
    }
    sy.x_sync();
    const int hi = t >> 4, lo = t & 15;
    if (SYNC::kExchange) {
#pragma unroll
        for (int n1 = 0; n1 < 16; ++n1) x[n1] = bufA[hi * 256 + n1 * 16 + lo];
    }
    sy.x_end();
    // pass 2
    radix16<S, false, false>(x);
#pragma unroll
    for (int k = 1; k < 16; ++k) {
        const float2 w = SYNC::kFakeTw2 ? tw1.at(k) : tw2[k * 16 + lo];
        x[k] = (S > 0) ? pcmulc(x[k], w.x, w.y) : pcmul(x[k], w.x, w.y);
    }
    sy.x_begin();
    if (SYNC::kExchange) {
        const int base = hi * 16 + (lo ^ hi);
#pragma unroll
        for (int k = 0; k < 16; ++k) bufB[k * 256 + base] = x[k];
    }
    sy.x_sync();
    if (SYNC::kExchange) {
        const int base = hi * 256 + lo * 16;
#pragma unroll
        for (int n0 = 0; n0 < 16; ++n0) x[n0] = bufB[base + (n0 ^ lo)];
    }
    sy.x_end();
    before_pass3();
    // pass 3
    if (PRUNE_OUT) {
        if ((t >> 5) == 0) radix16<S, false, false>(x);  // thread 0 also needs k2 = 4 (bin 1024)
        else radix16<S, false, true>(x);
    } else {
        radix16<S, false, false>(x);
    }
}

}  // namespace mmimo
