// Shared device helpers of the link kernels: native slot order, QAM mapping / slicing,
// PA models, Philox4x32-10.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "fft4096.cuh"

namespace mmimo {

// ---------------------------------------------------------------------------------------------
// Native slot order of all per-sub-carrier device arrays (n_fft = 4096, n_sub_carr = 2048).
//   slot p = jp*256 + t  (jp = 0..7, t = 0..255)  <->  FFT register j = (jp<4 ? jp : jp+8) of
//   thread t  <->  FFT bin t + 256 j, except slot 0 which carries bin 1024 (bin 0 = DC is unused,
//   modulation.py:264-267).  Data-sub-carrier index ("negative half then positive half"):
//   s = bin-3072 for bin >= 3072, s = 1023+bin for 1 <= bin <= 1024.
// ---------------------------------------------------------------------------------------------
constexpr int N_SC = 2048;
constexpr int SLOTS_PER_THREAD = 8;

__host__ __device__ __forceinline__ int slot_to_bin(int p) {
    if (p == 0) return 1024;
    const int jp = p >> 8, t = p & 255;
    return t + 256 * (jp < 4 ? jp : jp + 8);
}
__host__ __device__ __forceinline__ int bin_to_sc(int bin) { return bin >= 3072 ? bin - 3072 : 1023 + bin; }
__host__ __device__ __forceinline__ int slot_to_sc(int p) { return bin_to_sc(slot_to_bin(p)); }
__host__ __device__ __forceinline__ int sc_to_slot(int s) {
    const int bin = s < 1024 ? s + 3072 : s - 1023;
    if (bin == 1024) return 0;
    const int j = bin >> 8, t = bin & 255;
    return (j < 4 ? j : j - 8) * 256 + t;
}
__device__ __forceinline__ constexpr int jp_to_j(int jp) { return jp < 4 ? jp : jp + 8; }

// ---------------------------------------------------------------------------------------------
// Square QAM, Gray-reordered table of modulation.py:230-244 / :110-114 in closed form:
// symbol of integer b is base[gray_decode(b)], base[n] = pam[n/side] + 1j*zig(n).
// ---------------------------------------------------------------------------------------------
struct Qam {
    int log2side;  // 1,2,3,4 for 4/16/64/256-QAM
    __device__ __forceinline__ int side() const { return 1 << log2side; }
};

__device__ __forceinline__ float2 qam_symbol(const Qam q, uint32_t b) {
    uint32_t n = b;
    n ^= n >> 1;
    n ^= n >> 2;
    n ^= n >> 4;  // gray decode (<= 8 bits)
    const int side = q.side();
    const int li = n >> q.log2side, row = n & (side - 1);
    const int qi = (li & 1) ? (side - 1 - row) : row;
    return make_float2((float)(2 * li - side + 1), (float)(2 * qi - side + 1));
}

// nearest constellation point by per-axis slicing == argmin of modulation.py:145 away from exact ties
__device__ __forceinline__ uint32_t qam_detect(const Qam q, float2 r) {
    const int side = q.side();
    const float fs = (float)side;
    int li = (int)floorf((r.x + fs) * 0.5f);
    int qi = (int)floorf((r.y + fs) * 0.5f);
    li = min(max(li, 0), side - 1);
    qi = min(max(qi, 0), side - 1);
    const int row = (li & 1) ? (side - 1 - qi) : qi;
    const uint32_t n = (uint32_t)((li << q.log2side) + row);
    return n ^ (n >> 1);
}

// ---------------------------------------------------------------------------------------------
// PA models, distortion.py:9-19 (soft limiter), :102-113 (Rapp), :202-211 (third order)
// ---------------------------------------------------------------------------------------------
enum PaKind : int { PA_NONE = 0, PA_SOFT_LIMITER = 1, PA_RAPP = 2, PA_TOI = 3 };

struct PaParams {
    int kind;
    float p0;  // soft limiter / Rapp: saturation power; TOI: cubic coefficient
    float p1;  // Rapp: hardness p
};

template <int KIND>
__device__ __forceinline__ float2 pa_apply(float2 x, const float p0, const float sqrt_p0, const float p1) {
    if (KIND == PA_SOFT_LIMITER) {
        const float p = fmaf(x.x, x.x, x.y * x.y);
        if (p > p0) {
            const float g = sqrt_p0 * rsqrtf(p);
            x.x *= g;
            x.y *= g;
        }
        return x;
    } else if (KIND == PA_RAPP) {
        const float p = fmaf(x.x, x.x, x.y * x.y);
        // (|x|/sqrt(Psat))^(2p) = (p/Psat)^p ; gain = (1 + that)^(-1/(2p))
        const float r = __powf(p / p0, p1);
        const float g = __powf(1.0f + r, -0.5f / p1);
        return make_float2(x.x * g, x.y * g);
    } else if (KIND == PA_TOI) {
        const float p = fmaf(x.x, x.x, x.y * x.y);
        const float g = 1.0f - p0 * p;
        return make_float2(x.x * g, x.y * g);
    }
    return x;
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 counter-based RNG (Salmon et al. 2011), one 128-bit block per call.
// ---------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                       uint32_t k1, uint32_t (&out)[4]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// cos / sin of (pi/2)(n + u), |u| <= 1/2: two degree-3 polynomials in u^2 (max error 1.3e-7), the quadrant n by sign flips
__device__ __forceinline__ float2 cis_quadrant(const int n, const float u) {
    const float z = u * u;
    const float sn = u * fmaf(z, fmaf(z, fmaf(z, -0.004604428540915251f, 0.0796811580657959f), -0.6459636092185974f), 1.5707963705062866f);
    const float cs = fmaf(z, fmaf(z, fmaf(z, -0.02042119763791561f, 0.2536039352416992f), -1.2336976528167725f), 1.0f);
    const bool sw = n & 1;
    return make_float2(__int_as_float(__float_as_int(sw ? sn : cs) ^ (((n + 1) & 2) << 30)),
                       __int_as_float(__float_as_int(sw ? cs : sn) ^ ((n & 2) << 30)));
}

// two 32-bit uniforms -> one CN(0,1)-shaped pair (each component N(0,1)), Box-Muller (the frame kernels' noise)
__device__ __forceinline__ float2 box_muller(uint32_t u0, uint32_t u1) {
    const float a = ((float)u0 + 1.0f) * 2.3283064365386963e-10f;  // (0,1]
    const float r = sqrtf(-2.0f * logf(a));
    float s, c;
    sincospif((float)u1 * (2.0f * 2.3283064365386963e-10f), &s, &c);
    return make_float2(r * c, r * s);
}
// the same for the kernels that draw K x 2048 normals per realisation (Rayleigh coefficients, CSI error): radius from
// MUFU.LG2 / MUFU.SQRT, angle 2 pi u1 / 2^32 reduced in integer arithmetic (nearest quadrant + remainder in [-1/2, 1/2)),
// ~25 instructions.  (In the frame kernel the draw is 1/3000 of a frame and the precise form keeps its register allocation.)
__device__ __forceinline__ float2 box_muller_fast(uint32_t u0, uint32_t u1) {
    const float a = ((float)u0 + 1.0f) * 2.3283064365386963e-10f;  // (0,1]
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(-1.3862943611198906f * __log2f(a)));  // sqrt(-2 ln a)
    const uint32_t w = u1 + 0x20000000u;  // + 1/8 turn: w >> 30 is the NEAREST quadrant (mod 4)
    const float u = (float)((int)(w & 0x3fffffffu) - 0x20000000) * 9.313225746154785e-10f;  // [-1/2, 1/2)
    const float2 cs = cis_quadrant((int)(w >> 30), u);
    return make_float2(r * cs.x, r * cs.y);
}

// RNG stream ids (counter word c2)
enum RngStream : uint32_t { RNG_BITS = 1, RNG_NOISE = 2, RNG_POS = 3, RNG_CLEAN_BITS = 4, RNG_CLEAN_NOISE = 5,
                            RNG_RAYLEIGH = 6, RNG_CSI = 7 };

}  // namespace mmimo
