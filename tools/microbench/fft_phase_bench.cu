// Microbenchmark of the two resources of the fused frame kernel's antenna chain (IFFT -> PA -> FFT on 16 points per
// thread): the butterfly arithmetic (FMA pipe) and the shared-memory exchanges (LSU pipe), alone and together, with
// 1 and 2 CTAs of 256 threads per SM.  Prints SM cycles per antenna chain per CTA.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o fft_phase_bench fft_phase_bench.cu
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <vector>

#include "../../m_mimo_ofdm_with_nonlinear_pa_sim_b200/csrc/frame_kernel.cuh"
using namespace mmimo;

struct SyncFakeTw2 : SyncCta {  // full chain, but pass-2 twiddles from registers (wrong values, timing only)
    static constexpr bool kFakeTw2 = true;
};
struct SyncNone {  // compute only: no data movement, no barriers
    static constexpr bool kFakeTw2 = false;
    static constexpr bool kExchange = false;
    __device__ __forceinline__ void x_begin() const {}
    __device__ __forceinline__ void x_sync() const {}
    __device__ __forceinline__ void x_end() const {}
};

struct SyncNoBar {  // data movement without any barrier (results are wrong; timing experiment only)
    static constexpr bool kFakeTw2 = false;
    static constexpr bool kExchange = true;
    __device__ __forceinline__ void x_begin() const {}
    __device__ __forceinline__ void x_sync() const {}
    __device__ __forceinline__ void x_end() const {}
};

// MODE 4: full chain without barriers
// MODE 0: full chain; 1: compute only; 2: exchange only (same addressing, 4 exchanges); 3: exchange without barriers
template <int MODE>
__global__ void __launch_bounds__(256, 2) k(float2* out, const float2* tw4096, const float2* tw256, int iters, int delay = 0) {
    extern __shared__ float2 smem[];
    pf2* bufA = reinterpret_cast<pf2*>(smem);
    pf2* bufB = reinterpret_cast<pf2*>(smem + 4096);
    float2* tw2 = smem + 8192;
    const int t = threadIdx.x;
    tw2[t] = tw256[t];
    Tw1 tw1;
    load_tw1(tw1, tw4096, t);
    __syncthreads();
    pf2 x[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = pk((float)(t + j) * 1e-3f, (float)(t - j) * 1e-3f);
    const int hi = t >> 4, lo = t & 15;
    if (delay > 0 && blockIdx.x >= gridDim.x / 2) {  // stagger the second CTA of every SM
        const long long t0 = clock64();
        while (clock64() - t0 < delay) {}
    }
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
            fft4096<+1, true, false>(x, bufA, bufB, tw2, tw1, t);
#pragma unroll
            for (int j = 0; j < 16; ++j) x[j] = pa_apply_pk<PA_SOFT_LIMITER>(x[j], 1.0f, 1.0f, 0.f);
            fft4096<-1, false, true>(x, bufA, bufB, tw2, tw1, t);
#pragma unroll
            for (int j = 4; j < 12; ++j) x[j] = pmul(x[j - 4], pbc(1e-3f));
        } else if (MODE == 5) {
            fft4096<+1, true, false, SyncFakeTw2>(x, bufA, bufB, tw2, tw1, t);
#pragma unroll
            for (int j = 0; j < 16; ++j) x[j] = pa_apply_pk<PA_SOFT_LIMITER>(x[j], 1.0f, 1.0f, 0.f);
            fft4096<-1, false, true, SyncFakeTw2>(x, bufA, bufB, tw2, tw1, t);
#pragma unroll
            for (int j = 4; j < 12; ++j) x[j] = pmul(x[j - 4], pbc(1e-3f));
        } else if (MODE == 4) {
            fft4096<+1, true, false, SyncNoBar>(x, bufA, bufB, tw2, tw1, t);
#pragma unroll
            for (int j = 0; j < 16; ++j) x[j] = pa_apply_pk<PA_SOFT_LIMITER>(x[j], 1.0f, 1.0f, 0.f);
            fft4096<-1, false, true, SyncNoBar>(x, bufA, bufB, tw2, tw1, t);
#pragma unroll
            for (int j = 4; j < 12; ++j) x[j] = pmul(x[j - 4], pbc(1e-3f));
        } else if (MODE == 1) {
            fft4096<+1, true, false, SyncNone>(x, bufA, bufB, tw2, tw1, t);
#pragma unroll
            for (int j = 0; j < 16; ++j) x[j] = pa_apply_pk<PA_SOFT_LIMITER>(x[j], 1.0f, 1.0f, 0.f);
            fft4096<-1, false, true, SyncNone>(x, bufA, bufB, tw2, tw1, t);
#pragma unroll
            for (int j = 4; j < 12; ++j) x[j] = pmul(x[j - 4], pbc(1e-3f));
        } else {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
#pragma unroll
                for (int k2 = 0; k2 < 16; ++k2) bufA[k2 * 256 + t] = x[k2];
                if (MODE == 2) __syncthreads();
#pragma unroll
                for (int n1 = 0; n1 < 16; ++n1) x[n1] = bufA[hi * 256 + n1 * 16 + lo];
                {
                    const int base = hi * 16 + (lo ^ hi);
#pragma unroll
                    for (int k2 = 0; k2 < 16; ++k2) bufB[k2 * 256 + base] = x[k2];
                }
                if (MODE == 2) __syncthreads();
                {
                    const int base = hi * 256 + lo * 16;
#pragma unroll
                    for (int n0 = 0; n0 < 16; ++n0) x[n0] = bufB[base + (n0 ^ lo)];
                }
            }
        }
    }
    float2 s = make_float2(0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 16; ++j) { const float2 v = unpk(x[j]); s.x += v.x; s.y += v.y; }
    if (s.x == 123.456f) out[0] = s;
}

// ---- ping-pong of two 256-thread groups in ONE 512-thread CTA: the shared-memory exchanges of the two groups are
// serialised by a token (mbarrier pair), so that one group's exchange runs behind the other group's butterflies.
// mbarrier waits are real spin loops (basic-block boundaries ptxas does not schedule across); the state returned by the
// releasing arrive is folded into the data (token & zero) so that the arrive cannot sink below the next butterflies.
__device__ __forceinline__ unsigned long long mbar_arrive(unsigned addr) {
    unsigned long long st;
    asm volatile("mbarrier.arrive.shared::cta.b64 %0, [%1];" : "=l"(st) : "r"(addr) : "memory");
    return st;
}
__device__ __forceinline__ void mbar_wait(unsigned addr, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@!p bra WAIT_LOOP;\n"
        "}\n" ::"r"(addr), "r"(parity) : "memory");
}
struct SyncPingPong {
    static constexpr bool kExchange = true;
    static constexpr bool kFakeTw2 = false;
    unsigned my_tok, other_tok;   // shared addresses of the mbarriers: wait on mine, release the other group's
    unsigned* parity;             // phase parity of my token (per-thread register copy)
    int bar_id;                   // named barrier of my group
    unsigned zero;                // runtime zero (kernel argument)
    mutable unsigned fold;        // accumulates (token & zero); OR-ed into the data by the caller
    __device__ __forceinline__ void x_begin() const {
        mbar_wait(my_tok, *parity);
        *parity ^= 1u;
    }
    __device__ __forceinline__ void x_sync() const { asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory"); }
    __device__ __forceinline__ void x_end() const {
        const unsigned long long st = mbar_arrive(other_tok);
        fold |= (unsigned)st & zero;
    }
};

__global__ void __launch_bounds__(512, 1) kpp(float2* out, const float2* tw4096, const float2* tw256, int iters, unsigned zero, int mode) {
    extern __shared__ float2 smem[];
    __shared__ unsigned long long mbar[2];
    const int g = threadIdx.x >> 8, t = threadIdx.x & 255;
    pf2* bufA = reinterpret_cast<pf2*>(smem) + g * 8192;
    pf2* bufB = bufA + 4096;
    float2* tw2 = smem + 16384;
    if (threadIdx.x < 256) tw2[t] = tw256[t];
    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            const unsigned a = (unsigned)__cvta_generic_to_shared(&mbar[i]);
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(256));
        }
    }
    Tw1 tw1;
    load_tw1(tw1, tw4096, t);
    __syncthreads();
    pf2 x[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = pk((float)(t + j) * 1e-3f, (float)(t - j) * 1e-3f);
    unsigned parity = g == 0 ? 1u : 0u;  // group 0 owns the token first (parity 1 of a fresh barrier is complete)
    SyncPingPong sy{(unsigned)__cvta_generic_to_shared(&mbar[g]), (unsigned)__cvta_generic_to_shared(&mbar[g ^ 1]), &parity, 1 + g, zero, 0u};
    for (int it = 0; it < iters; ++it) {
        fft4096<+1, true, false, SyncPingPong>(x, bufA, bufB, tw2, tw1, t, sy);
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = pa_apply_pk<PA_SOFT_LIMITER>(x[j], 1.0f, 1.0f, 0.f);
        fft4096<-1, false, true, SyncPingPong>(x, bufA, bufB, tw2, tw1, t, sy);
#pragma unroll
        for (int j = 4; j < 12; ++j) x[j] = pmul(x[j - 4], pbc(1e-3f));
        x[0] = x[0] | (unsigned long long)sy.fold;
    }
    // let the partner finish: release once more so that nobody spins forever
    mbar_arrive(sy.other_tok);
    float2 s = make_float2(0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 16; ++j) { const float2 v = unpk(x[j]); s.x += v.x; s.y += v.y; }
    if (s.x == 123.456f) out[0] = s;
}

// co-residency experiment: half of the CTAs run the butterflies only, the other half the exchanges only; each CTA
// reports its own duration (clock64) and SM id.  cls: 0 = mixed, 1 = all compute, 2 = all exchange
__global__ void __launch_bounds__(256, 2) kmix(long long* dur, int* smid, const float2* tw4096, const float2* tw256, int iters, int cls) {
    extern __shared__ float2 smem[];
    pf2* bufA = reinterpret_cast<pf2*>(smem);
    pf2* bufB = reinterpret_cast<pf2*>(smem + 4096);
    float2* tw2 = smem + 8192;
    const int t = threadIdx.x;
    tw2[t] = tw256[t];
    Tw1 tw1;
    load_tw1(tw1, tw4096, t);
    __syncthreads();
    pf2 x[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = pk((float)(t + j) * 1e-3f, (float)(t - j) * 1e-3f);
    const int hi = t >> 4, lo = t & 15;
    const bool compute = cls == 1 || (cls == 0 && blockIdx.x < gridDim.x / 2);
    const long long t0 = clock64();
    if (compute) {
        for (int it = 0; it < iters; ++it) {
            fft4096<+1, true, false, SyncNone>(x, bufA, bufB, tw2, tw1, t);
#pragma unroll
            for (int j = 0; j < 16; ++j) x[j] = pa_apply_pk<PA_SOFT_LIMITER>(x[j], 1.0f, 1.0f, 0.f);
            fft4096<-1, false, true, SyncNone>(x, bufA, bufB, tw2, tw1, t);
#pragma unroll
            for (int j = 4; j < 12; ++j) x[j] = pmul(x[j - 4], pbc(1e-3f));
        }
    } else {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
#pragma unroll
                for (int k2 = 0; k2 < 16; ++k2) bufA[k2 * 256 + t] = x[k2];
                __syncthreads();
#pragma unroll
                for (int n1 = 0; n1 < 16; ++n1) x[n1] = bufA[hi * 256 + n1 * 16 + lo];
                {
                    const int base = hi * 16 + (lo ^ hi);
#pragma unroll
                    for (int k2 = 0; k2 < 16; ++k2) bufB[k2 * 256 + base] = x[k2];
                }
                __syncthreads();
                {
                    const int base = hi * 256 + lo * 16;
#pragma unroll
                    for (int n0 = 0; n0 < 16; ++n0) x[n0] = bufB[base + (n0 ^ lo)];
                }
            }
        }
    }
    const long long t1 = clock64();
    float2 s = make_float2(0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 16; ++j) { const float2 v = unpk(x[j]); s.x += v.x; s.y += v.y; }
    if (t == 0) {
        dur[blockIdx.x] = t1 - t0 + (s.x == 123.456f ? 1 : 0);
        unsigned id; asm("mov.u32 %0, %%smid;" : "=r"(id));
        smid[blockIdx.x] = (int)id;
    }
}

template <int NCTA>
__global__ void __launch_bounds__(256, NCTA) kc(float2* out, const float2* tw4096, const float2* tw256, int iters) {
    extern __shared__ float2 smem[];
    pf2* bufA = reinterpret_cast<pf2*>(smem);
    pf2* bufB = reinterpret_cast<pf2*>(smem + 4096);
    float2* tw2 = smem + 8192;
    const int t = threadIdx.x;
    tw2[t] = tw256[t];
    Tw1Compact tw1;
    load_tw1(tw1, tw4096, t);
    __syncthreads();
    pf2 x[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = pk((float)(t + j) * 1e-3f, (float)(t - j) * 1e-3f);
    for (int it = 0; it < iters; ++it) {
        fft4096<+1, true, false, SyncCta, NoHook, Tw1Compact>(x, bufA, bufB, tw2, tw1, t);
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = pa_apply_pk<PA_SOFT_LIMITER>(x[j], 1.0f, 1.0f, 0.f);
        fft4096<-1, false, true, SyncCta, NoHook, Tw1Compact>(x, bufA, bufB, tw2, tw1, t);
#pragma unroll
        for (int j = 4; j < 12; ++j) x[j] = pmul(x[j - 4], pbc(1e-3f));
    }
    float2 s = make_float2(0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 16; ++j) { const float2 v = unpk(x[j]); s.x += v.x; s.y += v.y; }
    if (s.x == 123.456f) out[0] = s;
}
template <int NCTA>
static void run_c(float2* d, const float2* t4, const float2* t2, int nsm) {
    const int iters = 2000, smem = (8192 + 256) * 8;
    cudaFuncSetAttribute(kc<NCTA>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    int occ = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kc<NCTA>, 256, smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        kc<NCTA><<<nsm * NCTA, 256, smem>>>(d, t4, t2, iters);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep) best = fminf(best, ms);
    }
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double cyc = best * 1e-3 * clk * 1e3;
    printf("%-34s %d CTA/SM (occupancy %d)  %8.3f ms  %8.0f SM-cycles/chain\n", "full chain, compact twiddles", NCTA, occ, best, cyc / iters / NCTA);
}

template <int MODE>
static void run(const char* name, float2* d, const float2* t4, const float2* t2, int nsm, int per_sm, int delay = 0) {
    const int iters = 2000, smem = (8192 + 256) * 8 + (per_sm == 1 ? 110 * 1024 : 0);  // extra smem forces 1 CTA/SM
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        k<MODE><<<nsm * per_sm, 256, smem>>>(d, t4, t2, iters, delay);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep) best = fminf(best, ms);
    }
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double cyc = best * 1e-3 * clk * 1e3;
    printf("%-34s %d CTA/SM  %8.3f ms  %8.0f cycles/chain/CTA  %8.0f SM-cycles/chain\n", name, per_sm, best, cyc / iters, cyc / iters / per_sm);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(e));
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs, clockRate %d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
    std::vector<float2> t4(4096), t2(256);
    for (int i = 0; i < 4096; ++i) t4[i] = make_float2((float)cos(-2 * M_PI * i / 4096), (float)sin(-2 * M_PI * i / 4096));
    for (int k1 = 0; k1 < 16; ++k1) for (int n0 = 0; n0 < 16; ++n0) t2[k1 * 16 + n0] = make_float2((float)cos(-2 * M_PI * k1 * n0 / 256), (float)sin(-2 * M_PI * k1 * n0 / 256));
    float2 *d, *d4, *d2;
    cudaMalloc(&d, 64); cudaMalloc(&d4, 4096 * 8); cudaMalloc(&d2, 256 * 8);
    cudaMemcpy(d4, t4.data(), 4096 * 8, cudaMemcpyHostToDevice); cudaMemcpy(d2, t2.data(), 256 * 8, cudaMemcpyHostToDevice);
    const int nsm = p.multiProcessorCount;
    for (int per_sm = 1; per_sm <= 2; ++per_sm) {
        run<0>("full chain (IFFT, PA, FFT)", d, d4, d2, nsm, per_sm);
        run<1>("butterflies only (no exchange)", d, d4, d2, nsm, per_sm);
        run<2>("4 exchanges only (STS, bar, LDS)", d, d4, d2, nsm, per_sm);
        run<3>("4 exchanges, no barrier", d, d4, d2, nsm, per_sm);
        run<4>("full chain, no barrier", d, d4, d2, nsm, per_sm);
    }
    run<5>("full chain, pass-2 twiddles in regs", d, d4, d2, nsm, 1);
    run<5>("full chain, pass-2 twiddles in regs", d, d4, d2, nsm, 2);
    {
        const int iters = 2000, smem = (8192 + 256) * 8, grid = nsm * 2;
        cudaFuncSetAttribute(kmix, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        long long* ddur; int* dsm;
        cudaMalloc(&ddur, grid * 8); cudaMalloc(&dsm, grid * 4);
        std::vector<long long> dur(grid); std::vector<int> sm(grid);
        for (int cls = 0; cls < 3; ++cls) {
            for (int rep = 0; rep < 2; ++rep) kmix<<<grid, 256, smem>>>(ddur, dsm, d4, d2, iters, cls);
            cudaMemcpy(dur.data(), ddur, grid * 8, cudaMemcpyDeviceToHost);
            cudaMemcpy(sm.data(), dsm, grid * 4, cudaMemcpyDeviceToHost);
            double a = 0, b = 0; int paired = 0;
            for (int i = 0; i < grid / 2; ++i) { a += dur[i]; b += dur[i + grid / 2]; }
            for (int i = 0; i < grid / 2; ++i) for (int j = grid / 2; j < grid; ++j) if (sm[i] == sm[j]) { ++paired; break; }
            printf("co-residency cls %d (0 mixed, 1 all compute, 2 all exchange): first half %8.0f cycles/chain/CTA, second half %8.0f; %d of %d first-half CTAs share an SM with a second-half CTA  [%s]\n",
                   cls, a / (grid / 2) / iters, b / (grid / 2) / iters, paired, grid / 2, cudaGetErrorString(cudaGetLastError()));
        }
    }
    {
        const int iters = 2000, smem = (16384 + 256) * 8;
        cudaFuncSetAttribute(kpp, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        float best = 1e30f;
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(e0);
            kpp<<<nsm, 512, smem>>>(d, d4, d2, iters, 0u, 0);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (rep) best = fminf(best, ms);
        }
        int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
        printf("ping-pong groups (mbarrier token, 512 threads, 1 CTA/SM): %8.3f ms  %8.0f SM-cycles/chain  [%s]\n", best,
               best * 1e-3 * clk * 1e3 / iters / 2, cudaGetErrorString(cudaGetLastError()));
    }
    run_c<2>(d, d4, d2, nsm);
    run_c<3>(d, d4, d2, nsm);
    return 0;
}
