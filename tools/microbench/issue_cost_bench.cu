// Issue-slot cost of memory / integer instructions next to packed-fp32 arithmetic on sm_100a.
// 2 CTAs x 256 threads per SM (4 warps per SM sub-partition, as in the frame kernel).  Per loop iteration and
// thread: 64 independent FFMA2 plus M instructions of one other kind; M is small enough that the shared-memory
// data path (128 B/clk/SM) is never the bound.  Prints SM sub-partition cycles per iteration per warp and the
// implied cost of the added instruction in issue cycles.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o issue_cost_bench issue_cost_bench.cu
#include <cuda_runtime.h>
#include <stdio.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ void sts64(unsigned a, u64 v) { asm volatile("st.shared.b64 [%0], %1;" ::"r"(a), "l"(v) : "memory"); }
__device__ __forceinline__ void sts32(unsigned a, unsigned v) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void sts128(unsigned a, u64 v, u64 w) { asm volatile("st.shared.v2.b64 [%0], {%1, %2};" ::"r"(a), "l"(v), "l"(w) : "memory"); }
__device__ __forceinline__ u64 lds64(unsigned a) { u64 v; asm volatile("ld.shared.b64 %0, [%1];" : "=l"(v) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ unsigned lds32(unsigned a) { unsigned v; asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ void lds128(unsigned a, u64& v, u64& w) { asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(v), "=l"(w) : "r"(a) : "memory"); }

enum { NONE, STS32, STS64, STS128, LDS32, LDS64, LDS128, LOP, FADD, LDS64_BCAST, SHFL };

template <int KIND, int M>
__global__ void __launch_bounds__(256, 2) k(float* out, int iters, float a, float b) {
    extern __shared__ unsigned char smem[];
    const unsigned base = (unsigned)__cvta_generic_to_shared(smem);
    const int t = threadIdx.x;
    u64 y[16], m[16], w64[16];
    unsigned z[16];
    float f[16];
    for (int i = 0; i < 16; ++i) {
        float2 v = make_float2((float)(t + i), (float)(t - i));
        y[i] = *reinterpret_cast<u64*>(&v);
        float2 w = make_float2(a, b);
        m[i] = *reinterpret_cast<u64*>(&w);
        z[i] = t * 7 + i;
        w64[i] = i;
        f[i] = (float)i;
    }
    for (int i = t; i < 8192; i += 256) reinterpret_cast<float2*>(smem)[i] = make_float2(a, b);
    __syncthreads();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                y[i] = fma2(y[i], m[i], m[(i + 1) & 15]);
                const int slot = r * 16 + i;  // 0..63
                if (slot % (64 / (M ? M : 1)) == 0 && M) {
                    const int q = slot / (64 / M);  // 0..M-1
                    if (KIND == STS32) sts32(base + (q * 256 + t) * 4, z[i]);
                    if (KIND == STS64) sts64(base + (q * 256 + t) * 8, m[i]);
                    if (KIND == STS128) sts128(base + (q * 256 + t) * 16, m[i], m[(i + 8) & 15]);
                    if (KIND == LDS32) z[q & 15] = lds32(base + (q * 256 + t) * 4);
                    if (KIND == LDS64) w64[q & 15] = lds64(base + (q * 256 + t) * 8);
                    if (KIND == LDS64_BCAST) w64[q & 15] = lds64(base + (q * 16 + (t & 15)) * 8);
                    if (KIND == LDS128) lds128(base + (q * 256 + t) * 16, w64[q & 15], w64[(q + 8) & 15]);
                    if (KIND == LOP) z[q & 15] = (z[q & 15] ^ (unsigned)it) + 0x9e3779b9u;
                    if (KIND == FADD) f[q & 15] += a;
                    if (KIND == SHFL) z[q & 15] = __shfl_xor_sync(0xffffffffu, z[q & 15], 5);
                }
            }
        }
    }
    float s = 0;
    for (int i = 0; i < 16; ++i) {
        float2 v = *reinterpret_cast<float2*>(&y[i]);
        float2 w = *reinterpret_cast<float2*>(&m[i]);
        s += v.x + v.y + w.x + w.y + (float)z[i] + f[i] + (float)w64[i];
    }
    if (s == 123.456f) out[0] = s;
}

static double g_base = 0;
template <int KIND, int M>
static void run(const char* name, float* d, int nsm, int n_inst_per_op) {
    const int iters = 4096, smem = 65536;
    cudaFuncSetAttribute(k<KIND, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        k<KIND, M><<<nsm * 2, 256, smem>>>(d, iters, 0.999f, 0.001f);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep) best = fminf(best, ms);
    }
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double cyc = best * 1e-3 * clk * 1e3 / iters;  // SM cycles per iteration (4 warps per sub-partition)
    const double per_warp = cyc / 4;
    if (KIND == NONE) g_base = per_warp;
    printf("%-28s M=%2d  %8.1f cycles/iter/warp", name, M, per_warp);
    if (KIND != NONE) printf("   added cost %.2f cycles per op (%d instr)", (per_warp - g_base) / M, n_inst_per_op);
    printf("   [%s]\n", cudaGetErrorString(cudaGetLastError()));
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs, clockRate %d kHz; 64 FFMA2 per iteration per thread = 128 issue cycles\n", p.name, p.multiProcessorCount, p.clockRate);
    float* d; cudaMalloc(&d, 64);
    const int n = p.multiProcessorCount;
    run<NONE, 0>("64 FFMA2 only", d, n, 0);
    run<LOP, 8>("+ LOP3,IADD", d, n, 2);   run<LOP, 16>("+ LOP3,IADD", d, n, 2);
    run<FADD, 8>("+ FADD", d, n, 1);       run<FADD, 16>("+ FADD", d, n, 1);
    run<STS32, 4>("+ STS.32", d, n, 1);    run<STS32, 8>("+ STS.32", d, n, 1);
    run<STS64, 4>("+ STS.64", d, n, 1);    run<STS64, 8>("+ STS.64", d, n, 1);
    run<STS128, 2>("+ STS.128", d, n, 1);  run<STS128, 4>("+ STS.128", d, n, 1);
    run<LDS32, 4>("+ LDS.32", d, n, 1);    run<LDS32, 8>("+ LDS.32", d, n, 1);
    run<LDS64, 4>("+ LDS.64", d, n, 1);    run<LDS64, 8>("+ LDS.64", d, n, 1);
    run<LDS64_BCAST, 4>("+ LDS.64 (half-warp dup)", d, n, 1); run<LDS64_BCAST, 8>("+ LDS.64 (half-warp dup)", d, n, 1);
    run<LDS128, 2>("+ LDS.128", d, n, 1);  run<LDS128, 4>("+ LDS.128", d, n, 1);
    run<LDS64, 16>("+ LDS.64", d, n, 1);   run<STS64, 16>("+ STS.64", d, n, 1); run<LDS32, 16>("+ LDS.32", d, n, 1); run<LDS32, 32>("+ LDS.32", d, n, 1);
    run<SHFL, 4>("+ SHFL.32", d, n, 1);    run<SHFL, 8>("+ SHFL.32", d, n, 1);
    return 0;
}
