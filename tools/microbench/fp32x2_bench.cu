// Microbenchmark: issue / pipe throughput of scalar FP32 ops vs the sm_100 packed f32x2 forms (FADD2/FMUL2/FFMA2).
// Prints warp-instructions per clock per SM and the implied fp32 lane-ops per clock per SM.
#include <cuda_runtime.h>
#include <stdio.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm volatile("add.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm volatile("mul.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float a, float b) {
    constexpr int N = 16;
    float x[N]; u64 y[N]; unsigned z[N];
    __shared__ unsigned sm[1024];
    u64* sm64 = reinterpret_cast<u64*>(sm);
    for (int i = threadIdx.x; i < 1024; i += 256) sm[i] = i;
    __syncthreads();
    for (int i = 0; i < N; ++i) z[i] = threadIdx.x * 7 + i;
    for (int i = 0; i < N; ++i) { x[i] = threadIdx.x + i; float2 t = make_float2(x[i], x[i] + 1); y[i] = *reinterpret_cast<u64*>(&t); }
    float2 ab = make_float2(a, b); u64 ab2 = *reinterpret_cast<u64*>(&ab);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
#pragma unroll
            for (int i = 0; i < N; ++i) {
                if (MODE == 0) x[i] = fmaf(x[i], a, b);
                if (MODE == 1) x[i] = x[i] + a;
                if (MODE == 2) y[i] = fma2(y[i], ab2, ab2);
                if (MODE == 3) y[i] = add2(y[i], ab2);
                if (MODE == 4) y[i] = mul2(y[i], ab2);
                if (MODE == 5) { if (i & 1) y[i] = add2(y[i], ab2); else x[i] = fmaf(x[i], a, b); }   // mix FADD2 + FFMA
                if (MODE == 6) { if (i & 1) y[i] = fma2(y[i], ab2, ab2); else x[i] = x[i] + a; }       // mix FFMA2 + FADD
                if (MODE == 7) { if (i & 1) y[i] = fma2(y[i], ab2, ab2); else z[i] = (z[i] ^ (unsigned)it) + 0x9e3779b9u; }  // FFMA2 + (LOP3, IADD)
                if (MODE == 8) { if (i & 1) y[i] = fma2(y[i], ab2, ab2); else z[i] ^= sm[(threadIdx.x + z[i]) & 255]; }    // FFMA2 + LDS (+LOP3s)
                if (MODE == 9) { y[i] = fma2(y[i], ab2, ab2); sm[(threadIdx.x + 32 * i) & 1023] = (unsigned)it; }           // FFMA2 + STS 1:1
                if (MODE == 11) { y[i] = fma2(y[i], ab2, ab2); y[(i + 1) & 15] = fma2(y[(i + 1) & 15], ab2, ab2); sm[(threadIdx.x + 32 * i) & 1023] = (unsigned)it; }  // 2 FFMA2 : 1 STS.32
                if (MODE == 12) { y[i] = fma2(y[i], ab2, ab2); y[(i + 1) & 15] = fma2(y[(i + 1) & 15], ab2, ab2); y[(i + 2) & 15] = fma2(y[(i + 2) & 15], ab2, ab2); y[(i + 3) & 15] = fma2(y[(i + 3) & 15], ab2, ab2); sm64[(threadIdx.x + 32 * i) & 511] = y[i]; }  // 4 FFMA2 : 1 STS.64
                if (MODE == 13) { y[i] = fma2(y[i], ab2, ab2); y[(i + 1) & 15] = fma2(y[(i + 1) & 15], ab2, ab2); y[(i + 2) & 15] = fma2(y[(i + 2) & 15], ab2, ab2); y[(i + 3) & 15] = fma2(y[(i + 3) & 15], ab2, ab2); }  // 4 FFMA2 only
                if (MODE == 14) { sm64[(threadIdx.x + 32 * i) & 511] = y[i]; y[i] += 1; }  // STS.64 only
                if (MODE == 10) { sm[(threadIdx.x + 32 * i) & 1023] = (unsigned)it; }                                        // STS only
            }
        }
    }
    float s = 0; for (int i = 0; i < N; ++i) { float2 t = *reinterpret_cast<float2*>(&y[i]); s += x[i] + t.x + t.y + (float)z[i]; }
    s += (float)sm[threadIdx.x];
    if (s == 123.456f) out[0] = s;
}
template <int MODE> void run(const char* name, float* d, int nsm, double lanes_per_inst) {
    const int iters = 2048, grid = nsm * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0); k<MODE><<<grid, 256>>>(d, iters, 0.999f, 0.001f); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep) best = ms < best ? ms : best;
    }
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double winst = 16.0 * 8 * iters * 8.0 * grid;       // warp instructions
    const double cyc = best * 1e-3 * clk * 1e3;
    printf("%-22s %8.3f ms  %6.2f warp-inst/clk/SM  %7.1f fp32-lane-ops/clk/SM\n", name, best, winst / cyc / nsm, winst / cyc / nsm * 32 * lanes_per_inst);
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs, clockRate %d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
    float* d; cudaMalloc(&d, 64);
    run<0>("FFMA", d, p.multiProcessorCount, 1);
    run<1>("FADD", d, p.multiProcessorCount, 1);
    run<2>("FFMA2", d, p.multiProcessorCount, 2);
    run<3>("FADD2", d, p.multiProcessorCount, 2);
    run<4>("FMUL2", d, p.multiProcessorCount, 2);
    run<5>("FADD2+FFMA 1:1", d, p.multiProcessorCount, 1.5);
    run<6>("FFMA2+FADD 1:1", d, p.multiProcessorCount, 1.5);
    run<7>("FFMA2 + (LOP3+IADD) 8:8x2", d, p.multiProcessorCount, 1);
    run<8>("FFMA2 + LDS chain 8:8", d, p.multiProcessorCount, 1);
    run<9>("FFMA2 + STS 16:16", d, p.multiProcessorCount, 1);
    run<10>("STS only 16", d, p.multiProcessorCount, 1);
    run<11>("2 FFMA2 : 1 STS.32 (balanced)", d, p.multiProcessorCount, 1);
    run<12>("4 FFMA2 : 1 STS.64 (balanced)", d, p.multiProcessorCount, 1);
    run<13>("4 FFMA2 only", d, p.multiProcessorCount, 1);
    run<14>("STS.64 only", d, p.multiProcessorCount, 1);
    return 0;
}
