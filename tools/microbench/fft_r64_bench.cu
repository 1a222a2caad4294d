// Microbenchmark + self-check of the two-pass radix-64 transform (fft4096_r64.cuh): 64 threads per transform,
// 4 independent groups per CTA of 256 threads, 1 CTA per SM.  Prints SM cycles per antenna chain
// (IFFT -> soft limiter -> FFT) for comparison with fft_phase_bench.cu (three-pass radix-16 form).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o fft_r64_bench fft_r64_bench.cu
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <complex>
#include <vector>

#include "../../m_mimo_ofdm_with_nonlinear_pa_sim_b200/csrc/frame_kernel.cuh"
#include "../../m_mimo_ofdm_with_nonlinear_pa_sim_b200/csrc/fft4096_r64.cuh"
using namespace mmimo;

__device__ __forceinline__ pf2 pconj(pf2 a) { return pmul(a, pk(1.f, -1.f)); }
struct SyncNone64 {
    static constexpr bool kExchange = false;
    __device__ __forceinline__ void x_sync() const {}
};
struct SyncNoBar64 {
    static constexpr bool kExchange = true;
    __device__ __forceinline__ void x_sync() const {}
};

// one transform by group 0 of one CTA: in/out natural bin order
template <int S>
__global__ void __launch_bounds__(256, 1) k_check(const float2* in, float2* out, const float2* twg) {
    extern __shared__ float2 smem[];
    float2* tw = smem;
    pf2* buf = reinterpret_cast<pf2*>(smem + 4096) + (threadIdx.x >> 6) * 4096;
    for (int i = threadIdx.x; i < 4096; i += 256) tw[i] = twg[i];
    __syncthreads();
    const int lane = threadIdx.x & 63, g = threadIdx.x >> 6;
    pf2 x[64];
#pragma unroll
    for (int n1 = 0; n1 < 64; ++n1) x[n1] = pk2(in[lane + 64 * n1]);
    fft4096_r64<S, false, false>(x, buf, TwSmem64{tw}, lane, SyncGroup{g + 1});
    if (g == 0) {
#pragma unroll
        for (int k1 = 0; k1 < 64; ++k1) out[lane + 64 * k1] = unpk(x[rev64(k1)]);
    }
}

// pruned chain check: data bins only in, data bins only out
__global__ void __launch_bounds__(256, 1) k_check_chain(const float2* in, float2* out, const float2* twg) {
    extern __shared__ float2 smem[];
    float2* tw = smem;
    pf2* buf = reinterpret_cast<pf2*>(smem + 4096) + (threadIdx.x >> 6) * 4096;
    for (int i = threadIdx.x; i < 4096; i += 256) tw[i] = twg[i];
    __syncthreads();
    const int lane = threadIdx.x & 63, g = threadIdx.x >> 6;
    pf2 x[64];
#pragma unroll
    for (int n1 = 0; n1 < 64; ++n1) x[n1] = pk2(in[lane + 64 * n1]);
    fft4096_r64<+1, true, false>(x, buf, TwSmem64{tw}, lane, SyncGroup{g + 1});
    pf2 y[64];
#pragma unroll
    for (int n1 = 0; n1 < 64; ++n1) y[n1] = pa_apply_pk<PA_SOFT_LIMITER>(x[rev64(n1)], 1.0f, 1.0f, 0.f);
    fft4096_r64<-1, false, true>(y, buf, TwSmem64{tw}, lane, SyncGroup{g + 1});
    if (g == 0) {
#pragma unroll
        for (int k1 = 0; k1 < 64; ++k1)
            if (k1 <= 16 || k1 >= 48) out[lane + 64 * k1] = unpk(y[rev64(k1)]);
    }
}

template <int MODE, bool TWR = false, int NT = 256>
__global__ void __launch_bounds__(NT, 1) k(float2* out, const float2* twg, int iters, int delay, const float2* tw4096g) {
    extern __shared__ float2 smem[];
    float2* tw = smem;
    pf2* buf = reinterpret_cast<pf2*>(smem + 4096) + (threadIdx.x >> 6) * 4096;
    for (int i = threadIdx.x; i < 4096; i += NT) tw[i] = twg[i];
    __syncthreads();
    const int lane = threadIdx.x & 63, g = threadIdx.x >> 6;
    TwReg64 twr;
    if (TWR) load_tw64(twr, tw4096g, lane);
    const TwSmem64 tws{tw};
    pf2 x[64];
#pragma unroll
    for (int j = 0; j < 64; ++j) x[j] = pk((float)(lane + j) * 1e-3f, (float)(lane - j) * 1e-3f);
    if (g >= 2 && delay > 0) {  // stagger the groups that share SM sub-partitions with groups 0 / 1
        const long long t0 = clock64();
        while (clock64() - t0 < delay) {}
    }
    for (int it = 0; it < iters; ++it) {
        pf2 y[64];
        if (MODE == 0) {
            if (TWR) fft4096_r64<+1, true, false>(x, buf, twr, lane, SyncGroup{g + 1}); else fft4096_r64<+1, true, false>(x, buf, tws, lane, SyncGroup{g + 1});
#pragma unroll
            for (int n1 = 0; n1 < 64; ++n1) y[n1] = pa_apply_pk<PA_SOFT_LIMITER>(x[rev64(n1)], 1.0f, 1.0f, 0.f);
            if (TWR) fft4096_r64<-1, false, true>(y, buf, twr, lane, SyncGroup{g + 1}); else fft4096_r64<-1, false, true>(y, buf, tws, lane, SyncGroup{g + 1});
        } else if (MODE == 3 || MODE == 4) {
            constexpr int D = MODE == 3 ? 8 : 16;
            fft4096_r64_pipelined<+1, true, false, D>(x, buf, tws, lane, SyncGroup{g + 1});
#pragma unroll
            for (int n1 = 0; n1 < 64; ++n1) y[n1] = pa_apply_pk<PA_SOFT_LIMITER>(x[rev64(n1)], 1.0f, 1.0f, 0.f);
            fft4096_r64_pipelined<-1, false, true, D>(y, buf, tws, lane, SyncGroup{g + 1});
        } else if (MODE == 5 || MODE == 6) {
            // ONE transform body for both directions: IFFT(x) = conj(FFT(conj x)), the limiter commutes with conj,
            // no pruning; the digit-reversal between the transforms is folded into the limiter's destination registers
            constexpr int D = MODE == 5 ? 8 : 4;
#pragma unroll 1
            for (int half = 0; half < 2; ++half) {
                fft4096_r64_pipelined<-1, false, false, D>(x, buf, tws, lane, SyncGroup{g + 1});
                if (half == 0) {
#pragma unroll
                    for (int n1 = 0; n1 < 64; ++n1) {
                        if (n1 <= rev64(n1)) {
                            const pf2 a = pconj(pa_apply_pk<PA_SOFT_LIMITER>(x[rev64(n1)], 1.0f, 1.0f, 0.f));
                            const pf2 b = pconj(pa_apply_pk<PA_SOFT_LIMITER>(x[n1], 1.0f, 1.0f, 0.f));
                            x[n1] = a;
                            x[rev64(n1)] = b;
                        }
                    }
                }
            }
#pragma unroll
            for (int n1 = 0; n1 < 64; ++n1) y[n1] = x[n1];
        } else if (MODE == 1) {
            if (TWR) fft4096_r64<+1, true, false>(x, buf, twr, lane, SyncNone64()); else fft4096_r64<+1, true, false>(x, buf, tws, lane, SyncNone64());
#pragma unroll
            for (int n1 = 0; n1 < 64; ++n1) y[n1] = pa_apply_pk<PA_SOFT_LIMITER>(x[rev64(n1)], 1.0f, 1.0f, 0.f);
            if (TWR) fft4096_r64<-1, false, true>(y, buf, twr, lane, SyncNone64()); else fft4096_r64<-1, false, true>(y, buf, tws, lane, SyncNone64());
        } else {
            if (TWR) fft4096_r64<+1, true, false>(x, buf, twr, lane, SyncNoBar64()); else fft4096_r64<+1, true, false>(x, buf, tws, lane, SyncNoBar64());
#pragma unroll
            for (int n1 = 0; n1 < 64; ++n1) y[n1] = pa_apply_pk<PA_SOFT_LIMITER>(x[rev64(n1)], 1.0f, 1.0f, 0.f);
            if (TWR) fft4096_r64<-1, false, true>(y, buf, twr, lane, SyncNoBar64()); else fft4096_r64<-1, false, true>(y, buf, tws, lane, SyncNoBar64());
        }
#pragma unroll
        for (int n1 = 0; n1 < 64; ++n1)
            if (n1 <= 16 || n1 >= 48) x[n1] = pmul(y[rev64(n1)], pbc(1e-3f));
        if (lane != 0) x[16] = pk(0.f, 0.f);
    }
    float2 s = make_float2(0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 64; ++j) { const float2 v = unpk(x[j]); s.x += v.x; s.y += v.y; }
    if (s.x == 123.456f) out[0] = s;
}

typedef std::complex<double> cd;
static std::vector<cd> dft(const std::vector<cd>& x, int sign) {
    const int n = (int)x.size();
    std::vector<cd> w(n), o(n);
    for (int i = 0; i < n; ++i) w[i] = std::polar(1.0, sign * 2 * M_PI * i / n);
    for (int k = 0; k < n; ++k) {
        cd a = 0;
        for (int i = 0; i < n; ++i) a += x[i] * w[(size_t)i * k % n];
        o[k] = a;
    }
    return o;
}

static const float2* g_tw4096 = nullptr;
template <int MODE, bool TWR = false, int NT = 256>
static void run(const char* name, float2* d, const float2* tw, int nsm, int delay = 0) {
    const int iters = 1000, smem = (4096 + (NT / 64) * 4096) * 8;
    cudaFuncSetAttribute(k<MODE, TWR, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        k<MODE, TWR, NT><<<nsm, NT, smem>>>(d, tw, iters, delay, g_tw4096);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep) best = fminf(best, ms);
    }
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double cyc = best * 1e-3 * clk * 1e3;
    printf("%-34s dly %5d %8.3f ms  %8.0f SM-cycles/chain (4 chains per SM in flight)\n", name, delay, best, cyc / iters / (NT / 64));
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(e));
}

int main(int argc, char** argv) {
    const bool prof = argc > 1;
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs, clockRate %d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
    std::vector<float2> tw(4096);
    for (int k0 = 0; k0 < 64; ++k0) for (int n0 = 0; n0 < 64; ++n0)
        tw[k0 * 64 + n0] = make_float2((float)cos(-2 * M_PI * k0 * n0 / 4096), (float)sin(-2 * M_PI * k0 * n0 / 4096));
    float2 *d, *dtw, *din, *dout;
    cudaMalloc(&d, 64); cudaMalloc(&dtw, 4096 * 8); cudaMalloc(&din, 4096 * 8); cudaMalloc(&dout, 4096 * 8);
    cudaMemcpy(dtw, tw.data(), 4096 * 8, cudaMemcpyHostToDevice);
    {
        std::vector<float2> t4(4096);
        for (int i = 0; i < 4096; ++i) t4[i] = make_float2((float)cos(-2 * M_PI * i / 4096), (float)sin(-2 * M_PI * i / 4096));
        float2* d4; cudaMalloc(&d4, 4096 * 8); cudaMemcpy(d4, t4.data(), 4096 * 8, cudaMemcpyHostToDevice);
        g_tw4096 = d4;
    }
    const int smem = (4096 + 4 * 4096) * 8;
    cudaFuncSetAttribute(k_check<+1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(k_check<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(k_check_chain, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (prof) {
        const int smem2 = (4096 + 4 * 4096) * 8;
        cudaFuncSetAttribute(k<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2);
        k<0, true><<<p.multiProcessorCount, 256, smem2>>>(d, dtw, 100, 0, g_tw4096);
        cudaDeviceSynchronize();
        printf("prof run: %s\n", cudaGetErrorString(cudaGetLastError()));
        return 0;
    }
    // ---- correctness -------------------------------------------------------------------------------
    std::vector<float2> in(4096), out(4096);
    std::vector<cd> xin(4096);
    srand(1);
    for (int i = 0; i < 4096; ++i) {
        in[i] = make_float2(rand() / (float)RAND_MAX - 0.5f, rand() / (float)RAND_MAX - 0.5f);
        xin[i] = cd(in[i].x, in[i].y);
    }
    cudaMemcpy(din, in.data(), 4096 * 8, cudaMemcpyHostToDevice);
    for (int sgn = -1; sgn <= 1; sgn += 2) {
        if (sgn < 0) k_check<-1><<<1, 256, smem>>>(din, dout, dtw); else k_check<+1><<<1, 256, smem>>>(din, dout, dtw);
        cudaMemcpy(out.data(), dout, 4096 * 8, cudaMemcpyDeviceToHost);
        std::vector<cd> ref = dft(xin, sgn);
        double num = 0, den = 0;
        for (int i = 0; i < 4096; ++i) { num += std::norm(cd(out[i].x, out[i].y) - ref[i]); den += std::norm(ref[i]); }
        printf("transform sign %+d: rel L2 error %.3e  (%s)\n", sgn, sqrt(num / den), cudaGetErrorString(cudaGetLastError()));
    }
    {   // pruned chain: data bins 1..1024, 3072..4095
        std::vector<cd> xd(4096, cd(0, 0));
        for (int i = 0; i < 4096; ++i) {
            const bool data = (i >= 1 && i <= 1024) || i >= 3072;
            if (!data) in[i] = make_float2(0.f, 0.f);
            in[i].x *= 0.05f; in[i].y *= 0.05f;
            xd[i] = cd(in[i].x, in[i].y);
        }
        cudaMemcpy(din, in.data(), 4096 * 8, cudaMemcpyHostToDevice);
        cudaMemset(dout, 0, 4096 * 8);
        k_check_chain<<<1, 256, smem>>>(din, dout, dtw);
        cudaMemcpy(out.data(), dout, 4096 * 8, cudaMemcpyDeviceToHost);
        std::vector<cd> td = dft(xd, +1);
        int nclip = 0;
        for (auto& v : td) { const double pw = std::norm(v); if (pw > 1.0) { v *= 1.0 / sqrt(pw); ++nclip; } }
        std::vector<cd> ref = dft(td, -1);
        double num = 0, den = 0;
        for (int i = 0; i < 4096; ++i) {
            const bool data = (i >= 1 && i <= 1024) || i >= 3072;
            if (!data) continue;
            num += std::norm(cd(out[i].x, out[i].y) - ref[i]); den += std::norm(ref[i]);
        }
        printf("pruned chain (IFFT, limiter [%d clipped], FFT): rel L2 error on data bins %.3e  (%s)\n", nclip, sqrt(num / den),
               cudaGetErrorString(cudaGetLastError()));
    }
    // ---- timing ------------------------------------------------------------------------------------
    const int nsm = p.multiProcessorCount;
    run<0>("r64 full chain (IFFT, PA, FFT)", d, dtw, nsm);
    run<1>("r64 butterflies + twiddle loads only", d, dtw, nsm);
    run<2>("r64 full chain, no barriers", d, dtw, nsm);
    run<3>("r64 chain, pipelined loads (8)", d, dtw, nsm);
    run<4>("r64 chain, pipelined loads (16)", d, dtw, nsm);
    run<3>("r64 chain, pipelined (8), groups 2,3 late", d, dtw, nsm, 3500);
    run<3>("r64 chain, pipelined (8), groups 2,3 late", d, dtw, nsm, 7000);
    run<4>("r64 chain, pipelined (16), groups 2,3 late", d, dtw, nsm, 7000);
    run<5>("r64 one shared transform body (8)", d, dtw, nsm);
    run<6>("r64 one shared transform body (4)", d, dtw, nsm);
    run<5>("r64 shared body (8), groups 2,3 late", d, dtw, nsm, 3500);
    run<0, true>("r64 full chain, register twiddles", d, dtw, nsm);
    run<0, false, 384>("r64 full chain, 6 groups/SM", d, dtw, nsm);
    run<0, false, 320>("r64 full chain, 5 groups/SM", d, dtw, nsm);
    return 0;
}
