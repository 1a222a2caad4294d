// Signal-to-distortion evaluator: the per-snapshot powers that
// main_beampatterns_plotting/sdr_mp.py:124-151 computes from AntennaArray.transmit(return_both=True) and
// Miso*Fd.propagate(sum_signals=False):
//     desired    = sum_s | sum_k alpha_k h_k[s] X_clean_k[s] |^2
//     distortion = sum_s | sum_k h_k[s] (X_k[s] - alpha_k X_clean_k[s]) |^2
// fused like the frame kernel: one CTA per snapshot, antennas looped, both antenna sums in registers, the [K, n_fft]
// signal matrices of the reference are never materialised.
#pragma once
#include "frame_kernel.cuh"

namespace mmimo {

struct SdrArgs {
    TxSide tx;                       // channel, precoder source, symbol scale (as in FrameArgs::tx)
    const FrameScalars* scal;        // [F]
    const float* alpha;              // [F][K] per-antenna Bussgang alpha (k_prep_realisation)
    float pa_p1;
    Qam qam;
    int n_frames;
    const uint8_t* tx_idx;           // [F][2048] native slots or nullptr (Philox)
    uint64_t seed, frame_id0;
    double* desired;                 // [F]
    double* distortion;              // [F]
};

template <int PA>
__global__ void __launch_bounds__(FFT_THREADS, 2) k_sdr_frame(const SdrArgs a, const float2* __restrict__ tw4096,
                                                              const float2* __restrict__ tw256) {
    extern __shared__ float2 smem[];
    pf2* bufA = reinterpret_cast<pf2*>(smem);
    pf2* bufB = reinterpret_cast<pf2*>(smem + 4096);
    float2* tw2 = smem + 8192;
    float2* symn_s = tw2 + 256;
    double* red_s = reinterpret_cast<double*>(symn_s + 2048);
    const int t = threadIdx.x;
    tw2[t] = tw256[t];
    Tw1 tw1;
    load_tw1(tw1, tw4096, t);
    __syncthreads();

    for (int f = blockIdx.x; f < a.n_frames; f += gridDim.x) {
        const FrameScalars sc = a.scal[f];
        uint32_t txw[2];
        if (a.tx_idx) {
            const uint8_t* p = a.tx_idx + (size_t)f * N_SC;
            txw[0] = txw[1] = 0;
#pragma unroll
            for (int jp = 0; jp < 8; ++jp) txw[jp >> 2] |= (uint32_t)p[jp * 256 + t] << (8 * (jp & 3));
        } else {
            uint32_t r[4];
            const uint64_t fid = a.frame_id0 + f;
            philox4x32_10((uint32_t)fid, (uint32_t)(fid >> 32), RNG_BITS, (uint32_t)t, (uint32_t)a.seed, (uint32_t)(a.seed >> 32), r);
            const uint32_t m = (1u << (2 * a.qam.log2side)) - 1u;
            const uint32_t m4 = m | (m << 8) | (m << 16) | (m << 24);
            txw[0] = r[0] & m4;
            txw[1] = r[1] & m4;
        }
        const TxSide& sd = a.tx;
        const float2* __restrict__ Hm = sd.h_mac + f * sd.h_mac_fs;
        const float2* __restrict__ Hp = sd.v ? sd.v + f * sd.v_fs : sd.h_pre + f * sd.h_pre_fs;
        const bool explicit_v = sd.v != nullptr;
        const float p0 = sc.pa_p0_tx, sqrt_p0 = sqrtf(p0);
        {
            const float* __restrict__ ss = sd.sscale + f * sd.sscale_fs;
#pragma unroll
            for (int jp = 0; jp < 8; ++jp) {
                const int p = jp * 256 + t;
                const float2 s = qam_symbol(a.qam, (txw[jp >> 2] >> (8 * (jp & 3))) & 0xffu);
                const float g = ss[p];
                symn_s[p] = make_float2(s.x * g, s.y * g);
            }
        }
        pf2 des[8], dis[8];
#pragma unroll
        for (int jp = 0; jp < 8; ++jp) des[jp] = dis[jp] = pk(0.f, 0.f);
        for (int k = 0; k < sd.k; ++k) {
            const float2* __restrict__ pr = Hp + (size_t)k * N_SC;
            const float2* __restrict__ hk = Hm + (size_t)k * N_SC;
            const float al64 = a.alpha[(size_t)f * sd.k + k] * 64.f;  // x carries 1/64 (FFT scaling), the clean signal does not
            pf2 x[16];
#pragma unroll
            for (int jp = 0; jp < 8; ++jp) {
                const float2 s = symn_s[jp * 256 + t];
                const pf2 r = pk2(__ldg(pr + jp * 256 + t));
                const pf2 xv = explicit_v ? pcmul(r, s.x, s.y) : pfma(pswap(r), pbc(s.y), pmul(r, pk(s.x, -s.x)));
                x[jp_to_j(jp)] = xv;
                // alpha_k * h_k * X_clean_k joins the desired sum and leaves the distortion sum
                const pf2 c = pcfma(pmul(xv, pbc(al64)), pk2(__ldg(hk + jp * 256 + t)), pk(0.f, 0.f));
                des[jp] = padd(des[jp], c);
                dis[jp] = psub(dis[jp], c);
            }
#pragma unroll
            for (int j = 4; j < 12; ++j) x[j] = pk(0.f, 0.f);
            if (t == 0) { x[4] = x[0]; x[0] = pk(0.f, 0.f); }
            antenna_chain<PA>(x, bufA, bufB, tw2, tw1, t, p0, sqrt_p0, a.pa_p1);
            if (t == 0) x[0] = x[4];
#pragma unroll
            for (int jp = 0; jp < 8; ++jp)
                dis[jp] = pcfma(pmul(x[jp_to_j(jp)], pbc(0.015625f)), pk2(__ldg(hk + jp * 256 + t)), dis[jp]);
        }
        double pd = 0.0, pi = 0.0;
#pragma unroll
        for (int jp = 0; jp < 8; ++jp) {
            const float2 d = unpk(des[jp]), e = unpk(dis[jp]);
            pd += (double)d.x * d.x + (double)d.y * d.y;
            pi += (double)e.x * e.x + (double)e.y * e.y;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            pd += __shfl_xor_sync(0xffffffffu, pd, o);
            pi += __shfl_xor_sync(0xffffffffu, pi, o);
        }
        __syncthreads();
        if ((t & 31) == 0) { red_s[t >> 5] = pd; red_s[8 + (t >> 5)] = pi; }
        __syncthreads();
        if (t == 0) {
            double sd2 = 0.0, si2 = 0.0;
            for (int w = 0; w < 8; ++w) { sd2 += red_s[w]; si2 += red_s[8 + w]; }
            a.desired[f] = sd2;
            a.distortion[f] = si2;
        }
    }
}
constexpr int SDR_SMEM_BYTES = (4096 * 2 + 256 + 2048) * 8 + 16 * 8;

}  // namespace mmimo
