// Per-realisation kernels: channel generation, MRT precoder + AGC / alpha / noise bookkeeping,
// and the FFT-free "no distortion" frame.  All per-sub-carrier arrays use the native slot order.
#pragma once
#include "sdr_kernel.cuh"

namespace mmimo {

constexpr double C_LIGHT = 299792458.0;  // scipy.constants.c (channel.py:61,66)
constexpr int MAX_ANT = 256;

// uniform double in (0,1) from two 32-bit words
__device__ __forceinline__ double u01(uint32_t hi, uint32_t lo) {
    return ((double)hi * 4294967296.0 + (double)lo + 0.5) * (1.0 / 18446744073709551616.0);
}

// ---------------------------------------------------------------------------------------------
// k_gen_channel: MisoLosFd / MisoTwoPathFd.calc_channel_mat (channel.py:35-72, :116-167) on the data
// sub-carriers only; receiver position injected or drawn as in mp_model.py:143-148 (both offsets are
// added to rx_loc_x - the reference's cord_y quirk).  grid = (K, F), block = 256.
// ---------------------------------------------------------------------------------------------
struct ChanArgs {
    const double* tx_pos;     // [K][3]
    const double* rx_pos_in;  // [F][3] or nullptr -> Philox
    double* rx_pos_out;       // [F][3] or nullptr
    const double* freq_slot;  // [2048] sub-carrier frequency per native slot (float32-rounded values, channel.py:51-52)
    double rx0[3];            // nominal receiver position
    double loc_var;           // rx_loc_var
    double gain_lin;          // sqrt(10^((G_tx+G_rx)/10))
    int kind;                 // 0 LOS, 1 two-path
    int skip_attenuation;
    uint64_t seed, frame_id0;
    uint32_t pos_stream;      // RNG_POS or a clean-run stream
    float2* h_out;            // [F][K][2048]
    int n_ant;
    // multi-user scripts (main_multiuser_mcnc_ber_vs_ebn0.py:546-556): frame f belongs to user f % n_usr whose nominal
    // position is rx0_tab[user]; here x and y are offset around their OWN nominal values (no cord_y quirk)
    const double* rx0_tab;    // [n_usr][3] or nullptr
    int n_usr;
    // fused first pass of k_prep_realisation (MRT, precoder derived from this very channel): with grid.x == 1 the CTA
    // sees all K rows of the realisation and also writes the symbol scale 1 / (64 ||h[:, s]||) (antenna_array.py:170-171)
    float* sscale_out;        // [F][2048] or nullptr
};

// one propagation path: amp / f * exp(+j 2 pi d f / c), amp = gain c / (4 pi d) (channel.py:56-66); the phase (thousands
// of cycles) is range-reduced in double, the amplitude needs fp32 only.  d_over_c and amp are per (frame, antenna), rf is
// the fp32 reciprocal of the sub-carrier frequency (per slot, loop-invariant in the callers).
#ifndef CHAN_FAST
#define CHAN_FAST 1
#endif
#if CHAN_FAST
// x = d f / c in turns.  4x = n + u with n integer and |u| <= 1/2 comes out of three double operations (round-to-integer
// by the 1.5 * 2^52 constant, the quadrant n mod 4 sits in the low word of the sum, the second FMA returns the EXACT
// remainder of the product), then cos / sin of (pi/2) u are two degree-3 polynomials in u^2 (max error 1.2e-7) and the
// quadrant turns them by j^n.  24 instructions per coefficient instead of ~40 through sincospif and a division.
__device__ __forceinline__ float2 path_coeff(double d_over_c, double f, float rf, float amp, bool skip_att) {
    const double t4 = 4.0 * d_over_c;
    const double big = fma(t4, f, 6755399441055744.0);
    const int q = __double2loint(big);
    const float u = (float)fma(t4, f, -(big - 6755399441055744.0));
    const float2 cs = cis_quadrant(q, u);
    const float a = skip_att ? 1.0f : amp * rf;
    const float re = cs.x, im = cs.y;
    return make_float2(a * re, a * im);
}
#else
__device__ __forceinline__ float2 path_coeff(double d_over_c, double f, float rf, float amp, bool skip_att) {
    const double x = d_over_c * f;
    const float r = (float)(x - rint(x));
    float s, c;
    sincospif(2.0f * r, &s, &c);
    if (skip_att) return make_float2(c, s);
    const float a = amp / (float)f;
    return make_float2(a * c, a * s);
}
#endif
__device__ __forceinline__ float2 path_coeff(double d_over_c, double f, float amp, bool skip_att) {
    return path_coeff(d_over_c, f, 1.0f / (float)f, amp, skip_att);
}

// the same from distance and linear gain (stage kernels: one realisation per call)
__device__ __forceinline__ float2 path_coeff(double d, double f, double gain, bool skip_att) {
    return path_coeff(d * (1.0 / C_LIGHT), f, (float)(gain * C_LIGHT / (4.0 * 3.14159265358979323846 * d)), skip_att);
}

constexpr int CHAN_ANT_PER_CTA = 8;  // antennas per CTA: the receiver draw and the geometry set-up are shared

// grid (ceil(K / CHAN_ANT_PER_CTA), F): one CTA writes the rows of up to 8 antennas of one realisation;
// grid (1, F) with sscale_out: one CTA writes all K rows and the MRT symbol scale
__global__ void __launch_bounds__(256) k_gen_channel(const ChanArgs a) {
    __shared__ double t1_s[MAX_ANT], t2_s[MAX_ANT];
    __shared__ float amp1_s[MAX_ANT], amp2_s[MAX_ANT];
    const int per_cta = a.sscale_out ? a.n_ant : CHAN_ANT_PER_CTA;
    const int kb = blockIdx.x * per_cta, f = blockIdx.y, t = threadIdx.x;
    if (t < per_cta && kb + t < a.n_ant) {  // geometry of antenna kb + t (channel.py:56-66, :118-140)
        const int k = kb + t;
        double rx[3];
        if (a.rx_pos_in) {
            rx[0] = a.rx_pos_in[f * 3 + 0]; rx[1] = a.rx_pos_in[f * 3 + 1]; rx[2] = a.rx_pos_in[f * 3 + 2];
        } else {
            uint32_t r[4];
            const uint64_t fid = a.frame_id0 + f;
            philox4x32_10((uint32_t)fid, (uint32_t)(fid >> 32), a.pos_stream, 0u, (uint32_t)a.seed, (uint32_t)(a.seed >> 32), r);
            if (a.rx0_tab) {
                const double* n0 = a.rx0_tab + (size_t)((a.frame_id0 + f) % (uint64_t)a.n_usr) * 3;
                rx[0] = n0[0] + (u01(r[0], r[1]) - 0.5) * a.loc_var;
                rx[1] = n0[1] + (u01(r[2], r[3]) - 0.5) * a.loc_var;
                rx[2] = n0[2];
            } else {
                rx[0] = a.rx0[0] + (u01(r[0], r[1]) - 0.5) * a.loc_var;
                rx[1] = a.rx0[0] + (u01(r[2], r[3]) - 0.5) * a.loc_var;  // quirk: rx_loc_x, mp_model.py:146,196
                rx[2] = a.rx0[2];
            }
        }
        if (a.rx_pos_out && k == 0) { a.rx_pos_out[f * 3 + 0] = rx[0]; a.rx_pos_out[f * 3 + 1] = rx[1]; a.rx_pos_out[f * 3 + 2] = rx[2]; }
        const double dx = a.tx_pos[k * 3 + 0] - rx[0], dy = a.tx_pos[k * 3 + 1] - rx[1], dz = a.tx_pos[k * 3 + 2] - rx[2];
        const double d1 = sqrt(dx * dx + dy * dy + dz * dz);
        double d2 = 0.0;
        if (a.kind == 1) {
            const double dxy = sqrt(dx * dx + dy * dy);
            const double el = atan((a.tx_pos[k * 3 + 2] + rx[2]) / dxy);
            const double sn = sin(el);
            d2 = a.tx_pos[k * 3 + 2] / sn + rx[2] / sn;
        }
        const double amp0 = a.gain_lin * C_LIGHT / (4.0 * 3.14159265358979323846);
        t1_s[t] = d1 * (1.0 / C_LIGHT);
        t2_s[t] = d2 * (1.0 / C_LIGHT);
        amp1_s[t] = (float)(amp0 / d1);
        amp2_s[t] = a.kind == 1 ? (float)(amp0 / d2) : 0.f;
    }
    __syncthreads();
    // two adjacent slots per thread and step: 16-byte stores, 16-byte table loads (the table stays in registers)
    double2 fr[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) fr[j] = *reinterpret_cast<const double2*>(a.freq_slot + j * 512 + 2 * t);
    float2 rf[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) rf[j] = make_float2(1.0f / (float)fr[j].x, 1.0f / (float)fr[j].y);
    const int nk = min(per_cta, a.n_ant - kb);
    float nrm[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) nrm[j] = 0.f;
    for (int i = 0; i < nk; ++i) {
        float2* out = a.h_out + ((size_t)f * a.n_ant + kb + i) * N_SC;
        const double t1 = t1_s[i], t2 = t2_s[i];
        const float amp1 = amp1_s[i], amp2 = amp2_s[i];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int p = j * 512 + 2 * t;
            float2 h0 = path_coeff(t1, fr[j].x, rf[j].x, amp1, a.skip_attenuation);
            float2 h1 = path_coeff(t1, fr[j].y, rf[j].y, amp1, a.skip_attenuation);
            if (a.kind == 1) {
                const float2 g0 = path_coeff(t2, fr[j].x, rf[j].x, amp2, a.skip_attenuation);
                const float2 g1 = path_coeff(t2, fr[j].y, rf[j].y, amp2, a.skip_attenuation);
                h0.x -= g0.x; h0.y -= g0.y;
                h1.x -= g1.x; h1.y -= g1.y;
            }
            *reinterpret_cast<float4*>(out + p) = make_float4(h0.x, h0.y, h1.x, h1.y);
            // same operation order as the first pass of k_prep_realisation: bit-identical scale
            nrm[2 * j] = fmaf(h0.x, h0.x, fmaf(h0.y, h0.y, nrm[2 * j]));
            nrm[2 * j + 1] = fmaf(h1.x, h1.x, fmaf(h1.y, h1.y, nrm[2 * j + 1]));
        }
    }
    if (a.sscale_out) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int p = j * 512 + 2 * t;
            *reinterpret_cast<float2*>(a.sscale_out + (size_t)f * N_SC + p) =
                make_float2((1.0f / sqrtf(nrm[2 * j])) * 0.015625f, (1.0f / sqrtf(nrm[2 * j + 1])) * 0.015625f);
        }
    }
}

// k_gen_channel_bg: the same rows as k_gen_channel's (K/8, F) form, written by 64-thread CTAs that need 64 registers per
// thread and 200 bytes of shared memory - small enough to be resident NEXT TO a CTA of the frame kernel (240 registers x 256
// threads, 221.5 KB), so that the channel rows of batch n + 1 are generated in the issue slots the frame kernel of batch n
// leaves idle (link_host.inl, "pipelined realisations").  Same arithmetic, same operation order per element: bit-identical rows.
constexpr int CHAN_BG_THREADS = 64;
// no shared memory and no barrier: every thread derives the receiver position and the geometry of the CTA's antennas itself
__global__ void __launch_bounds__(CHAN_BG_THREADS, 16) k_gen_channel_bg(const ChanArgs a) {
    const int kb = blockIdx.x * CHAN_ANT_PER_CTA, f = blockIdx.y, t = threadIdx.x;
    double rx[3];
    if (a.rx_pos_in) {
        rx[0] = a.rx_pos_in[f * 3 + 0]; rx[1] = a.rx_pos_in[f * 3 + 1]; rx[2] = a.rx_pos_in[f * 3 + 2];
    } else {
        uint32_t r[4];
        const uint64_t fid = a.frame_id0 + f;
        philox4x32_10((uint32_t)fid, (uint32_t)(fid >> 32), a.pos_stream, 0u, (uint32_t)a.seed, (uint32_t)(a.seed >> 32), r);
        if (a.rx0_tab) {
            const double* n0 = a.rx0_tab + (size_t)((a.frame_id0 + f) % (uint64_t)a.n_usr) * 3;
            rx[0] = n0[0] + (u01(r[0], r[1]) - 0.5) * a.loc_var;
            rx[1] = n0[1] + (u01(r[2], r[3]) - 0.5) * a.loc_var;
            rx[2] = n0[2];
        } else {
            rx[0] = a.rx0[0] + (u01(r[0], r[1]) - 0.5) * a.loc_var;
            rx[1] = a.rx0[0] + (u01(r[2], r[3]) - 0.5) * a.loc_var;  // quirk: rx_loc_x, mp_model.py:146,196
            rx[2] = a.rx0[2];
        }
    }
    if (a.rx_pos_out && kb == 0 && t == 0) { a.rx_pos_out[f * 3 + 0] = rx[0]; a.rx_pos_out[f * 3 + 1] = rx[1]; a.rx_pos_out[f * 3 + 2] = rx[2]; }
    const int nk = min(CHAN_ANT_PER_CTA, a.n_ant - kb);
    for (int i = 0; i < nk; ++i) {
        const int k = kb + i;  // geometry of antenna k, exactly as in k_gen_channel (channel.py:56-66, :118-140)
        const double dx = a.tx_pos[k * 3 + 0] - rx[0], dy = a.tx_pos[k * 3 + 1] - rx[1], dz = a.tx_pos[k * 3 + 2] - rx[2];
        const double d1 = sqrt(dx * dx + dy * dy + dz * dz);
        double d2 = 0.0;
        if (a.kind == 1) {
            const double dxy = sqrt(dx * dx + dy * dy);
            const double el = atan((a.tx_pos[k * 3 + 2] + rx[2]) / dxy);
            const double sn = sin(el);
            d2 = a.tx_pos[k * 3 + 2] / sn + rx[2] / sn;
        }
        const double amp0 = a.gain_lin * C_LIGHT / (4.0 * 3.14159265358979323846);
        const double t1 = d1 * (1.0 / C_LIGHT), t2 = d2 * (1.0 / C_LIGHT);
        const float amp1 = (float)(amp0 / d1), amp2 = a.kind == 1 ? (float)(amp0 / d2) : 0.f;
        float2* out = a.h_out + ((size_t)f * a.n_ant + k) * N_SC;
#pragma unroll 2
        for (int p = 2 * t; p < N_SC; p += 2 * CHAN_BG_THREADS) {
            const double2 fr = *reinterpret_cast<const double2*>(a.freq_slot + p);
            float2 h0 = path_coeff(t1, fr.x, amp1, a.skip_attenuation);
            float2 h1 = path_coeff(t1, fr.y, amp1, a.skip_attenuation);
            if (a.kind == 1) {
                const float2 g0 = path_coeff(t2, fr.x, amp2, a.skip_attenuation);
                const float2 g1 = path_coeff(t2, fr.y, amp2, a.skip_attenuation);
                h0.x -= g0.x; h0.y -= g0.y;
                h1.x -= g1.x; h1.y -= g1.y;
            }
            *reinterpret_cast<float4*>(out + p) = make_float4(h0.x, h0.y, h1.x, h1.y);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// k_prep_realisation: AntennaArray.set_precoding_matrix(MRT) (antenna_array.py:163-171),
// AntennaArray.update_distortion (:313-360) and Link.recalculate_agc (mp_model.py:290-329) for one
// channel realisation per CTA.
// ---------------------------------------------------------------------------------------------
struct PrepArgs {
    const float2* h_loop; long long h_fs;  // channel knowledge [F][K][2048]
    const float2* v_in;                    // optional externally supplied precoder [F][K][2048] (unscaled)
    float2* v_out;                         // [F][K][2048] precoder (unscaled) or nullptr
    float* sscale;                         // [F][2048] symbol scale of the frame kernel: 1/(64 ||h||) (MRT) or 1/64
    float2* agc_inv;                       // [F][2048]  1 / (64 * sum_k alpha_k h_k v_k)
    float2* hkvk_inv;                      // [F][2048]  1 / (sum_k h_k v_k)        (clean run, no FFT scale)
    FrameScalars* scal;                    // [F]
    float* alpha_out;                      // [F][K] or nullptr
    int n_ant, n_frames;
    int pa_kind, is_mcnc;
    double ibo_lin;            // 10^(ibo_db/10)  (TOI: 10^(toi_db/10))
    double avg_symbol_power;   // 42 for 64-QAM (modulation.py:218)
    double avg_sample_power;   // avg_symbol_power * n_sub_carr / n_fft (modulation.py:424)
    double snr_lin;            // 10^(snr_db/10)
};
// multi-point batches (mmimo_link_sweep): frame f of this launch belongs to sweep point (frame0 + f) / fpp, whose IBO / SNR
// replace the scalars of PrepArgs.  A separate kernel parameter on purpose: growing PrepArgs itself by these 24 bytes moves
// ptxas from 72 to 84 registers for the whole kernel (2 instead of 3 resident CTAs, 2.1 -> 4.8 ms per 4096 frames).
struct SweepArgs {
    const double* snr_lin_tab; const double* ibo_lin_tab;
    int fpp; unsigned frame0;   // 32-bit: a 64-bit division is an out-of-line call
    int sscale_ready;           // the channel kernel has written sscale already: skip the norm pass over H
};

__device__ __forceinline__ double block_sum_256(double v, double* red_s, int t) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((t & 31) == 0) red_s[t >> 5] = v;
    __syncthreads();
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red_s[w];
    return s;
}

template <int NW>
__device__ __forceinline__ double block_sum_n(double v, double* red_s, int t) {  // NW warps
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((t & 31) == 0) red_s[t >> 5] = v;
    __syncthreads();
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < NW; ++w) s += red_s[w];
    return s;
}

#ifndef PREP_T
#define PREP_T 512  // threads per realisation CTA; N_SC / PREP_T sub-carrier slots per thread
#endif
constexpr int PREP_NS = N_SC / PREP_T, PREP_W = PREP_T / 32;
template <bool SWEEP>
__global__ void __launch_bounds__(PREP_T) k_prep_realisation(const PrepArgs a, const SweepArgs sw) {
    __shared__ float part_s[MAX_ANT][PREP_W];
    __shared__ float alpha_s[MAX_ANT];
    __shared__ double vkpow_s[MAX_ANT];
    __shared__ double red_s[PREP_W];
    const int f = blockIdx.x, t = threadIdx.x, K = a.n_ant;
    const float2* __restrict__ H = a.h_loop + f * a.h_fs;
    const float2* __restrict__ Vin = a.v_in ? a.v_in + (size_t)f * K * N_SC : nullptr;
    float2* Vout = a.v_out ? a.v_out + (size_t)f * K * N_SC : nullptr;

    float ninv[PREP_NS];
    if (!Vin && sw.sscale_ready) {
#pragma unroll
        for (int jp = 0; jp < PREP_NS; ++jp) ninv[jp] = a.sscale[(size_t)f * N_SC + jp * PREP_T + t] * 64.0f;  // exact
    } else if (!Vin) {
        float nrm[PREP_NS];
#pragma unroll
        for (int jp = 0; jp < PREP_NS; ++jp) nrm[jp] = 0.f;
        for (int k = 0; k < K; ++k) {
#pragma unroll
            for (int jp = 0; jp < PREP_NS; ++jp) {
                const float2 h = H[(size_t)k * N_SC + jp * PREP_T + t];
                nrm[jp] = fmaf(h.x, h.x, fmaf(h.y, h.y, nrm[jp]));
            }
        }
#pragma unroll
        for (int jp = 0; jp < PREP_NS; ++jp) ninv[jp] = 1.0f / sqrtf(nrm[jp]);
    }
#pragma unroll
    for (int jp = 0; jp < PREP_NS; ++jp) a.sscale[(size_t)f * N_SC + jp * PREP_T + t] = Vin ? 0.015625f : ninv[jp] * 0.015625f;
    // precoder out + per-antenna power
    for (int k = 0; k < K; ++k) {
        float pw = 0.f;
#pragma unroll
        for (int jp = 0; jp < PREP_NS; ++jp) {
            const int p = jp * PREP_T + t;
            float2 v;
            if (Vin) v = Vin[(size_t)k * N_SC + p];
            else {
                const float2 h = H[(size_t)k * N_SC + p];
                v = make_float2(h.x * ninv[jp], -h.y * ninv[jp]);
            }
            pw = fmaf(v.x, v.x, fmaf(v.y, v.y, pw));
            if (Vout) Vout[(size_t)k * N_SC + p] = v;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) pw += __shfl_xor_sync(0xffffffffu, pw, o);
        if ((t & 31) == 0) part_s[k][t >> 5] = pw;
    }
    __syncthreads();
    if (t < K) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < PREP_W; ++w) s += (double)part_s[t][w];
        vkpow_s[t] = s;
        // mp_model.py:315-317 + modulation.py:178-189
        const double g2 = ((SWEEP && sw.ibo_lin_tab) ? sw.ibo_lin_tab[(sw.frame0 + (unsigned)f) / (unsigned)sw.fpp] : a.ibo_lin) * (double)N_SC / (s * (double)K);
        const double g = sqrt(g2);
        const double al = 1.0 - exp(-g2) + (1.7724538509055160273 * g * 0.5) * erfc(g);
        alpha_s[t] = (float)al;
        if (a.alpha_out) a.alpha_out[(size_t)f * K + t] = (float)al;
    }
    __syncthreads();
    // AGC vectors
    float2 agc[PREP_NS], hv[PREP_NS];
#pragma unroll
    for (int jp = 0; jp < PREP_NS; ++jp) agc[jp] = hv[jp] = make_float2(0.f, 0.f);
    for (int k = 0; k < K; ++k) {
        const float al = alpha_s[k];
#pragma unroll
        for (int jp = 0; jp < PREP_NS; ++jp) {
            const int p = jp * PREP_T + t;
            const float2 h = H[(size_t)k * N_SC + p];
            float2 v;
            if (Vin) v = Vin[(size_t)k * N_SC + p];
            else v = make_float2(h.x * ninv[jp], -h.y * ninv[jp]);
            const float2 m = cmul(h, v);
            hv[jp] = cadd(hv[jp], m);
            agc[jp] = make_float2(fmaf(al, m.x, agc[jp].x), fmaf(al, m.y, agc[jp].y));
        }
    }
    double s_agc = 0.0, s_hv = 0.0;
#pragma unroll
    for (int jp = 0; jp < PREP_NS; ++jp) {
        const int p = jp * PREP_T + t;
        const double a2 = (double)agc[jp].x * agc[jp].x + (double)agc[jp].y * agc[jp].y;
        const double h2 = (double)hv[jp].x * hv[jp].x + (double)hv[jp].y * hv[jp].y;
        s_agc += a2; s_hv += h2;
        // 1/z = conj(z)/|z|^2
        a.agc_inv[(size_t)f * N_SC + p] = make_float2((float)(agc[jp].x / (64.0 * a2)), (float)(-agc[jp].y / (64.0 * a2)));
        a.hkvk_inv[(size_t)f * N_SC + p] = make_float2((float)(hv[jp].x / h2), (float)(-hv[jp].y / h2));
    }
    s_agc = block_sum_n<PREP_W>(s_agc, red_s, t);
    s_hv = block_sum_n<PREP_W>(s_hv, red_s, t);
    if (t == 0) {
        double tot = 0.0;
        for (int k = 0; k < K; ++k) tot += vkpow_s[k];
        const double gain = tot / ((double)K * N_SC);           // antenna_array.py:328-341
        const double avg = a.avg_sample_power * gain;
        FrameScalars sc;
        // per-point IBO / SNR of a multi-point batch (mmimo_link_sweep)
        const double ibo_lin = (SWEEP && sw.ibo_lin_tab) ? sw.ibo_lin_tab[(sw.frame0 + (unsigned)f) / (unsigned)sw.fpp] : a.ibo_lin;
        const double snr_lin = (SWEEP && sw.snr_lin_tab) ? sw.snr_lin_tab[(sw.frame0 + (unsigned)f) / (unsigned)sw.fpp] : a.snr_lin;
        const double p0_tx = (a.pa_kind == PA_TOI) ? (1.0 / ibo_lin) / avg : ibo_lin * avg;  // distortion.py:37,228
        // CNC loop PA: avg_samp_pow = avg_symbol_power * n_sub_carr / n_fft (corrector.py:34-35)
        const double p0_cnc = (a.pa_kind == PA_TOI) ? (1.0 / ibo_lin) / a.avg_sample_power : ibo_lin * a.avg_sample_power;
        sc.pa_p0_tx = (float)p0_tx;
        sc.pa_p0_loop = (float)(a.is_mcnc ? p0_tx : p0_cnc);
        // noise.py:64-66: component std = 0.5*sqrt(2 P / snr), P = avg_symbol_power * noise_scaler
        sc.noise_sigma = (float)(0.5 * sqrt(2.0 * a.avg_symbol_power * (s_agc / N_SC) / snr_lin));
        sc.noise_sigma_clean = (float)(0.5 * sqrt(2.0 * a.avg_symbol_power * (s_hv / N_SC) / snr_lin));
        a.scal[f] = sc;
    }
}

// ---------------------------------------------------------------------------------------------
// k_prep_los: the work of k_prep_realisation in CLOSED FORM for a line-of-sight channel with MRT precoding from the
// true channel (the headline configuration: main_mp_miso_mcnc_ber_vs_ebn0.py).  With h_k[s] = a_k / f_s * e^(j phi)
// (channel.py:56-66, a_k = gain c / (4 pi d_k)) every quantity of antenna_array.py:163-171, :313-360 and
// mp_model.py:290-329 depends on the K amplitudes only:
//     ||h[:, s]||^2 = A / f_s^2,  A = sum_k a_k^2          v_k[s] = conj(h_k[s]) f_s / sqrt(A),  |v_k[s]|^2 = a_k^2 / A
//     vk_pow[k] = N_sc a_k^2 / A   ->  ibo_k, alpha_k      sum_k h_k v_k = sqrt(A) / f_s
//     sum_k alpha_k h_k v_k = S1 / (f_s sqrt(A)),  S1 = sum_k alpha_k a_k^2        mean |v|^2 = 1 / K
// so no pass over the [F][K][2048] channel rows is needed (k_prep_realisation reads them twice: 1.6 ms per 4144
// frames at K = 64).  One CTA per realisation; receiver positions come from the buffer k_gen_channel read or wrote.
// ---------------------------------------------------------------------------------------------
struct PrepLosArgs {
    const double* tx_pos;     // [K][3]
    const double* rx_pos;     // [F][3]
    const double* freq_slot;  // [2048]
    double gain_lin;
    float* sscale; float2* agc_inv; float2* hkvk_inv; FrameScalars* scal; float* alpha_out;
    int n_ant, pa_kind, is_mcnc;
    double ibo_lin, avg_symbol_power, avg_sample_power, snr_lin;
};

template <bool SWEEP>
__global__ void __launch_bounds__(256) k_prep_los(const PrepLosArgs a, const SweepArgs sw) {
    __shared__ double red_s[8];
    const int f = blockIdx.x, t = threadIdx.x, K = a.n_ant;
    const double ibo_lin = (SWEEP && sw.ibo_lin_tab) ? sw.ibo_lin_tab[(sw.frame0 + (unsigned)f) / (unsigned)sw.fpp] : a.ibo_lin;
    const double snr_lin = (SWEEP && sw.snr_lin_tab) ? sw.snr_lin_tab[(sw.frame0 + (unsigned)f) / (unsigned)sw.fpp] : a.snr_lin;
    double a2 = 0.0;
    if (t < K) {
        const double dx = a.tx_pos[t * 3 + 0] - a.rx_pos[f * 3 + 0], dy = a.tx_pos[t * 3 + 1] - a.rx_pos[f * 3 + 1],
                     dz = a.tx_pos[t * 3 + 2] - a.rx_pos[f * 3 + 2];
        // the amplitude the channel kernel uses (rounded to float there)
        const double amp = (double)(float)(a.gain_lin * C_LIGHT / (4.0 * 3.14159265358979323846) / sqrt(dx * dx + dy * dy + dz * dz));
        a2 = amp * amp;
    }
    const double A = block_sum_256(a2, red_s, t);
    double al_a2 = 0.0;
    if (t < K) {  // mp_model.py:315-317 + modulation.py:178-189 with vk_pow[k] = N_sc a_k^2 / A
        const double g2 = ibo_lin * A / (a2 * (double)K);
        const double g = sqrt(g2);
        const double al = 1.0 - exp(-g2) + (1.7724538509055160273 * g * 0.5) * erfc(g);
        if (a.alpha_out) a.alpha_out[(size_t)f * K + t] = (float)al;
        al_a2 = al * a2;
    }
    const double S1 = block_sum_256(al_a2, red_s, t);
    const double rA = sqrt(A);
    double s_f = 0.0;
#pragma unroll
    for (int jp = 0; jp < 8; ++jp) {
        const int p = jp * 256 + t;
        const double fr = a.freq_slot[p];
        s_f += 1.0 / (fr * fr);
        a.sscale[(size_t)f * N_SC + p] = (float)(fr / (64.0 * rA));
        a.agc_inv[(size_t)f * N_SC + p] = make_float2((float)(fr * rA / (64.0 * S1)), 0.f);
        a.hkvk_inv[(size_t)f * N_SC + p] = make_float2((float)(fr / rA), 0.f);
    }
    s_f = block_sum_256(s_f, red_s, t);
    if (t == 0) {
        const double avg = a.avg_sample_power / (double)K;   // antenna_array.py:328-341: mean |v|^2 = 1 / K
        FrameScalars sc;
        const double p0_tx = (a.pa_kind == PA_TOI) ? (1.0 / ibo_lin) / avg : ibo_lin * avg;
        const double p0_cnc = (a.pa_kind == PA_TOI) ? (1.0 / ibo_lin) / a.avg_sample_power : ibo_lin * a.avg_sample_power;
        sc.pa_p0_tx = (float)p0_tx;
        sc.pa_p0_loop = (float)(a.is_mcnc ? p0_tx : p0_cnc);
        const double s_agc = S1 * S1 / A * s_f, s_hv = A * s_f;
        sc.noise_sigma = (float)(0.5 * sqrt(2.0 * a.avg_symbol_power * (s_agc / N_SC) / snr_lin));
        sc.noise_sigma_clean = (float)(0.5 * sqrt(2.0 * a.avg_symbol_power * (s_hv / N_SC) / snr_lin));
        a.scal[f] = sc;
    }
}

// ---------------------------------------------------------------------------------------------
// k_clean_frame: the "no distortion" loop body (mp_model.py:158-173).  FFT(IFFT(x)) == x, so it is
// precode -> channel -> noise -> / sum_k h_k v_k -> slice -> count.  One CTA per frame.
// ---------------------------------------------------------------------------------------------
struct CleanArgs {
    const float2* h_true; long long h_fs;
    const float2* h_pre;                    // channel knowledge: v_k = conj(h_pre_k) * sscale (MRT, antenna_array.py:170-171)
    const float* sscale;                    // [F][2048] 1/(64 ||h_pre||)
    const float2* hkvk_inv;                 // [F][2048]
    const FrameScalars* scal;
    int n_ant, n_frames;
    Qam qam;
    const uint8_t* tx_idx; const float2* noise; int noise_unit;
    uint64_t seed, frame_id0;
    uint8_t* det_out;            // [F][2048] or nullptr
    uint32_t* err_out;           // [F] or nullptr
    unsigned long long* err_total;  // [1]
};

__global__ void __launch_bounds__(256) k_clean_frame(const CleanArgs a) {
    __shared__ unsigned int err_s;
    const int f = blockIdx.x, t = threadIdx.x;
    if (t == 0) err_s = 0;
    __syncthreads();
    const FrameScalars sc = a.scal[f];
    uint32_t txw[2] = {0u, 0u};
    if (a.tx_idx) {
#pragma unroll
        for (int jp = 0; jp < 8; ++jp) txw[jp >> 2] |= (uint32_t)a.tx_idx[(size_t)f * N_SC + jp * 256 + t] << (8 * (jp & 3));
    } else {
        uint32_t r[4];
        const uint64_t fid = a.frame_id0 + f;
        philox4x32_10((uint32_t)fid, (uint32_t)(fid >> 32), RNG_CLEAN_BITS, (uint32_t)t, (uint32_t)a.seed, (uint32_t)(a.seed >> 32), r);
        const uint32_t m = (1u << (2 * a.qam.log2side)) - 1u;
        const uint32_t m4 = m | (m << 8) | (m << 16) | (m << 24);
        txw[0] = r[0] & m4; txw[1] = r[1] & m4;
    }
    float2 g[8];
#pragma unroll
    for (int jp = 0; jp < 8; ++jp) g[jp] = make_float2(0.f, 0.f);
    const float2* __restrict__ H = a.h_true + f * a.h_fs;
    const float2* __restrict__ HP = a.h_pre + f * a.h_fs;
    for (int k = 0; k < a.n_ant; ++k) {
#pragma unroll
        for (int jp = 0; jp < 8; ++jp) g[jp] = cadd(g[jp], cmulc(H[(size_t)k * N_SC + jp * 256 + t], HP[(size_t)k * N_SC + jp * 256 + t]));
    }
#pragma unroll
    for (int jp = 0; jp < 8; ++jp) {
        const float ssc = a.sscale[(size_t)f * N_SC + jp * 256 + t];
        g[jp].x *= ssc; g[jp].y *= ssc;
    }
    float2 nz[8];
    if (a.noise) {
#pragma unroll
        for (int jp = 0; jp < 8; ++jp) {
            const float2 u = a.noise[(size_t)f * N_SC + jp * 256 + t];
            const float ns = a.noise_unit ? sc.noise_sigma_clean : 1.0f;
            nz[jp] = make_float2(u.x * ns, u.y * ns);
        }
    } else {
        const uint64_t fid = a.frame_id0 + f;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t r[4];
            philox4x32_10((uint32_t)fid, (uint32_t)(fid >> 32), RNG_CLEAN_NOISE, (uint32_t)(t * 4 + q), (uint32_t)a.seed, (uint32_t)(a.seed >> 32), r);
            const float2 g0 = box_muller(r[0], r[1]), g1 = box_muller(r[2], r[3]);
            nz[2 * q] = make_float2(g0.x * sc.noise_sigma_clean, g0.y * sc.noise_sigma_clean);
            nz[2 * q + 1] = make_float2(g1.x * sc.noise_sigma_clean, g1.y * sc.noise_sigma_clean);
        }
    }
    uint32_t detw[2] = {0u, 0u};
#pragma unroll
    for (int jp = 0; jp < 8; ++jp) {
        const int p = jp * 256 + t;
        const float2 s = qam_symbol(a.qam, (txw[jp >> 2] >> (8 * (jp & 3))) & 0xffu);
        // g holds sum h v / 64
        const float2 y = cadd(cmul(make_float2(64.f * g[jp].x, 64.f * g[jp].y), s), nz[jp]);
        const uint32_t idx = qam_detect(a.qam, cmul(y, a.hkvk_inv[(size_t)f * N_SC + p]));
        detw[jp >> 2] |= idx << (8 * (jp & 3));
        if (a.det_out) a.det_out[(size_t)f * N_SC + p] = (uint8_t)idx;
    }
    unsigned int nerr = __popc(detw[0] ^ txw[0]) + __popc(detw[1] ^ txw[1]);
    nerr = __reduce_add_sync(0xffffffffu, nerr);
    if ((t & 31) == 0) atomicAdd(&err_s, nerr);
    __syncthreads();
    if (t == 0) {
        if (a.err_out) a.err_out[f] = err_s;
        atomicAdd(a.err_total, (unsigned long long)err_s);
    }
}

// ---------------------------------------------------------------------------------------------
// k_gen_rayleigh: MisoRayleighFd (channel.py:187-292): IID CN(0,1) coefficients times the free-space
// loss evaluated at the NOMINAL receiver position (the attenuation is fixed at construction).
// ---------------------------------------------------------------------------------------------
struct RayleighArgs {
    const double* tx_pos; const double* freq_slot;
    double rx0[3]; double gain_lin; int skip_attenuation;
    uint64_t seed, frame_id0; uint32_t stream;
    float2* h_out; int n_ant;
    float* sscale_out;  // as ChanArgs::sscale_out: with grid.x == 1 the CTA also writes the MRT symbol scale
};

// grid (ceil(K / 8), F) or (1, F) with sscale_out; the coefficient of (frame, antenna k, slot p) is Box-Muller of Philox
// block (frame id, stream + (k << 8), 4 t + q), p = (2 q + e) 256 + t - the assignment of round 1, so the generated
// channels are the same up to rounding.  Attenuation = (path amplitude of antenna k) x (1 / f per slot): the per-slot
// reciprocals are computed once per thread and serve all antennas of the CTA.
__global__ void __launch_bounds__(256) k_gen_rayleigh(const RayleighArgs a) {
    __shared__ float att_s[MAX_ANT];
    const int per_cta = a.sscale_out ? a.n_ant : CHAN_ANT_PER_CTA;
    const int kb = blockIdx.x * per_cta, f = blockIdx.y, t = threadIdx.x;
    if (t < per_cta && kb + t < a.n_ant) {
        const int k = kb + t;
        const double dx = a.tx_pos[k * 3 + 0] - a.rx0[0], dy = a.tx_pos[k * 3 + 1] - a.rx0[1], dz = a.tx_pos[k * 3 + 2] - a.rx0[2];
        const double d = sqrt(dx * dx + dy * dy + dz * dz);
        att_s[t] = a.skip_attenuation ? 0.70710678118654752440f
                                      : (float)(0.70710678118654752440 * a.gain_lin * C_LIGHT / (4.0 * 3.14159265358979323846 * d));
    }
    float rf[8], nrm[8];
#pragma unroll
    for (int jp = 0; jp < 8; ++jp) {
        rf[jp] = a.skip_attenuation ? 1.0f : (float)(1.0 / a.freq_slot[jp * 256 + t]);
        nrm[jp] = 0.f;
    }
    __syncthreads();
    const uint64_t fid = a.frame_id0 + f;
    const int nk = min(per_cta, a.n_ant - kb);
    for (int i = 0; i < nk; ++i) {
        const int k = kb + i;
        const float att = att_s[i];
        float2* out = a.h_out + ((size_t)f * a.n_ant + k) * N_SC + t;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t r[4];
            philox4x32_10((uint32_t)fid, (uint32_t)(fid >> 32), a.stream + ((uint32_t)k << 8), (uint32_t)(t * 4 + q),
                          (uint32_t)a.seed, (uint32_t)(a.seed >> 32), r);
            const float2 g0 = box_muller_fast(r[0], r[1]), g1 = box_muller_fast(r[2], r[3]);
            const float s0 = att * rf[2 * q], s1 = att * rf[2 * q + 1];
            const float2 h0 = make_float2(g0.x * s0, g0.y * s0), h1 = make_float2(g1.x * s1, g1.y * s1);
            out[(2 * q) * 256] = h0;
            out[(2 * q + 1) * 256] = h1;
            // same operation order as the first pass of k_prep_realisation: bit-identical scale
            nrm[2 * q] = fmaf(h0.x, h0.x, fmaf(h0.y, h0.y, nrm[2 * q]));
            nrm[2 * q + 1] = fmaf(h1.x, h1.x, fmaf(h1.y, h1.y, nrm[2 * q + 1]));
        }
    }
    if (a.sscale_out) {
#pragma unroll
        for (int jp = 0; jp < 8; ++jp) a.sscale_out[(size_t)f * N_SC + jp * 256 + t] = (1.0f / sqrtf(nrm[jp])) * 0.015625f;
    }
}

// ---------------------------------------------------------------------------------------------
// k_csi_error: mp_model.py:264-284, h_hat = sqrt(1-eps^2) h + eps CN(0, mean_s |h_k[s]|^2) per antenna.
// ---------------------------------------------------------------------------------------------
struct CsiArgs {
    const float2* noise_in;  // [F][K][2048] injected unit normals (native slots) or nullptr -> Philox
    const float2* h_in; float2* h_out; float eps, scale;
    uint64_t seed, frame_id0; uint32_t stream; int n_ant;
    float* sscale_out;       // as ChanArgs::sscale_out (norm of the PERTURBED rows: the precoder is derived from them)
};

// grid (K, F), or (1, F) with sscale_out: the CTA walks all antennas of the realisation and also writes the symbol scale
#ifndef CSI_T
#define CSI_T 256  // threads per CTA; N_SC / CSI_T slots per thread (512 measured slower: 2.33 vs 1.96 ms per 4144 realisations)
#endif
constexpr int CSI_NS = N_SC / CSI_T, CSI_W = CSI_T / 32;
__global__ void __launch_bounds__(CSI_T) k_csi_error(const CsiArgs a) {
    __shared__ double red_s[CSI_W];
    const int f = blockIdx.y, t = threadIdx.x;
    // slot of this thread's element jp: (2 q + e) 256 + t' with t' = t mod 256, e = jp mod 2, q = t / 256 + (CSI_T / 256) (jp / 2):
    // both halves of a Philox block (round 1: block 4 t' + q -> slots (2 q) 256 + t' and (2 q + 1) 256 + t') stay in one thread
    auto slot = [&](const int jp) { return (2 * ((t >> 8) + (CSI_T / 256) * (jp >> 1)) + (jp & 1)) * 256 + (t & 255); };
    const int k0 = a.sscale_out ? 0 : blockIdx.x, k1 = a.sscale_out ? a.n_ant : blockIdx.x + 1;
    const uint64_t fid = a.frame_id0 + f;
    float nrm[CSI_NS];
#pragma unroll
    for (int jp = 0; jp < CSI_NS; ++jp) nrm[jp] = 0.f;
    for (int k = k0; k < k1; ++k) {
        const float2* in = a.h_in + ((size_t)f * a.n_ant + k) * N_SC;
        float2* out = a.h_out + ((size_t)f * a.n_ant + k) * N_SC;
        float2 h[CSI_NS];
        double pw = 0.0;
#pragma unroll
        for (int jp = 0; jp < CSI_NS; ++jp) {
            h[jp] = in[slot(jp)];
            pw += (double)h[jp].x * h[jp].x + (double)h[jp].y * h[jp].y;
        }
        pw = block_sum_n<CSI_W>(pw, red_s, t) / N_SC;
        const float sig = a.eps * (float)sqrt(0.5 * pw);
        float2 g[CSI_NS];
        if (a.noise_in) {
            const float2* nz = a.noise_in + ((size_t)f * a.n_ant + k) * N_SC;
#pragma unroll
            for (int jp = 0; jp < CSI_NS; ++jp) g[jp] = nz[slot(jp)];
        } else {
#pragma unroll
            for (int j2 = 0; j2 < CSI_NS / 2; ++j2) {
                const int q = (t >> 8) + (CSI_T / 256) * j2;
                uint32_t r[4];
                philox4x32_10((uint32_t)fid, (uint32_t)(fid >> 32), a.stream + ((uint32_t)k << 8), (uint32_t)((t & 255) * 4 + q),
                              (uint32_t)a.seed, (uint32_t)(a.seed >> 32), r);
                g[2 * j2] = box_muller_fast(r[0], r[1]);
                g[2 * j2 + 1] = box_muller_fast(r[2], r[3]);
            }
        }
#pragma unroll
        for (int jp = 0; jp < CSI_NS; ++jp) {
            const float2 o = make_float2(fmaf(a.scale, h[jp].x, sig * g[jp].x), fmaf(a.scale, h[jp].y, sig * g[jp].y));
            out[slot(jp)] = o;
            nrm[jp] = fmaf(o.x, o.x, fmaf(o.y, o.y, nrm[jp]));  // operation order of k_prep_realisation's first pass
        }
    }
    if (a.sscale_out) {
#pragma unroll
        for (int jp = 0; jp < CSI_NS; ++jp) a.sscale_out[(size_t)f * N_SC + slot(jp)] = (1.0f / sqrtf(nrm[jp])) * 0.015625f;
    }
}

}  // namespace mmimo
