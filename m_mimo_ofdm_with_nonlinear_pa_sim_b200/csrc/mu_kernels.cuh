// Two-user realisation kernel: multi-user MRT / zero-forcing precoding (antenna_array.py:202-257), the array's PA
// re-normalisation (antenna_array.py:313-360) and the per-user AGC / noise scalers of
// main_multiuser/main_multiuser_mcnc_ber_vs_ebn0.py:580-622 (same maths as corrector.py:374-400) for one channel
// realisation (both users) per CTA.  Arrays are ordered by receiver INSTANCE = (realisation, user).
#pragma once
#include "prep_kernels.cuh"

namespace mmimo {

enum MuPrecoding : int { MU_MRT = 0, MU_ZF = 1 };

struct MuPrepArgs {
    const float2* h;        // [R][2][K][2048] channel of user u towards the array (native slots)
    float2* v;              // [R][2][K][2048] precoder of user u (unscaled; the frame kernel applies 1/64)
    float2* agc_inv;        // [R][2][2048]  1 / (64 * sum_k alpha_k h_{u,k} v_{u,k})
    FrameScalars* scal;     // [R][2]
    float* alpha_out;       // [R][K] or nullptr
    int n_ant, n_real, precoding, pa_kind;
    int loop_cnc;           // receiver loop = single-antenna CNC model: its PA sees avg_symbol_power * n_sc / n_fft (corrector.py:34-35)
    double ibo_lin, avg_symbol_power, avg_sample_power, snr_lin;
};

// 512 threads, four sub-carrier slots per thread (MU_NS): half the fp64 accumulators per thread of the 256 x 8 form, so that two
// CTAs = 1024 threads per SM are resident and twice as many row loads are in flight
constexpr int MU_T = 512, MU_NS = N_SC / MU_T, MU_W = MU_T / 32;
__global__ void __launch_bounds__(MU_T, 2) k_prep_mu(const MuPrepArgs a) {
    __shared__ float part_s[MAX_ANT][MU_W];
    __shared__ float alpha_s[MAX_ANT];
    __shared__ double vkpow_s[MAX_ANT];
    __shared__ double red_s[MU_W];
    const int r = blockIdx.x, t = threadIdx.x, K = a.n_ant;
    const float2* __restrict__ H0 = a.h + (size_t)(2 * r) * K * N_SC;
    const float2* __restrict__ H1 = H0 + (size_t)K * N_SC;
    float2* V0 = a.v + (size_t)(2 * r) * K * N_SC;
    float2* V1 = V0 + (size_t)K * N_SC;

    // per sub-carrier Gram matrix of T = [h_0 h_1] (K x 2): G[a][b] = sum_k h_a conj(h_b)   (antenna_array.py:236-240)
    double g00[MU_NS], g11[MU_NS], g01r[MU_NS], g01i[MU_NS];
#pragma unroll
    for (int jp = 0; jp < MU_NS; ++jp) g00[jp] = g11[jp] = g01r[jp] = g01i[jp] = 0.0;
    for (int k = 0; k < K; ++k) {
#pragma unroll
        for (int jp = 0; jp < MU_NS; ++jp) {
            const float2 x = H0[(size_t)k * N_SC + jp * MU_T + t], y = H1[(size_t)k * N_SC + jp * MU_T + t];
            g00[jp] += (double)x.x * x.x + (double)x.y * x.y;
            g11[jp] += (double)y.x * y.x + (double)y.y * y.y;
            g01r[jp] += (double)x.x * y.x + (double)x.y * y.y;   // x conj(y)
            g01i[jp] += (double)x.y * y.x - (double)x.x * y.y;
        }
    }
    // per sub-carrier coefficients: v_0 = c00 conj(h_0) + c10 conj(h_1), v_1 = c01 conj(h_0) + c11 conj(h_1)
    double c00[MU_NS], c11[MU_NS], c01r[MU_NS], c01i[MU_NS];  // c10 = conj(c01); c00, c11 real
#pragma unroll
    for (int jp = 0; jp < MU_NS; ++jp) {
        if (a.precoding == MU_MRT) {  // common normalisation over users, antenna_array.py:202-220
            c00[jp] = c11[jp] = 1.0 / sqrt(g00[jp] + g11[jp]);
            c01r[jp] = c01i[jp] = 0.0;
        } else {
            // inv(G) = 1/det [[g11, -g01], [-conj(g01), g00]];  V = conj(T) inv(G), then unit Frobenius norm per
            // sub-carrier (antenna_array.py:240-257; the sqrt(K-U) factor cancels in the normalisation).
            // ||V||_F^2 = trace(inv(G)^H G inv(G)) = trace(inv(G)) = (g00 + g11) / det       (G is Hermitian)
            const double det = g00[jp] * g11[jp] - (g01r[jp] * g01r[jp] + g01i[jp] * g01i[jp]);
            const double tr = g00[jp] + g11[jp];
            if (det > 1e-15 * tr * tr) {
                const double nrm = 1.0 / (det * sqrt(tr / det));
                c00[jp] = g11[jp] * nrm;
                c11[jp] = g00[jp] * nrm;
                c01r[jp] = -g01r[jp] * nrm;   // inv(G)[0][1] = -g01 / det ; inv(G)[1][0] is its conjugate
                c01i[jp] = -g01i[jp] * nrm;
            } else {
                // Collinear users (e.g. co-located): G is singular, np.linalg.inv raises and the reference falls back to
                // np.linalg.pinv (antenna_array.py:243-251; smallest / largest eigenvalue ~ det / tr^2 under pinv's 1e-15
                // cut-off).  Rank one: G = tr u u^H, so pinv(G) = G / tr^2 and ||V||_F^2 = trace(pinv(G)) = 1 / tr.
                const double nrm = 1.0 / (tr * sqrt(tr));
                c00[jp] = g00[jp] * nrm;
                c11[jp] = g11[jp] * nrm;
                c01r[jp] = g01r[jp] * nrm;
                c01i[jp] = g01i[jp] * nrm;
            }
        }
    }
    // precoders out + per-antenna power summed over users (corrector.py:383)
    for (int k = 0; k < K; ++k) {
        float pw = 0.f;
#pragma unroll
        for (int jp = 0; jp < MU_NS; ++jp) {
            const int p = jp * MU_T + t;
            const float2 x = H0[(size_t)k * N_SC + p], y = H1[(size_t)k * N_SC + p];
            // v_0 = c00 conj(h_0) + conj(c01) conj(h_1),  v_1 = c01 conj(h_0) + c11 conj(h_1); formed in double: the two
            // terms of a zero-forcing precoder nearly cancel
            const double xr = x.x, xi = -(double)x.y, yr = y.x, yi = -(double)y.y;
            const float2 v0 = make_float2((float)(c00[jp] * xr + (c01r[jp] * yr + c01i[jp] * yi)),
                                          (float)(c00[jp] * xi + (c01r[jp] * yi - c01i[jp] * yr)));
            const float2 v1 = make_float2((float)((c01r[jp] * xr - c01i[jp] * xi) + c11[jp] * yr),
                                          (float)((c01r[jp] * xi + c01i[jp] * xr) + c11[jp] * yi));
            V0[(size_t)k * N_SC + p] = v0;
            V1[(size_t)k * N_SC + p] = v1;
            pw += v0.x * v0.x + v0.y * v0.y + v1.x * v1.x + v1.y * v1.y;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) pw += __shfl_xor_sync(0xffffffffu, pw, o);
        if ((t & 31) == 0) part_s[k][t >> 5] = pw;
    }
    __syncthreads();
    if (t < K) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < MU_W; ++w) s += (double)part_s[t][w];
        vkpow_s[t] = s;
        const double g2 = a.ibo_lin * (double)N_SC / (s * (double)K);   // corrector.py:390-394
        const double g = sqrt(g2);
        const double al = 1.0 - exp(-g2) + (1.7724538509055160273 * g * 0.5) * erfc(g);
        alpha_s[t] = (float)al;
        if (a.alpha_out) a.alpha_out[(size_t)r * K + t] = (float)al;
    }
    __syncthreads();
    // per-user AGC vectors  sum_k alpha_k h_{u,k} v_{u,k}  (script :598-611).  With v_0 = c00 conj(h_0) + conj(c01) conj(h_1)
    // and v_1 = c01 conj(h_0) + c11 conj(h_1) the sums are combinations of the alpha-weighted Gram entries
    //   S00 = sum_k alpha_k |h_0|^2, S11 = sum_k alpha_k |h_1|^2, S01 = sum_k alpha_k h_0 conj(h_1):
    //   a_0 = c00 S00 + conj(c01) S01,   a_1 = c01 conj(S01) + c11 S11
    // - one more pass over the channel rows only (the precoder rows are not read back: 4.4 GB less per 2072 realisations)
    double a0r[MU_NS], a0i[MU_NS], a1r[MU_NS], a1i[MU_NS];
    {
        double s00[MU_NS], s11[MU_NS], s01r[MU_NS], s01i[MU_NS];
#pragma unroll
        for (int jp = 0; jp < MU_NS; ++jp) s00[jp] = s11[jp] = s01r[jp] = s01i[jp] = 0.0;
        for (int k = 0; k < K; ++k) {
            const double al = alpha_s[k];
#pragma unroll
            for (int jp = 0; jp < MU_NS; ++jp) {
                const int p = jp * MU_T + t;
                const float2 x = H0[(size_t)k * N_SC + p], y = H1[(size_t)k * N_SC + p];
                s00[jp] += al * ((double)x.x * x.x + (double)x.y * x.y);
                s11[jp] += al * ((double)y.x * y.x + (double)y.y * y.y);
                s01r[jp] += al * ((double)x.x * y.x + (double)x.y * y.y);
                s01i[jp] += al * ((double)x.y * y.x - (double)x.x * y.y);
            }
        }
#pragma unroll
        for (int jp = 0; jp < MU_NS; ++jp) {
            a0r[jp] = c00[jp] * s00[jp] + (c01r[jp] * s01r[jp] + c01i[jp] * s01i[jp]);   // conj(c01) S01
            a0i[jp] = c01r[jp] * s01i[jp] - c01i[jp] * s01r[jp];
            a1r[jp] = (c01r[jp] * s01r[jp] + c01i[jp] * s01i[jp]) + c11[jp] * s11[jp];   // c01 conj(S01)
            a1i[jp] = c01i[jp] * s01r[jp] - c01r[jp] * s01i[jp];
        }
    }
    double s0 = 0.0, s1 = 0.0;
#pragma unroll
    for (int jp = 0; jp < MU_NS; ++jp) {
        const int p = jp * MU_T + t;
        const double a0 = a0r[jp] * a0r[jp] + a0i[jp] * a0i[jp];
        const double a1 = a1r[jp] * a1r[jp] + a1i[jp] * a1i[jp];
        s0 += a0; s1 += a1;
        a.agc_inv[(size_t)(2 * r) * N_SC + p] = make_float2((float)(a0r[jp] / (64.0 * a0)), (float)(-a0i[jp] / (64.0 * a0)));
        a.agc_inv[(size_t)(2 * r + 1) * N_SC + p] = make_float2((float)(a1r[jp] / (64.0 * a1)), (float)(-a1i[jp] / (64.0 * a1)));
    }
    s0 = block_sum_n<MU_W>(s0, red_s, t);
    s1 = block_sum_n<MU_W>(s1, red_s, t);
    if (t == 0) {
        double tot = 0.0;
        for (int k = 0; k < K; ++k) tot += vkpow_s[k];
        const double gain = tot / ((double)K * N_SC);  // antenna_array.py:328-341: users summed per antenna
        const double avg = a.avg_sample_power * gain;
        const double p0 = (a.pa_kind == PA_TOI) ? (1.0 / a.ibo_lin) / avg : a.ibo_lin * avg;
        FrameScalars sc;
        sc.pa_p0_tx = sc.pa_p0_loop = (float)p0;
        if (a.loop_cnc) sc.pa_p0_loop = (float)((a.pa_kind == PA_TOI) ? (1.0 / a.ibo_lin) / a.avg_sample_power : a.ibo_lin * a.avg_sample_power);
        sc.noise_sigma_clean = 0.f;
        sc.noise_sigma = (float)(0.5 * sqrt(2.0 * a.avg_symbol_power * (s0 / N_SC) / a.snr_lin));  // noise.py:64-66
        a.scal[2 * r] = sc;
        sc.noise_sigma = (float)(0.5 * sqrt(2.0 * a.avg_symbol_power * (s1 / N_SC) / a.snr_lin));
        a.scal[2 * r + 1] = sc;
    }
}

// loop-side PA parameter of the single-user receivers: scal_mu[i].pa_p0_loop = scal_su[i].pa_p0_tx, or `fixed` (> 0) for all
__global__ void k_copy_loop_p0(FrameScalars* mu, const FrameScalars* su, int n, float fixed) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) mu[i].pa_p0_loop = fixed > 0.f ? fixed : su[i].pa_p0_tx;
}

}  // namespace mmimo
