// Fused Monte-Carlo frame kernel: one CTA (256 threads) runs whole frames
//   bits -> QAM -> per-antenna precode -> IFFT -> PA -> FFT -> channel MAC over antennas -> AWGN -> AGC
//   -> [detect -> count errors -> regenerate all antennas -> subtract distortion estimate] x (I+1)
// i.e. the loop body of mp_model.py:208-222 with McncReceiver.receive (corrector.py:165-207) or
// CncReceiver.receive (corrector.py:52-112) inlined.  The sum over antennas stays in registers: every
// thread owns the same 8 data sub-carriers (native slot order, link_common.cuh) for every antenna.
//
// Precoder: either explicit rows V[k] (any precoder), or - the MRT case of antenna_array.py:170-171 -
// derived on the fly from the channel-knowledge rows, v_k[s] = conj(h_k[s]) * sscale[s] with
// sscale[s] = 1 / (64 ||h[:,s]||) written once per realisation by k_prep_realisation.  The derived form
// halves the HBM traffic of the kernel (one 16 KB row per antenna and pass instead of two).
#pragma once
#include "link_common.cuh"

namespace mmimo {

struct FrameScalars {   // written per frame by k_prep_realisation
    float pa_p0_tx;     // PA parameter of the real transmission (sat. power or cubic coefficient)
    float pa_p0_loop;   // PA parameter inside the receiver's regeneration loop
    float noise_sigma;  // std of each real noise component on a data sub-carrier (noise.py:64-66)
    float noise_sigma_clean;
};

// one "array transmit + propagate" description (antenna_array.py:58-140 + channel.py:74-89)
struct TxSide {
    const float2* h_mac; long long h_mac_fs;   // [F][K][2048] channel applied to the transmitted signal
    const float2* h_pre; long long h_pre_fs;   // [F][K][2048] channel knowledge the MRT precoder is derived from
    const float2* v;     long long v_fs;       // [F][K][2048] explicit precoder (unscaled) or nullptr -> derive
    const float* sscale; long long sscale_fs;  // [F][2048] symbol scale: 1/64 (explicit) or 1/(64 ||h||) (derived)
    const float2* agc_inv; long long agc_fs;   // [F][2048] 1 / (64 * AGC)
    int k;                                     // antennas
};

struct FrameArgs {
    TxSide tx;    // pass 0: the real transmission + mp_model.py:214
    TxSide loop;  // passes 1..I: the receiver's model of the transmitter (MCNC: K antennas; CNC: 1, h = 1)
    const FrameScalars* scal; long long scal_fs;  // per frame (stride 1) or shared (stride 0)
    float pa_p1;                                     // Rapp hardness
    int n_iter;                                      // I: detections 0..I are produced
    Qam qam;
    int n_frames;
    // inputs: injected or generated (Philox)
    const uint8_t* tx_idx;       // [F][2048] native slots, or nullptr
    const float2* noise;         // [F][2048] native slots (added before the AGC), or nullptr
    int noise_unit;              // injected noise holds unit normals: scale by the frame's noise_sigma (noise.py:64-66)
    const float2* rx_eq_in;      // [F][2048] equalised received signal: skips pass 0 (receiver-only use), or nullptr
    uint64_t seed; uint64_t frame_id0;
    // outputs
    uint8_t* det_out;            // [F][I+1][2048] native slots, or nullptr
    float2* rx_eq_out;           // [F][2048] equalised received data sub-carriers, or nullptr
    float2* corr_out;            // [F][I+1][2048] corrected symbols fed to the slicer, or nullptr
    uint32_t* err_out;           // [F][I+1], or nullptr
    unsigned long long* err_total;  // [I+1] accumulated with atomics
    float2* rxeq_scratch;        // [gridDim.x][2048] per-CTA scratch for the equalised received signal (stays in L2)
    // Two-user mode (MU instantiation only; corrector.py:348-489 McncMuReceiver, antenna_array.py:202-257): frames are
    // receiver INSTANCES ordered (realisation, user); instance f ^ 1 is the other user of the same realisation.  Both
    // sides use explicit precoders v = [inst][K][2048]; every array transmit sends own + other user
    // (modulation.py:373), the other user's symbols being the TRUE ones in the receiver loop (corrector.py:437-440).
    // loop_single_user != 0: the loop side models a single-user array instead (main_multiuser_mcnc_ber_vs_ebn0.py:563-572).
    int loop_single_user;
    // multi-point batches (mmimo_link_sweep): frame f counts into err_total[64 * ((frame0 + f) / fpp) + iteration]
    int fpp; unsigned frame0;
    // k_link_frame_t64 only: cycles by which antenna groups 2 and 3 start every pass after groups 0 and 1.  The two
    // warps that share a scheduler belong to groups g and g + 2; started together they run the same instruction
    // stream in lock-step, so their shared-memory exchange phases (no FMA work) coincide.  The offset makes one warp's
    // exchange overlap the other's butterflies (frame_kernel_t64.cuh).
    int stagger;
    // k_link_frame_t64 only - signal-to-distortion snapshots (main_beampatterns_plotting/sdr_mp.py:124-151) instead of the
    // receiver: with n_iter = 0 and these two set, pass 0 ends in
    //     desired[f]    = sum_s | sum_k alpha_k h_k X_clean_k |^2 = sum_s | sym[s] agc[s] |^2     (agc = the AGC vector of
    //     distortion[f] = sum_s | sum_k h_k X_k - sym[s] agc[s] |^2                               mp_model.py:290-329)
    // no noise is added, nothing is detected or counted.
    double* sdr_desired; double* sdr_distortion;
};

constexpr int FRAME_SMEM_FLOAT2 = 4096 * 2 + 256 + 2048;
constexpr int FRAME_SMEM_BYTES = FRAME_SMEM_FLOAT2 * 8 + 64 * 4;
constexpr int FRAME_SMEM_BYTES_MU = FRAME_SMEM_BYTES + N_SC * 8;  // + scaled symbols of the other user

// packed PA (distortion.py:9-19 soft limiter, :102-113 Rapp, :202-211 third order)
template <int KIND>
__device__ __forceinline__ pf2 pa_apply_pk(pf2 xp, const float p0, const float sqrt_p0, const float p1) {
    if (KIND == PA_NONE) return xp;
    const float2 x = unpk(xp);
    const float p = fmaf(x.x, x.x, x.y * x.y);
    float g;
    if (KIND == PA_SOFT_LIMITER) {
        float rs;  // MUFU.RSQ without the denormal fix-up of rsqrtf (p = 0 or denormal -> inf -> gain 1)
        asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rs) : "f"(p));
        g = fminf(1.0f, sqrt_p0 * rs);  // p <= p0 -> exactly 1
    } else if (KIND == PA_RAPP) {
        const float r = __powf(p / p0, p1);
        g = __powf(1.0f + r, -0.5f / p1);
    } else {
        g = 1.0f - p0 * p;
    }
    return pmul(xp, pbc(g));
}

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// One antenna: x (8 data slots, already precoded, 1/64 folded in) -> IFFT -> PA -> FFT; on return
// x[0..3], x[12..15] hold the data bins of this thread (thread 0: x[4] = bin 1024).
template <int PA, class HOOK = NoHook>
__device__ __forceinline__ void antenna_chain(pf2 (&x)[16], pf2* bufA, pf2* bufB, const float2* tw2, const Tw1& tw1,
                                              const int t, const float p0, const float sqrt_p0, const float p1,
                                              const HOOK before_last_pass = HOOK()) {
    fft4096<+1, true, false>(x, bufA, bufB, tw2, tw1, t);
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = pa_apply_pk<PA>(x[j], p0, sqrt_p0, p1);
    fft4096<-1, false, true>(x, bufA, bufB, tw2, tw1, t, SyncCta(), before_last_pass);
}

template <int PA, bool MU = false>
__global__ void __launch_bounds__(FFT_THREADS, 2) k_link_frame(const FrameArgs a, const float2* __restrict__ tw4096,
                                                               const float2* __restrict__ tw256) {
    extern __shared__ float2 smem[];
    pf2* bufA = reinterpret_cast<pf2*>(smem);
    pf2* bufB = reinterpret_cast<pf2*>(smem + 4096);
    float2* tw2 = smem + 8192;
    float2* symn_s = tw2 + 256;      // scaled current symbols, thread-private addresses jp*256+t
    // equalised received signal: written once per frame, read once per pass, thread-private addresses -> global
    // scratch (L2-resident) instead of 16 KB of shared memory, which leaves the L1 room for the channel rows
    float2* rxeq_s = a.rxeq_scratch + (size_t)blockIdx.x * N_SC;
    unsigned int* err_s = reinterpret_cast<unsigned int*>(symn_s + 2048);
    float2* symo_s = reinterpret_cast<float2*>(err_s + 64);  // MU only: scaled symbols of the other user

    const int t = threadIdx.x;
    tw2[t] = tw256[t];
    if (t < 64) err_s[t] = 0;
    Tw1 tw1;
    load_tw1(tw1, tw4096, t);
    __syncthreads();

    for (int f = blockIdx.x; f < a.n_frames; f += gridDim.x) {
        const FrameScalars sc = a.scal[f * a.scal_fs];
        // ---- transmitted symbols -------------------------------------------------------------
        uint32_t txw[2], otw[2] = {0u, 0u};
        auto load_tx = [&](const int g, uint32_t (&w)[2]) {  // 8 symbol values of this thread for instance g
            if (a.tx_idx) {
                const uint8_t* p = a.tx_idx + (size_t)g * N_SC;
                w[0] = w[1] = 0;
#pragma unroll
                for (int jp = 0; jp < 8; ++jp) w[jp >> 2] |= (uint32_t)p[jp * 256 + t] << (8 * (jp & 3));
            } else {
                uint32_t r[4];
                const uint64_t fid = a.frame_id0 + g;
                philox4x32_10((uint32_t)fid, (uint32_t)(fid >> 32), RNG_BITS, (uint32_t)t, (uint32_t)a.seed,
                              (uint32_t)(a.seed >> 32), r);
                const uint32_t m = (1u << (2 * a.qam.log2side)) - 1u;
                const uint32_t m4 = m | (m << 8) | (m << 16) | (m << 24);
                w[0] = r[0] & m4;
                w[1] = r[1] & m4;
            }
        };
        load_tx(f, txw);
        if (MU) load_tx(f ^ 1, otw);
        uint32_t curw[2] = {txw[0], txw[1]};  // symbols fed to the next array transmit

        for (int pass = 0; pass <= a.n_iter; ++pass) {
            const bool first = (pass == 0);
            const TxSide& sd = first ? a.tx : a.loop;
            const int K = (first && a.rx_eq_in) ? 0 : sd.k;  // receiver-only entry: no transmission in pass 0
            const float2* __restrict__ Hm = sd.h_mac + f * sd.h_mac_fs;
            const float2* __restrict__ Hp = sd.v ? sd.v + f * sd.v_fs : sd.h_pre + f * sd.h_pre_fs;
            const bool explicit_v = sd.v != nullptr;
            const bool two_usr = MU && !(a.loop_single_user && !first);
            const float2* __restrict__ Ho = (MU && sd.v) ? sd.v + (f ^ 1) * sd.v_fs : nullptr;  // precoder rows of the other user
            const float p0 = first ? sc.pa_p0_tx : sc.pa_p0_loop;
            const float sqrt_p0 = sqrtf(p0);
            {
                const float* __restrict__ ss = sd.sscale + f * sd.sscale_fs;
#pragma unroll
                for (int jp = 0; jp < 8; ++jp) {
                    const int p = jp * 256 + t;
                    const float2 s = qam_symbol(a.qam, (curw[jp >> 2] >> (8 * (jp & 3))) & 0xffu);
                    const float g = ss[p];
                    symn_s[p] = make_float2(s.x * g, s.y * g);
                    if (MU) {
                        const float2 o = qam_symbol(a.qam, (otw[jp >> 2] >> (8 * (jp & 3))) & 0xffu);
                        symo_s[p] = make_float2(o.x * g, o.y * g);
                    }
                }
            }
            pf2 acc[8];
#pragma unroll
            for (int jp = 0; jp < 8; ++jp) acc[jp] = pk(0.f, 0.f);

            pf2 rown[8];  // precoder / channel-knowledge row of the NEXT antenna, loaded one transform pass ahead
            if (K > 0) {
#pragma unroll
                for (int jp = 0; jp < 8; ++jp) rown[jp] = pk2(__ldg(Hp + jp * 256 + t));
            }
            for (int k = 0; k < K; ++k) {
                const float2* __restrict__ pr = Hp + (size_t)k * N_SC;
                const float2* __restrict__ hk = Hm + (size_t)k * N_SC;
                if (k + 2 < K) {  // pull the rows of antenna k+2 from HBM into L2 while this antenna computes
                    const float2* qa = pr + 2 * N_SC;
                    const float2* qb = hk + 2 * N_SC;
                    if (t < 128) prefetch_l2(qa + t * 16);
                    else if (qb != qa) prefetch_l2(qb + (t - 128) * 16);
                }
                pf2 x[16];
                if (explicit_v) {
#pragma unroll
                    for (int jp = 0; jp < 8; ++jp) {
                        const float2 s = symn_s[jp * 256 + t];
                        x[jp_to_j(jp)] = pcmul(rown[jp], s.x, s.y);
                        if (MU && two_usr)  // modulation.py:373: sum over users of symbol * precoder
                            x[jp_to_j(jp)] = pcfma(pk2(__ldg(Ho + (size_t)k * N_SC + jp * 256 + t)), pk2(symo_s[jp * 256 + t]), x[jp_to_j(jp)]);
                    }
                } else {
#pragma unroll
                    for (int jp = 0; jp < 8; ++jp) {  // conj(h) * (c + j s) = (h.x c + h.y s, h.x s - h.y c)
                        const float2 s = symn_s[jp * 256 + t];
                        const pf2 h = rown[jp];
                        x[jp_to_j(jp)] = pfma(pswap(h), pbc(s.y), pmul(h, pk(s.x, -s.x)));
                    }
                }
#pragma unroll
                for (int j = 4; j < 12; ++j) x[j] = pk(0.f, 0.f);
                if (t == 0) { x[4] = x[0]; x[0] = pk(0.f, 0.f); }
                const float2* __restrict__ prn = pr + (k + 1 < K ? N_SC : 0);
                antenna_chain<PA>(x, bufA, bufB, tw2, tw1, t, p0, sqrt_p0, a.pa_p1, [&]() {
#pragma unroll
                    for (int jp = 0; jp < 8; ++jp) rown[jp] = pk2(__ldg(prn + jp * 256 + t));
                });
                if (t == 0) x[0] = x[4];
#pragma unroll
                for (int jp = 0; jp < 8; ++jp) acc[jp] = pcfma(x[jp_to_j(jp)], pk2(__ldg(hk + jp * 256 + t)), acc[jp]);
            }

            // ---- end of pass: equalise / estimate distortion / slice / count --------------------
            uint32_t detw[2] = {0u, 0u};
            const float2* __restrict__ ai = sd.agc_inv + f * sd.agc_fs;
            float2 nz[8];
            if (first && !a.rx_eq_in) {
                if (a.noise) {
                    const float ns = a.noise_unit ? sc.noise_sigma : 1.0f;
#pragma unroll
                    for (int jp = 0; jp < 8; ++jp) {
                        const float2 u = a.noise[(size_t)f * N_SC + jp * 256 + t];
                        nz[jp] = make_float2(u.x * ns, u.y * ns);
                    }
                } else {
                    const uint64_t fid = a.frame_id0 + f;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        uint32_t r[4];
                        philox4x32_10((uint32_t)fid, (uint32_t)(fid >> 32), RNG_NOISE, (uint32_t)(t * 4 + q),
                                      (uint32_t)a.seed, (uint32_t)(a.seed >> 32), r);
                        const float2 g0 = box_muller(r[0], r[1]), g1 = box_muller(r[2], r[3]);
                        nz[2 * q] = make_float2(g0.x * sc.noise_sigma, g0.y * sc.noise_sigma);
                        nz[2 * q + 1] = make_float2(g1.x * sc.noise_sigma, g1.y * sc.noise_sigma);
                    }
                }
            }
            unsigned int nerr = 0;
#pragma unroll
            for (int jp = 0; jp < 8; ++jp) {
                const int p = jp * 256 + t;
                const float2 ac = unpk(acc[jp]);
                float2 corr;
                if (first) {
                    if (a.rx_eq_in) {  // McncReceiver.receive / CncReceiver.receive entry (corrector.py:176,66)
                        corr = a.rx_eq_in[(size_t)f * N_SC + p];
                    } else {
                        const float2 r = make_float2(fmaf(64.f, nz[jp].x, ac.x), fmaf(64.f, nz[jp].y, ac.y));
                        corr = cmul(r, ai[p]);
                    }
                    rxeq_s[p] = corr;
                    if (a.rx_eq_out) a.rx_eq_out[(size_t)f * N_SC + p] = corr;
                } else {
                    const float2 z = cmul(ac, ai[p]);
                    const float2 sy = qam_symbol(a.qam, (curw[jp >> 2] >> (8 * (jp & 3))) & 0xffu);
                    const float2 d = csub(z, sy);   // distortion estimate, corrector.py:205
                    corr = csub(rxeq_s[p], d);      // corrector.py:184
                }
                if (a.corr_out) a.corr_out[((size_t)f * (a.n_iter + 1) + pass) * N_SC + p] = corr;
                const uint32_t idx = qam_detect(a.qam, corr);
                detw[jp >> 2] |= idx << (8 * (jp & 3));
            }
            curw[0] = detw[0];
            curw[1] = detw[1];
            nerr = __popc(detw[0] ^ txw[0]) + __popc(detw[1] ^ txw[1]);
            nerr = __reduce_add_sync(0xffffffffu, nerr);
            if ((t & 31) == 0) atomicAdd(&err_s[pass], nerr);
            if (a.det_out) {
                uint8_t* dp = a.det_out + ((size_t)f * (a.n_iter + 1) + pass) * N_SC;
#pragma unroll
                for (int jp = 0; jp < 8; ++jp) dp[jp * 256 + t] = (uint8_t)((detw[jp >> 2] >> (8 * (jp & 3))) & 0xffu);
            }
        }
        __syncthreads();
        if (t <= a.n_iter) {
            const unsigned int e = err_s[t];
            err_s[t] = 0;
            if (a.err_out) a.err_out[(size_t)f * (a.n_iter + 1) + t] = e;
            atomicAdd(&a.err_total[t + (MU ? (f & 1) * 32 : 0) + (a.fpp ? 64u * ((a.frame0 + (unsigned)f) / (unsigned)a.fpp) : 0u)], (unsigned long long)e);  // MU: user 1 counts at [32..]
        }
        __syncthreads();
    }
}

}  // namespace mmimo
