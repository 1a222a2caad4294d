// CncReceiver.receive (corrector.py:52-112) as a kernel of its own: the receiver loop of the single-antenna model, one
// FRAME PER 64-THREAD GROUP.
//
// k_link_frame_t64 runs a CNC frame as pass 0 (the real transmission: K antenna chains over the CTA's four groups) followed
// by I loop passes of ONE chain each (h = v = 1), during which three of the four groups idle: 16 + 5 chain slots per frame
// at K = 64, I = 5 where 17.25 would do.  Here the frame kernel only runs pass 0 (n_iter = 0, rx_eq_out) and this kernel
// takes the equalised received signal from there (FrameArgs::rx_eq_in): every group owns a frame, runs its I regenerations
// (symbols -> transform -> PA -> transform, frame_kernel_t64.cuh's loop body) and all I + 1 detections on its own 32 slots
// per thread (slot lane + 64 i <-> transform register i or i + 32), with no CTA-wide barrier at all.  The four groups of a
// CTA keep the exchange token ring of the frame kernel (their stores into shared memory take turns).
// The arithmetic per slot is the frame kernel's, operation by operation (unit channel: the MAC and the meeting of the
// groups are exact identities), so both forms give identical decisions and counters (tests/test_gpu_link_parity.py).
#pragma once
#include "frame_kernel_t64.cuh"

namespace mmimo {

#ifndef CNC_SPLIT
#define CNC_SPLIT 1  // 0: CNC frames stay entirely in k_link_frame_t64 (round-2 form)
#endif
constexpr int CNC_LOOP_SMEM_BYTES = (4096 + R64_GROUPS * (t64::XBUF + N_SC)) * 8;  // table + per group: exchange buffer, received signal
static_assert(CNC_LOOP_SMEM_BYTES <= 227 * 1024, "shared-memory budget of one CTA");

template <int PA>
// uni_ss / uni_agc: the loop side's symbol scale and 1 / (64 alpha) when they are the same for every sub-carrier (they are for
// CncReceiver: h = v = 1, corrector.py:104-107); uniform == 0 reads a.loop.sscale / a.loop.agc_inv per slot instead.
// mode bit 0: two-user batch, frames are receiver instances (realisation, user) and user 1 counts at err_total[32 + pass];
// mode bit 1: CncMuReceiver (corrector.py:288-345): the regenerated signal is v (detected symbol) + v (TRUE symbol of the other
// user, instance f ^ 1) with the explicit loop-side precoder a.loop.v (uni_v when uniform: sqrt(2)/2, corrector.py:322).
__global__ void __launch_bounds__(FFT_THREADS, 1) k_cnc_loop_t64(const FrameArgs a, const float2* __restrict__ tab_g, const int uniform,
                                                                 const float uni_ss, const float2 uni_agc, const int mode,
                                                                 const float2 uni_v) {
    extern __shared__ __align__(128) float2 smem[];
    float2* tab = smem;
    const int t = threadIdx.x, lane = t & 63, g = t >> 6;
    pf2* buf = reinterpret_cast<pf2*>(smem + 4096) + g * t64::XBUF;
    float2* sym_s = reinterpret_cast<float2*>(buf) + N_SC;  // scaled symbols of the coming regeneration (idle half of the buffer)
    float2* rx_s = smem + 4096 + R64_GROUPS * t64::XBUF + g * N_SC + lane;  // equalised received signal of the group's frame
    const SyncGroup sy{g + 1};
    for (int i = t; i < t64::TAB_FLOAT2; i += FFT_THREADS) tab[i] = tab_g[i];
    __syncthreads();
    const int ring_next = (0x0321 >> (4 * g)) & 3;
    if (ring_next == 0) asm volatile("bar.arrive %0, 128;" ::"r"(5) : "memory");
    auto ring_acquire = [&]() { asm volatile("bar.sync %0, 128;" ::"r"(5 + g) : "memory"); };
    auto ring_release = [&]() { asm volatile("bar.arrive %0, 128;" ::"r"(5 + ring_next) : "memory"); };
    const t64::TabA tab_a = t64::load_tab_a(tab, lane);
    const int mir = 64 - lane;
    const int mir0 = lane == 0 ? 1024 : mir + 64 * 31;
    const int mir16 = lane == 0 ? 0 : mir + 64 * 15;
    const TxSide& sd = a.loop;

    for (int base = blockIdx.x * R64_GROUPS; base < a.n_frames; base += gridDim.x * R64_GROUPS) {
        const int f = base + g;
        if (f >= a.n_frames) {  // no frame for this group in the CTA's last round: the token still has to travel
            for (int e = 0; e < 2 * a.n_iter; ++e) { ring_acquire(); ring_release(); }
            continue;
        }
        const FrameScalars sc = a.scal[f * a.scal_fs];
        // symbol values of the 32 slots lane + 64 i: byte i & 3 of word i >> 2.  Slot p belongs to Philox block p & 255, byte p >> 8
        uint32_t txw[8], curw[8], otw[8];
        auto load_tx32 = [&](const int inst, uint32_t (&w)[8]) {
#pragma unroll
            for (int q = 0; q < 8; ++q) w[q] = 0;
            if (a.tx_idx) {
                const uint8_t* p = a.tx_idx + (size_t)inst * N_SC + lane;
#pragma unroll
                for (int i = 0; i < 32; ++i) w[i >> 2] |= (uint32_t)p[64 * i] << (8 * (i & 3));
            } else {
                const uint64_t fid = a.frame_id0 + inst;
                const uint32_t m = (1u << (2 * a.qam.log2side)) - 1u;
                const uint32_t m4 = m | (m << 8) | (m << 16) | (m << 24);
                uint32_t blk[4][2];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    uint32_t r[4];
                    philox4x32_10((uint32_t)fid, (uint32_t)(fid >> 32), RNG_BITS, (uint32_t)(lane + 64 * q), (uint32_t)a.seed,
                                  (uint32_t)(a.seed >> 32), r);
                    blk[q][0] = r[0] & m4;
                    blk[q][1] = r[1] & m4;
                }
#pragma unroll
                for (int i = 0; i < 32; ++i) {  // slot lane + 64 i: block i & 3, byte jp = i >> 2 of its two words
                    const int jp = i >> 2;
                    w[i >> 2] |= ((blk[i & 3][jp >> 2] >> (8 * (jp & 3))) & 0xffu) << (8 * (i & 3));
                }
            }
        };
        load_tx32(f, txw);
        if (mode & 2) load_tx32(f ^ 1, otw);
        {   // every thread reads back only what it wrote: no barrier needed
            const float2* __restrict__ rxe = a.rx_eq_in + (size_t)f * N_SC + lane;
#pragma unroll
            for (int i = 0; i < 32; ++i) rx_s[64 * i] = rxe[64 * i];
        }
        const float2* __restrict__ ai = sd.agc_inv + f * sd.agc_fs + lane;
        const float* __restrict__ ss = sd.sscale + f * sd.sscale_fs + lane;
        const float p0 = sc.pa_p0_loop;
        const float sqrt_p0 = sqrtf(p0);
        const float pa_out_scale = !T64_PA_UNIT || PA == PA_NONE ? 1.0f : PA == PA_TOI ? 1.0f / sqrt_p0 : sqrt_p0;
        const float pa_in_scale = !T64_PA_UNIT || PA == PA_NONE ? 1.0f : PA == PA_TOI ? sqrt_p0 : 1.0f / sqrt_p0;
        const size_t cnt0 = (a.fpp ? 64u * ((a.frame0 + (unsigned)f) / (unsigned)a.fpp) : 0u) + ((mode & 1) ? (f & 1) * 32 : 0);
        const float g_oth = (mode & 2) ? (uniform ? uni_ss : sd.sscale[f * sd.sscale_fs]) * pa_in_scale : 0.f;

        for (int pass = 0; pass <= a.n_iter; ++pass) {
            pf2 x[64];
            if (pass > 0) {  // ---- regenerate: symbols -> transform -> PA -> transform (frame_kernel_t64.cuh, unit channel) ----
                sy.x_sync();  // the symbols written after the previous detection are visible
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const int n1 = i < 16 ? i : i + 32;
                    const int src = i == 0 ? mir0 : i == 16 ? mir16 : mir + 64 * (31 - i);
                    x[n1] = pk2(sym_s[src]);
                }
                x[16] = pk(0.f, 0.f);
                if (lane == 0) { x[16] = x[0]; x[0] = pk(0.f, 0.f); }
                auto pa_all = [&]() {
#pragma unroll
                    for (int n1 = 0; n1 < 64; ++n1) {
                        if (n1 <= rev64(n1)) {
                            const pf2 u0 = pa_apply_t64<PA>(x[rev64(n1)], p0, sqrt_p0, a.pa_p1);
                            const pf2 u1 = pa_apply_t64<PA>(x[n1], p0, sqrt_p0, a.pa_p1);
                            x[n1] = u0;
                            x[rev64(n1)] = u1;
                        }
                    }
                };
                t64::p1_level_a_pruned(x);
                t64::Raw3 ring[T64_DEPTH];
#pragma unroll 1
                for (int half = 0;; ++half) {
                    t64::p1_levels_bc(x);
                    sy.x_sync();  // S1: nobody still reads the buffer (symbols / previous exchange)
                    ring_acquire();
                    t64::exchange_store(x, buf, lane);
                    sy.x_sync();  // S2: stores visible
                    ring_release();
                    t64::pass2_load(x, buf, lane);
                    t64::Raw3 rb[4];
                    t64::p2_level_a(x, rb, tab, tab_a, lane);
                    if (half == 1) sy.x_sync();  // S3: all loads done, the buffer is free for the next symbols
                    t64::p2_level_b_all<T64_DEPTH>(x, rb, ring, tab, lane);
                    if (half == 1) break;
                    t64::p2_level_c_all<T64_DEPTH, false>(x, ring, tab, lane);
                    pa_all();
                    t64::p1_level_a(x);
                }
                t64::p2_level_c_all<T64_DEPTH, T64_PRUNE != 0>(x, ring, tab, lane);
            }
            // ---- detection `pass` on the group's own slots (corrector.py:74-84, :104-110) ----
            uint32_t detw[8];
#pragma unroll
            for (int w = 0; w < 8; ++w) detw[w] = 0;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                float2 corr = rx_s[64 * i];
                if (pass > 0) {
                    const int n1 = i < 16 ? i : i + 32;
                    pf2 yv = x[rev64(n1)];
                    if (i == 0 && lane == 0) yv = x[rev64(16)];
                    if (T64_PA_UNIT) yv = pmul(yv, pbc(pa_out_scale));
                    const float2 z = cmul(unpk(yv), uniform ? uni_agc : ai[64 * i]);
                    const float2 sy0 = qam_symbol(a.qam, (curw[i >> 2] >> (8 * (i & 3))) & 0xffu);
                    const float2 d = csub(z, sy0);  // distortion estimate, corrector.py:110
                    corr = csub(corr, d);           // corrector.py:74
                }
                if (a.corr_out) a.corr_out[((size_t)f * (a.n_iter + 1) + pass) * N_SC + lane + 64 * i] = corr;
                detw[i >> 2] |= qam_detect(a.qam, corr) << (8 * (i & 3));
            }
            unsigned int nerr = 0;
#pragma unroll
            for (int w = 0; w < 8; ++w) {
                nerr += __popc(detw[w] ^ txw[w]);
                curw[w] = detw[w];
            }
            nerr = __reduce_add_sync(0xffffffffu, nerr);
            if ((t & 31) == 0) {
                if (a.err_total) atomicAdd(&a.err_total[pass + cnt0], (unsigned long long)nerr);
                if (a.err_out) atomicAdd(&a.err_out[(size_t)f * (a.n_iter + 1) + pass], nerr);  // zeroed by the host
            }
            if (a.det_out) {
                uint8_t* dp = a.det_out + ((size_t)f * (a.n_iter + 1) + pass) * N_SC + lane;
#pragma unroll
                for (int i = 0; i < 32; ++i) dp[64 * i] = (uint8_t)((detw[i >> 2] >> (8 * (i & 3))) & 0xffu);
            }
            if (pass < a.n_iter) {  // scaled symbols of the next regeneration (modulation.py:13-25 on the detected values)
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float2 s = qam_symbol(a.qam, (detw[i >> 2] >> (8 * (i & 3))) & 0xffu);
                    const float gsc = (uniform ? uni_ss : ss[64 * i]) * pa_in_scale;
                    float2 xs = make_float2(s.x * gsc, s.y * gsc);
                    if (mode & 2) {  // frame kernel, two users on an explicit-precoder side: v * s, then + v_other * s_other
                        const int p = lane + 64 * i;
                        const float2 v = uniform ? uni_v : sd.v[f * sd.v_fs + p];
                        const float2 vo = uniform ? uni_v : sd.v[(f ^ 1) * sd.v_fs + p];
                        const float2 o = qam_symbol(a.qam, (otw[i >> 2] >> (8 * (i & 3))) & 0xffu);
                        xs = unpk(pcfma(pk2(vo), pk(o.x * g_oth, o.y * g_oth), pcmul(pk2(v), xs.x, xs.y)));
                    }
                    sym_s[lane + 64 * i] = xs;
                }
            }
        }
    }
}

}  // namespace mmimo
