// Fused Monte-Carlo frame kernel, scaled-tangent radix-64 form: the frame semantics and the work decomposition of
// k_link_frame_r64 (frame_kernel_r64.cuh; mp_model.py:208-222 with corrector.py:165-207 / :52-112 inlined) on the
// transform of fft4096_t64.cuh, which needs 907 instead of 1208 packed FMA-pipe instructions per transform.
//   * Both transforms of an antenna chain are FORWARD transforms and nothing is conjugated:
//       IFFT(x)[n] = FFT(x)[-n] / N, so with the precoded symbols written MIRRORED (bin b at input position -b mod N)
//       the first transform yields the time signal in natural order, the PA acts on it sample by sample and the
//       second transform returns the spectrum in natural bin order: FFT(PA(FFT(mirror(x)))) = FFT(PA(N IFFT(x))).
//     The mirror is free: the scaled symbols and the precoder row are read from shared memory at the mirrored slot
//     (lane -> 64 - lane, register i -> 31 - i; lane 0 keeps its own column), conflict-free.
//   * One CTA of 256 threads per SM, four 64-thread groups, each group runs WHOLE antenna chains
//     (precode -> FFT -> PA -> FFT -> MAC) with 64 points per thread and ONE shared-memory exchange per transform;
//     channel rows arrive by cp.async.bulk (TMA, mbarrier completion) in the group's idle exchange buffer; the partial
//     sums of the four groups meet in the exchange buffers in a fixed order (bit-reproducible results).
// Thread (group g, lane n0) owns slots lane + 64 i, i = 0..31 <-> transform register n1 = i (i < 16) or i + 32
// <-> bin lane + 64 n1; slot 0 carries bin 1024 (register 16 of lane 0), bin 0 is unused (link_common.cuh).
#pragma once
#include "fft4096_t64.cuh"
#include "frame_kernel_r64.cuh"

namespace mmimo {

#ifndef T64_DEPTH
#define T64_DEPTH 4  // level-C table entries fetched this many butterfly classes ahead of their use
#endif
#ifndef T64_ACC_SMEM
#define T64_ACC_SMEM 11  // of the 32 per-thread accumulators, how many live in (thread-private) shared memory
#endif
#ifndef T64_PRUNE
#define T64_PRUNE 1  // guard-band pruning of the first transform's level A and of the second transform's last level
#endif
#ifndef T64_LOOP
#define T64_LOOP 1  // 1: guard-band form of level A of the first transform ahead of the loop; 2: loop starts with level A
#endif
#ifndef T64_SYNC
#define T64_SYNC 1  // group-barrier placement, see the kernel
#endif
#ifndef T64_RING
#define T64_RING 2  // the groups take turns in the shared-memory exchange (token ring over named barriers): 1 stores + loads,
                   // 2 stores only (measured best), 3 one token for the stores and one for the loads; 0 off
#endif
#ifndef T64_EPI_EARLY
#define T64_EPI_EARLY 1  // the AGC row of the pass epilogue is loaded before the antenna groups meet (L2 latency under the barriers)
#endif
#ifndef T64_FIRST_ROW
#define T64_FIRST_ROW 0  // where the copy of a group's FIRST precoder row of a pass starts: 0 after the symbols of the pass are
                         // written (round-2 form); 1 at the very top of the pass (under the symbol prologue); 2 in the group's
                         // last chain of the PREVIOUS pass (under the pass epilogue), else at the top of the pass
#endif
#ifndef T64_PA_UNIT
#define T64_PA_UNIT 1  // PA evaluated on the signal scaled to a unit threshold (one multiply less per sample)
#endif
constexpr int FRAME_T64_SMEM_BYTES =
    (4096 + R64_GROUPS * t64::XBUF + 2 * N_SC + T64_ACC_SMEM * FFT_THREADS) * 8 + 64 * 4 + 2 * R64_GROUPS * 8;
// two-user instantiation: + the other user's symbols as int8 pairs (2 bytes per sub-carrier)
constexpr int FRAME_T64_SMEM_BYTES_MU = FRAME_T64_SMEM_BYTES + N_SC * 2;
static_assert(FRAME_T64_SMEM_BYTES_MU <= 227 * 1024, "shared-memory budget of one CTA");

// PA on one packed sample.  T64_PA_UNIT: the sample arrives scaled so that the threshold is 1 (see the kernel):
// soft limiter g = rsqrt(max(|x|^2, 1)); Rapp g = (1 + |x|^(2p))^(-1/(2p)); third order g = 1 - |x|^2.
template <int KIND>
__device__ __forceinline__ pf2 pa_apply_t64(pf2 xp, const float p0, const float sqrt_p0, const float p1) {
    if (!T64_PA_UNIT) return pa_apply_pk<KIND>(xp, p0, sqrt_p0, p1);
    if (KIND == PA_NONE) return xp;
    const float2 x = unpk(xp);
    const float p = fmaf(x.x, x.x, x.y * x.y);
    float g;
    if (KIND == PA_SOFT_LIMITER) {
        asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(g) : "f"(fmaxf(p, 1.0f)));  // rsqrt(1) = 1 exactly: no clipping below 1
    } else if (KIND == PA_RAPP) {
        g = __powf(1.0f + __powf(p, p1), -0.5f / p1);
    } else {
        g = 1.0f - p;
    }
    return pmul(xp, pbc(g));
}

// MU (two users, FrameArgs: "Two-user mode"): every array transmit carries own + other user (modulation.py:373); the own
// precoder row arrives by TMA like the single-user one, the other user's row is read from L2 (prefetched a round ahead),
// the other user's (true) symbols sit in shared memory as int8 pairs.  Requires a uniform symbol scale on explicit-precoder
// sides (sscale_fs == 0, as mu_host.inl sets it).
template <int PA, bool MU = false>
__global__ void __launch_bounds__(FFT_THREADS, 1) k_link_frame_t64(const FrameArgs a, const float2* __restrict__ tab_g) {
    extern __shared__ __align__(128) float2 smem[];
    float2* tab = smem;                                         // per-lane twiddle parameters (t64::build_table)
    pf2* bufs = reinterpret_cast<pf2*>(smem + 4096);            // [group][t64::XBUF]
    float2* symn_s = smem + 4096 + R64_GROUPS * t64::XBUF;      // scaled current symbols, native slots
    unsigned int* err_s = reinterpret_cast<unsigned int*>(symn_s + N_SC + T64_ACC_SMEM * FFT_THREADS);
    unsigned long long* mbar_s = reinterpret_cast<unsigned long long*>(err_s + 64);
    float2* rxeq_s = reinterpret_cast<float2*>(mbar_s + 2 * R64_GROUPS);  // equalised received signal of the frame
    char2* symo_c = reinterpret_cast<char2*>(rxeq_s + N_SC);              // MU only: symbols of the other user

    const int t = threadIdx.x, lane = t & 63, g = t >> 6;
    pf2* accx_s = reinterpret_cast<pf2*>(symn_s + N_SC) + t;    // accumulators i < T64_ACC_SMEM at [i * 256]
    pf2* buf = bufs + g * t64::XBUF;
    const float2* rowA = reinterpret_cast<const float2*>(buf);   // channel row of the antenna being accumulated
    const float2* rowB = rowA + N_SC;                            // precoder / CSI row of the group's next antenna
    const unsigned mbA = smem_u32(mbar_s + 2 * g), mbB = mbA + 8;  // completion of the rowA / rowB copies
    const SyncGroup sy{g + 1};
    for (int i = t; i < t64::TAB_FLOAT2; i += FFT_THREADS) tab[i] = tab_g[i];
    if (t < 64) err_s[t] = 0;
    if (t < 2 * R64_GROUPS) mbar_init(smem_u32(mbar_s + t), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    unsigned phA = 0, phB = 0;  // parities of the mbarrier phases to wait for next
#ifndef T64_RING_ORDER
#define T64_RING_ORDER 0  // token order 0: 0,1,2,3 (measured best with the store-only token) ; 1: 0,2,1,3 ; 2: 0,1,3,2
#endif
    // successor of group g in the ring (one nibble per group)
    const int ring_next = ((T64_RING_ORDER == 1 ? 0x0132 : T64_RING_ORDER == 2 ? 0x2031 : 0x0321) >> (4 * g)) & 3;
    if (T64_RING && ring_next == 0) {  // the token(s) start with group 0
        asm volatile("bar.arrive %0, 128;" ::"r"(5) : "memory");
        if (T64_RING == 3) asm volatile("bar.arrive %0, 128;" ::"r"(9) : "memory");
    }
    const t64::TabA tab_a = t64::load_tab_a(tab, lane);
    // mirrored source slot of register i: (64 - lane) + 64 (31 - i); lane 0 keeps its column, where slot 0 holds
    // bin 1024: register 16 (i = 0, moved below) <- bin 3072 = slot 1024, register 48 (i = 16) <- bin 1024 = slot 0
    const int mir = 64 - lane;
    const int mir0 = lane == 0 ? 1024 : mir + 64 * 31;
    const int mir16 = lane == 0 ? 0 : mir + 64 * 15;

    for (int f = blockIdx.x; f < a.n_frames; f += gridDim.x) {
        const FrameScalars sc = a.scal[f * a.scal_fs];
        uint32_t txw[2];
        auto load_tx = [&](const int inst, uint32_t (&w)[2]) {  // 8 symbol values of this thread for instance inst
            if (a.tx_idx) {
                const uint8_t* p = a.tx_idx + (size_t)inst * N_SC;
                w[0] = w[1] = 0;
#pragma unroll
                for (int jp = 0; jp < 8; ++jp) w[jp >> 2] |= (uint32_t)p[jp * 256 + t] << (8 * (jp & 3));
            } else {
                uint32_t r[4];
                const uint64_t fid = a.frame_id0 + inst;
                philox4x32_10((uint32_t)fid, (uint32_t)(fid >> 32), RNG_BITS, (uint32_t)t, (uint32_t)a.seed,
                              (uint32_t)(a.seed >> 32), r);
                const uint32_t m = (1u << (2 * a.qam.log2side)) - 1u;
                const uint32_t m4 = m | (m << 8) | (m << 16) | (m << 24);
                w[0] = r[0] & m4;
                w[1] = r[1] & m4;
            }
        };
        load_tx(f, txw);
        if (MU) {  // the other user's true symbols (instance f ^ 1), constant over the frame; visible after the pass barrier
            uint32_t otw[2];
            load_tx(f ^ 1, otw);
#pragma unroll
            for (int jp = 0; jp < 8; ++jp) {
                const float2 o = qam_symbol(a.qam, (otw[jp >> 2] >> (8 * (jp & 3))) & 0xffu);
                symo_c[jp * 256 + t] = make_char2((signed char)o.x, (signed char)o.y);
            }
        }
        uint32_t curw[2] = {txw[0], txw[1]};

        for (int pass = 0; pass <= a.n_iter; ++pass) {
            const bool first = (pass == 0);
            const TxSide& sd = first ? a.tx : a.loop;
            const int K = (first && a.rx_eq_in) ? 0 : sd.k;
            const float2* __restrict__ Hm = sd.h_mac + f * sd.h_mac_fs;
            const float2* __restrict__ Hp = sd.v ? sd.v + f * sd.v_fs : sd.h_pre + f * sd.h_pre_fs;
            const bool explicit_v = sd.v != nullptr;
            const bool two_usr = MU && explicit_v && !(a.loop_single_user && !first);
            const float2* __restrict__ Ho = two_usr ? sd.v + (f ^ 1) * sd.v_fs : nullptr;  // precoder rows of the other user
            const float p0 = first ? sc.pa_p0_tx : sc.pa_p0_loop;
            const float sqrt_p0 = sqrtf(p0);
            // T64_PA_UNIT: the chain runs on x / sqrt(P) (P = saturation power; third order: 1 / cubic coefficient), so the
            // PA compares against 1 and needs no per-sample multiply by sqrt(P); the factor returns after the antenna sum
            const float pa_out_scale = !T64_PA_UNIT || PA == PA_NONE ? 1.0f : PA == PA_TOI ? 1.0f / sqrt_p0 : sqrt_p0;
            const float pa_in_scale = !T64_PA_UNIT || PA == PA_NONE ? 1.0f : PA == PA_TOI ? sqrt_p0 : 1.0f / sqrt_p0;
            if (T64_FIRST_ROW && g < K && lane == 0) {
                // T64_FIRST_ROW 2: was the row already requested by this group's last chain of the previous pass?  (it was if
                // that pass existed in this CTA and gave the group a chain; all quantities are uniform kernel arguments)
                bool ahead = false;
                if (T64_FIRST_ROW == 2 && !(first && f == (int)blockIdx.x)) {
                    const int kprev = pass >= 2 || (first && a.n_iter > 0) ? a.loop.k : (a.rx_eq_in ? 0 : a.tx.k);
                    ahead = g < kprev;
                }
                if (!ahead) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    mbar_expect_tx(mbB, ROW_BYTES);
                    bulk_g2s(smem_u32(rowB), Hp + (size_t)g * N_SC, ROW_BYTES, mbB);
                }
            }
            {
                const float* __restrict__ ss = sd.sscale + f * sd.sscale_fs;
#pragma unroll
                for (int jp = 0; jp < 8; ++jp) {
                    const int p = jp * 256 + t;
                    const float2 s = qam_symbol(a.qam, (curw[jp >> 2] >> (8 * (jp & 3))) & 0xffu);
                    const float gsc = ss[p] * pa_in_scale;
                    symn_s[p] = make_float2(s.x * gsc, s.y * gsc);
                }
            }
            const float g_oth = two_usr ? sd.sscale[f * sd.sscale_fs] * pa_in_scale : 0.f;  // uniform scale (see above)
            __syncthreads();  // symbols visible to all groups; every exchange buffer is idle from here
            // what the pass epilogue reads from global memory: on its way to L2 while the antennas are processed
            if (t < 128) prefetch_l2(sd.agc_inv + f * sd.agc_fs + t * 16);

            if (g < K) {  // precoder row of the group's first antenna
                if (lane == 0) {
                    if (!T64_FIRST_ROW) {
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        mbar_expect_tx(mbB, ROW_BYTES);
                        bulk_g2s(smem_u32(rowB), Hp + (size_t)g * N_SC, ROW_BYTES, mbB);
                    }
                    bulk_prefetch_l2(Hm + (size_t)g * N_SC, ROW_BYTES);
                    if (g + R64_GROUPS < K) bulk_prefetch_l2(Hp + (size_t)(g + R64_GROUPS) * N_SC, ROW_BYTES);
                    if (MU && two_usr) {
                        bulk_prefetch_l2(Ho + (size_t)g * N_SC, ROW_BYTES);
                        if (g + R64_GROUPS < K) bulk_prefetch_l2(Ho + (size_t)(g + R64_GROUPS) * N_SC, ROW_BYTES);
                    }
                }
            }
            if (a.stagger > 0 && g >= 2 && K > 2) {  // de-phase the two warps of every scheduler (FrameArgs::stagger)
                const long long t0 = clock64();
                while (clock64() - t0 < a.stagger) {}
            }
            pf2 acc[32];  // partial sum over this group's antennas (entries < T64_ACC_SMEM unused: in accx_s)
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                acc[i] = pk(0.f, 0.f);
                if (i < T64_ACC_SMEM) accx_s[i * FFT_THREADS] = pk(0.f, 0.f);
            }
            // T64_RING: the exchange phases (64 STS.64 + 64 LDS.64 per thread, no FMA work) of the four groups are serialised
            // in the fixed order g = 0, 1, 2, 3, 0, ... by a token that travels over named barriers 5 + g (128 = the 64
            // waiting threads of group g + the 64 arriving threads of group g - 1).  Started together, all eight warps
            // would hit the shared-memory pipe at once and leave the FMA pipe idle for that long; in the ring, one group
            // exchanges while the others compute.  Every group passes the token exactly 2 * ceil(K / 4) times per pass.
            auto ring_acquire = [&]() { asm volatile("bar.sync %0, 128;" ::"r"(5 + g) : "memory"); };
            auto ring_acquire2 = [&]() { asm volatile("bar.sync %0, 128;" ::"r"(9 + g) : "memory"); };
            auto ring_release2 = [&]() { asm volatile("bar.arrive %0, 128;" ::"r"(9 + ring_next) : "memory"); };
            auto ring_release = [&]() { asm volatile("bar.arrive %0, 128;" ::"r"(5 + ring_next) : "memory"); };
            for (int k = g; k < K; k += R64_GROUPS) {
                pf2 x[64];
                mbar_wait_parity(mbB, phB);
                phB ^= 1u;
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const int n1 = i < 16 ? i : i + 32;
                    const int src = i == 0 ? mir0 : i == 16 ? mir16 : mir + 64 * (31 - i);
                    const float2 s = symn_s[src];
                    const pf2 h = pk2(rowB[src]);
                    if (explicit_v) x[n1] = pcmul(h, s.x, s.y);                                 // v * s
                    else x[n1] = pfma(pswap(h), pbc(s.y), pmul(h, pk(s.x, -s.x)));               // conj(h) * s
                    if (MU && two_usr) {  // modulation.py:373: sum over users of symbol * precoder
                        const char2 o = symo_c[src];
                        x[n1] = pcfma(pk2(__ldg(Ho + (size_t)k * N_SC + src)), pk((float)o.x * g_oth, (float)o.y * g_oth), x[n1]);
                    }
                }
                if (T64_LOOP != 1) {
#pragma unroll
                    for (int n1 = 16; n1 < 48; ++n1) x[n1] = pk(0.f, 0.f);
                } else {
                    x[16] = pk(0.f, 0.f);
                }
                if (lane == 0) { x[16] = x[0]; x[0] = pk(0.f, 0.f); }
                // ONE loop body serves both transforms of the chain (two specialised bodies do not fit the instruction
                // cache next to each other, profiles/r01_r64_frame_kernel.txt); what differs between the transforms sits
                // outside the shared part: the guard-band form of level A of pass 1 (first transform, T64_LOOP 1), the
                // PA with the digit reversal folded into its destination registers, and the pruned last level of pass 2
                // (second transform), which follows the loop so that no register shuffle is needed on the way out.
                // Group barriers: S1 = everybody has finished READING the buffer (rows of the precode / previous
                // exchange) before it is overwritten; S2 = all stores visible; S3 (second transform only) = all loads done
                // before the TMA rows land in the buffer.  With T64_SYNC 1, S1 and S3 sit after long compute phases where
                // nobody waits, and the exchange loads overlap the first butterflies of pass 2.
                auto issue_rows = [&]() {
                    if (lane == 0) {  // the exchange buffer is idle from here to the next chain's stores
                        const bool more = k + R64_GROUPS < K;
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        mbar_expect_tx(mbA, ROW_BYTES);
                        bulk_g2s(smem_u32(rowA), Hm + (size_t)k * N_SC, ROW_BYTES, mbA);
                        if (more) {
                            mbar_expect_tx(mbB, ROW_BYTES);
                            bulk_g2s(smem_u32(rowB), Hp + (size_t)(k + R64_GROUPS) * N_SC, ROW_BYTES, mbB);
                            if (Hp != Hm) bulk_prefetch_l2(Hm + (size_t)(k + R64_GROUPS) * N_SC, ROW_BYTES);
                            if (k + 2 * R64_GROUPS < K) {
                                bulk_prefetch_l2(Hp + (size_t)(k + 2 * R64_GROUPS) * N_SC, ROW_BYTES);
                                if (MU && two_usr) bulk_prefetch_l2(Ho + (size_t)(k + 2 * R64_GROUPS) * N_SC, ROW_BYTES);
                            }
                        } else if (T64_FIRST_ROW == 2) {  // the group's last chain of this pass: first row of its next pass
                            const bool last = pass == a.n_iter;
                            const int nf = last ? f + (int)gridDim.x : f;
                            const TxSide& nsd = last ? a.tx : a.loop;
                            if (nf < a.n_frames && !(last && a.rx_eq_in) && g < nsd.k) {
                                const float2* __restrict__ Hn = nsd.v ? nsd.v + nf * nsd.v_fs : nsd.h_pre + nf * nsd.h_pre_fs;
                                mbar_expect_tx(mbB, ROW_BYTES);
                                bulk_g2s(smem_u32(rowB), Hn + (size_t)g * N_SC, ROW_BYTES, mbB);
                            }
                        }
                    }
                };
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
                if (!T64_SYNC) sy.x_sync();  // S1 of the first transform
                if (T64_LOOP == 1) t64::p1_level_a_pruned(x);
                t64::Raw3 ring[T64_DEPTH];
#pragma unroll 1
                for (int half = 0;; ++half) {
                    if (T64_LOOP != 1) t64::p1_level_a(x);
                    t64::p1_levels_bc(x);
                    if (T64_SYNC) sy.x_sync();  // S1
                    if (T64_RING) ring_acquire();
                    t64::exchange_store(x, buf, lane);
                    sy.x_sync();                // S2
                    if (T64_RING >= 2) ring_release();   // token covers the stores only (passing it on earlier, after 32 or 48
                                                         // of the 64 stores, measured slower)
                    if (T64_RING == 3) ring_acquire2();  // a second token for the loads
                    t64::pass2_load(x, buf, lane);
                    if (T64_RING == 1) ring_release();
                    if (T64_RING == 3) ring_release2();
                    t64::Raw3 rb[4];
                    if (!T64_SYNC) {
                        sy.x_sync();            // S3 (both transforms)
                        if (half == 1) issue_rows();
                        t64::p2_level_a(x, rb, tab, tab_a, lane);
                    } else {
                        t64::p2_level_a(x, rb, tab, tab_a, lane);
                        if (half == 1) {
                            sy.x_sync();        // S3
                            issue_rows();
                        }
                    }
                    t64::p2_level_b_all<T64_DEPTH>(x, rb, ring, tab, lane);
                    if (half == 1) break;
                    t64::p2_level_c_all<T64_DEPTH, false>(x, ring, tab, lane);
                    pa_all();
                    if (T64_LOOP == 1) t64::p1_level_a(x);
                }
                t64::p2_level_c_all<T64_DEPTH, T64_PRUNE != 0>(x, ring, tab, lane);
                pf2 (&y)[64] = x;
                mbar_wait_parity(mbA, phA);
                phA ^= 1u;
                {   // all loads before the first store: the compiler cannot prove rowA / accx_s disjoint
                    pf2 hv[32], av[T64_ACC_SMEM > 0 ? T64_ACC_SMEM : 1];
#pragma unroll
                    for (int i = 0; i < 32; ++i) hv[i] = pk2(rowA[lane + 64 * i]);
#pragma unroll
                    for (int i = 0; i < T64_ACC_SMEM; ++i) av[i] = accx_s[i * FFT_THREADS];
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const int n1 = i < 16 ? i : i + 32;
                        pf2 yv = y[rev64(n1)];
                        if (i == 0 && lane == 0) yv = y[rev64(16)];
                        if (i < T64_ACC_SMEM) av[i] = pcfma(yv, hv[i], av[i]);
                        else acc[i] = pcfma(yv, hv[i], acc[i]);
                    }
#pragma unroll
                    for (int i = 0; i < T64_ACC_SMEM; ++i) accx_s[i * FFT_THREADS] = av[i];
                }
            }
            const float2* __restrict__ ai = sd.agc_inv + f * sd.agc_fs;
            float2 aiv[8];
            if (T64_EPI_EARLY) {
#pragma unroll
                for (int jp = 0; jp < 8; ++jp) aiv[jp] = ai[jp * 256 + t];
            }
            if (T64_RING) {  // groups with one chain less than the others still pass the token on
                const int mine = K > g ? (K - g + R64_GROUPS - 1) / R64_GROUPS : 0, slots = (K + R64_GROUPS - 1) / R64_GROUPS;
                for (int e = 2 * mine; e < 2 * slots; ++e) {
                    ring_acquire(); ring_release();
                    if (T64_RING == 3) { ring_acquire2(); ring_release2(); }
                }
            }
            // ---- the four groups' partial sums meet in the (idle) exchange buffers: fixed summation order ----
            sy.x_sync();  // the group's last MAC has read its channel row
#pragma unroll
            for (int i = 0; i < 32; ++i) buf[lane + 64 * i] = i < T64_ACC_SMEM ? accx_s[i * FFT_THREADS] : acc[i];
            __syncthreads();
            pf2 accn[8];
#pragma unroll
            for (int jp = 0; jp < 8; ++jp) {
                const int p = jp * 256 + t;
                accn[jp] = padd(padd(bufs[p], bufs[t64::XBUF + p]), padd(bufs[2 * t64::XBUF + p], bufs[3 * t64::XBUF + p]));
                if (T64_PA_UNIT) accn[jp] = pmul(accn[jp], pbc(pa_out_scale));
            }

            if (!MU && a.sdr_desired) {  // ---- signal-to-distortion snapshot instead of the receiver (FrameArgs) ----
                double pd = 0.0, pi = 0.0;
#pragma unroll
                for (int jp = 0; jp < 8; ++jp) {
                    const int p = jp * 256 + t;
                    const float2 ac = unpk(accn[jp]);            // 64 x the noise-free received signal
                    const float2 iv = T64_EPI_EARLY ? aiv[jp] : ai[p];                   // 1 / (64 agc)
                    const float m = 1.0f / (64.0f * (iv.x * iv.x + iv.y * iv.y));
                    const float2 agc = make_float2(iv.x * m, -iv.y * m);
                    const float2 des = cmul(qam_symbol(a.qam, (curw[jp >> 2] >> (8 * (jp & 3))) & 0xffu), agc);
                    const float2 dis = make_float2(fmaf(ac.x, 0.015625f, -des.x), fmaf(ac.y, 0.015625f, -des.y));
                    pd += (double)des.x * des.x + (double)des.y * des.y;
                    pi += (double)dis.x * dis.x + (double)dis.y * dis.y;
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    pd += __shfl_xor_sync(0xffffffffu, pd, o);
                    pi += __shfl_xor_sync(0xffffffffu, pi, o);
                }
                double* red_s = reinterpret_cast<double*>(rxeq_s);
                if ((t & 31) == 0) { red_s[t >> 5] = pd; red_s[8 + (t >> 5)] = pi; }
                __syncthreads();
                if (t == 0) {
                    double sd2 = 0.0, si2 = 0.0;
                    for (int w = 0; w < 8; ++w) { sd2 += red_s[w]; si2 += red_s[8 + w]; }
                    a.sdr_desired[f] = sd2;
                    a.sdr_distortion[f] = si2;
                }
                continue;
            }
            // ---- end of pass: equalise / estimate distortion / slice / count (as k_link_frame) -----------
            uint32_t detw[2] = {0u, 0u};
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
#pragma unroll
            for (int jp = 0; jp < 8; ++jp) {
                const int p = jp * 256 + t;
                const float2 ac = unpk(accn[jp]);
                float2 corr;
                if (first) {
                    if (a.rx_eq_in) {
                        corr = a.rx_eq_in[(size_t)f * N_SC + p];
                    } else {
                        const float2 r = make_float2(fmaf(64.f, nz[jp].x, ac.x), fmaf(64.f, nz[jp].y, ac.y));
                        corr = cmul(r, T64_EPI_EARLY ? aiv[jp] : ai[p]);
                    }
                    rxeq_s[p] = corr;
                    if (a.rx_eq_out) a.rx_eq_out[(size_t)f * N_SC + p] = corr;
                } else {
                    const float2 z = cmul(ac, T64_EPI_EARLY ? aiv[jp] : ai[p]);
                    const float2 sy0 = qam_symbol(a.qam, (curw[jp >> 2] >> (8 * (jp & 3))) & 0xffu);
                    const float2 d = csub(z, sy0);   // distortion estimate, corrector.py:205
                    corr = csub(rxeq_s[p], d);       // corrector.py:184
                }
                if (a.corr_out) a.corr_out[((size_t)f * (a.n_iter + 1) + pass) * N_SC + p] = corr;
                const uint32_t idx = qam_detect(a.qam, corr);
                detw[jp >> 2] |= idx << (8 * (jp & 3));
            }
            curw[0] = detw[0];
            curw[1] = detw[1];
            unsigned int nerr = __popc(detw[0] ^ txw[0]) + __popc(detw[1] ^ txw[1]);
            nerr = __reduce_add_sync(0xffffffffu, nerr);
            if ((t & 31) == 0) atomicAdd(&err_s[pass], nerr);
            if (a.det_out) {
                uint8_t* dp = a.det_out + ((size_t)f * (a.n_iter + 1) + pass) * N_SC;
#pragma unroll
                for (int jp = 0; jp < 8; ++jp) dp[jp * 256 + t] = (uint8_t)((detw[jp >> 2] >> (8 * (jp & 3))) & 0xffu);
            }
        }
        __syncthreads();
        if (t <= a.n_iter && a.err_total) {
            const unsigned int e = err_s[t];
            err_s[t] = 0;
            if (a.err_out) a.err_out[(size_t)f * (a.n_iter + 1) + t] = e;
            atomicAdd(&a.err_total[t + (MU ? (f & 1) * 32 : 0) + (a.fpp ? 64u * ((a.frame0 + (unsigned)f) / (unsigned)a.fpp) : 0u)], (unsigned long long)e);  // MU: user 1 counts at [32..]
        }
        __syncthreads();
    }
}

}  // namespace mmimo
