// Host orchestration of Link.simulate (mp_model.py:89-222) - included by mmimo_b200.cu.

static double calc_alpha_host(double ibo_db) {  // modulation.py:178-189
    const double g = pow(10.0, ibo_db / 20.0);
    return 1.0 - exp(-g * g) + (sqrt(M_PI) * g / 2.0) * erfc(g);
}

template <int PA>
static void launch_frame(const FrameArgs& fa, const mmimo_ctx* c, int n) {
    // Default: the scaled-tangent radix-64 kernel (frame_kernel_t64.cuh).  MMIMO_FRAME_KERNEL=r64 / r16 (read once at
    // mmimo_create) select the round-1 radix-64 kernel or the three-pass radix-16 kernel: measurement knobs, same
    // results to rounding.
    if (c->frame_kernel == 0) {
        k_link_frame_t64<PA><<<std::min(n, c->n_sm), FFT_THREADS, FRAME_T64_SMEM_BYTES, c->stream>>>(fa, c->tw_t64.as<float2>());
        return;
    }
    if (c->frame_kernel == 1) {
        k_link_frame_r64<PA><<<std::min(n, c->n_sm), FFT_THREADS, FRAME64_SMEM_BYTES, c->stream>>>(fa, c->tw4096.as<float2>());
        return;
    }
    const int grid = std::min(n, c->ctas_per_sm_r16 * c->n_sm);
    k_link_frame<PA><<<grid, FFT_THREADS, FRAME_SMEM_BYTES, c->stream>>>(fa, c->tw4096.as<float2>(), c->tw256.as<float2>());
}

// upload a host array of `rows` x 2048 elements in data-sub-carrier order into `dst` (native slots)
template <class T>
static int upload_native(mmimo_ctx* c, const T* host, T* dst, size_t rows) {
    const size_t bytes = rows * N_SC * sizeof(T);
    if (c->stage.reserve(bytes)) return 1;
    CK(cudaMemcpyAsync(c->stage.p, host, bytes, cudaMemcpyHostToDevice, c->stream));
    k_to_native<T><<<nblk(rows * N_SC), 256, 0, c->stream>>>(c->stage.as<T>(), dst, rows);
    c->launches++;
    return 0;
}
template <class T>
static int download_native(mmimo_ctx* c, const T* src, T* host, size_t rows) {
    const size_t bytes = rows * N_SC * sizeof(T);
    if (c->stage.reserve(bytes)) return 1;
    k_from_native<T><<<nblk(rows * N_SC), 256, 0, c->stream>>>(src, c->stage.as<T>(), rows);
    c->launches++;
    CK(cudaMemcpyAsync(host, c->stage.p, bytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));  // stage is reused by the next transfer
    return 0;
}

struct RealisationIn {
    const double* rx_pos_dev;  // positions already on the device (staging slot of the copy stream) or nullptr
    const double* rx_pos;
    const float* h_true;
    const float* h_loop;
    const float* csi_noise;
    uint32_t pos_stream, ray_stream, csi_stream;
};

// channel -> (CSI error) -> MRT precoder + AGC for frames [c0, c0+n) of the request
static int run_realisation(mmimo_ctx* c, const mmimo_sim_request* rq, const RealisationIn& in, size_t c0, int n,
                           float2** h_true_out, float2** h_loop_out, bool want_v) {
    const mmimo_link_config& cfg = c->cfg;
    const int K = cfg.n_ant;
    const size_t hN = (size_t)n * K * N_SC;
    float2* h_true = c->h_true.as<float2>();
    float2* h_loop = h_true;
    int sscale_ready = 0;
    if (in.h_true) {
        if (upload_native<float2>(c, reinterpret_cast<const float2*>(in.h_true) + c0 * K * N_SC, h_true, (size_t)n * K)) return 1;
    } else if (cfg.channel_kind == MMIMO_CHAN_LOS || cfg.channel_kind == MMIMO_CHAN_TWO_PATH) {
        ChanArgs ca{};
        ca.tx_pos = c->tx_pos.as<double>();
        ca.rx_pos_in = nullptr;
        if (in.rx_pos_dev) {
            CK(cudaMemcpyAsync(c->rx_pos.p, in.rx_pos_dev, (size_t)n * 24, cudaMemcpyDeviceToDevice, c->stream));
            ca.rx_pos_in = c->rx_pos.as<double>();
        } else if (in.rx_pos) {
            CK(cudaMemcpyAsync(c->rx_pos.p, in.rx_pos + c0 * 3, (size_t)n * 24, cudaMemcpyHostToDevice, c->stream));
            ca.rx_pos_in = c->rx_pos.as<double>();
        }
        ca.rx_pos_out = in.rx_pos ? nullptr : c->rx_pos.as<double>();
        ca.freq_slot = c->freq_slot.as<double>();
        ca.rx0[0] = cfg.rx_pos[0]; ca.rx0[1] = cfg.rx_pos[1]; ca.rx0[2] = cfg.rx_pos[2];
        ca.loc_var = cfg.rx_loc_var;
        ca.gain_lin = sqrt(pow(10.0, cfg.ant_gain_db / 10.0));
        ca.kind = cfg.channel_kind == MMIMO_CHAN_TWO_PATH ? 1 : 0;
        ca.skip_attenuation = cfg.skip_attenuation;
        ca.seed = rq->seed; ca.frame_id0 = rq->frame_id0 + c0;
        ca.pos_stream = in.pos_stream;
        ca.h_out = h_true; ca.n_ant = K;
        // MRT from this very channel (no CSI error, no separate channel knowledge) and enough realisations to fill the
        // GPU with one CTA each: the channel kernel also writes the symbol scale, the realisation kernel skips a pass
        const bool los_closed_form = c->prep_analytic && cfg.channel_kind == MMIMO_CHAN_LOS && !cfg.skip_attenuation && !want_v;
        sscale_ready = !in.h_loop && !(rq->csi_eps > 0.0) && n >= 2 * c->n_sm && !los_closed_form;
        ca.sscale_out = sscale_ready ? c->sscale.as<float>() : nullptr;
        prof_begin(c, PROF_CHAN);
        if (c->stream == c->prep_stream && !sscale_ready)  // pipelined: the light form that fits beside a frame-kernel CTA
            k_gen_channel_bg<<<dim3((K + CHAN_ANT_PER_CTA - 1) / CHAN_ANT_PER_CTA, n), CHAN_BG_THREADS, 0, c->stream>>>(ca);
        else
            k_gen_channel<<<dim3(sscale_ready ? 1 : (K + CHAN_ANT_PER_CTA - 1) / CHAN_ANT_PER_CTA, n), 256, 0, c->stream>>>(ca);
        prof_end(c);
        c->launches++;
    } else if (cfg.channel_kind == MMIMO_CHAN_RAYLEIGH) {
        RayleighArgs ra{};
        ra.tx_pos = c->tx_pos.as<double>();
        ra.freq_slot = c->freq_slot.as<double>();
        ra.rx0[0] = cfg.rx_pos[0]; ra.rx0[1] = cfg.rx_pos[1]; ra.rx0[2] = cfg.rx_pos[2];
        ra.gain_lin = sqrt(pow(10.0, cfg.ant_gain_db / 10.0));
        ra.skip_attenuation = cfg.skip_attenuation;
        ra.seed = rq->seed; ra.frame_id0 = rq->frame_id0 + c0; ra.stream = in.ray_stream;
        ra.h_out = h_true; ra.n_ant = K;
        sscale_ready = !in.h_loop && !(rq->csi_eps > 0.0) && n >= 2 * c->n_sm;  // as above: fused norm pass
        ra.sscale_out = sscale_ready ? c->sscale.as<float>() : nullptr;
        prof_begin(c, PROF_CHAN);
        k_gen_rayleigh<<<dim3(sscale_ready ? 1 : (K + CHAN_ANT_PER_CTA - 1) / CHAN_ANT_PER_CTA, n), 256, 0, c->stream>>>(ra);
        prof_end(c);
        c->launches++;
    } else {
        return fail("mmimo_link: channel_kind INJECTED needs h_true in the request");
    }
    if (in.h_loop) {
        if (c->h_loop.reserve(hN * 8)) return 1;
        h_loop = c->h_loop.as<float2>();
        if (upload_native<float2>(c, reinterpret_cast<const float2*>(in.h_loop) + c0 * K * N_SC, h_loop, (size_t)n * K)) return 1;
    } else if (rq->csi_eps > 0.0) {
        if (c->h_loop.reserve(hN * 8)) return 1;
        h_loop = c->h_loop.as<float2>();
        CsiArgs xa{};
        if (in.csi_noise) {
            if (c->csi_nz.reserve(hN * 8)) return 1;
            if (upload_native<float2>(c, reinterpret_cast<const float2*>(in.csi_noise) + c0 * K * N_SC, c->csi_nz.as<float2>(), (size_t)n * K)) return 1;
            xa.noise_in = c->csi_nz.as<float2>();
        }
        xa.h_in = h_true; xa.h_out = h_loop; xa.eps = (float)rq->csi_eps; xa.scale = (float)sqrt(1.0 - rq->csi_eps * rq->csi_eps);
        xa.seed = rq->seed; xa.frame_id0 = rq->frame_id0 + c0; xa.stream = in.csi_stream; xa.n_ant = K;
        // the precoder is derived from the perturbed rows: with one CTA per realisation the kernel also writes their symbol scale
        sscale_ready = n >= 2 * c->n_sm;
        xa.sscale_out = sscale_ready ? c->sscale.as<float>() : nullptr;
        prof_begin(c, PROF_CHAN);
        k_csi_error<<<dim3(sscale_ready ? 1 : K, n), CSI_T, 0, c->stream>>>(xa);
        prof_end(c);
        c->launches++;
    }
    // line of sight + MRT from the true channel: closed form in the K path amplitudes, no pass over the channel rows
    const bool analytic = c->prep_analytic && !in.h_true && cfg.channel_kind == MMIMO_CHAN_LOS && !cfg.skip_attenuation &&
                          h_loop == h_true && !want_v;
    if (analytic) {
        PrepLosArgs la{};
        la.tx_pos = c->tx_pos.as<double>(); la.rx_pos = c->rx_pos.as<double>(); la.freq_slot = c->freq_slot.as<double>();
        la.gain_lin = sqrt(pow(10.0, cfg.ant_gain_db / 10.0));
        la.sscale = c->sscale.as<float>(); la.agc_inv = c->agc_inv.as<float2>(); la.hkvk_inv = c->hkvk_inv.as<float2>();
        la.scal = c->scal.as<FrameScalars>(); la.alpha_out = c->alpha.as<float>();
        la.n_ant = K; la.pa_kind = cfg.pa_kind; la.is_mcnc = cfg.receiver_kind == MMIMO_RX_MCNC;
        la.ibo_lin = pow(10.0, cfg.ibo_db / 10.0);
        la.avg_symbol_power = avg_symbol_power(cfg.constel_size);
        la.avg_sample_power = la.avg_symbol_power * ((double)cfg.n_sub_carr / cfg.n_fft);
        la.snr_lin = pow(10.0, rq->snr_db / 10.0);
        SweepArgs sw{c->sweep.snr_lin, c->sweep.ibo_lin, c->sweep.fpp, (unsigned)c0, 0};
        prof_begin(c, PROF_PREP);
        if (sw.fpp) k_prep_los<true><<<n, 256, 0, c->stream>>>(la, sw);
        else k_prep_los<false><<<n, 256, 0, c->stream>>>(la, sw);
        prof_end(c);
        c->launches++;
        *h_true_out = h_true;
        *h_loop_out = h_loop;
        return 0;
    }
    PrepArgs pa{};
    pa.h_loop = h_loop; pa.h_fs = (long long)K * N_SC;
    pa.v_in = nullptr;
    pa.v_out = want_v ? c->v.as<float2>() : nullptr;
    pa.sscale = c->sscale.as<float>();
    pa.agc_inv = c->agc_inv.as<float2>();
    pa.hkvk_inv = c->hkvk_inv.as<float2>();
    pa.scal = c->scal.as<FrameScalars>();
    pa.alpha_out = c->alpha.as<float>();
    pa.n_ant = K; pa.n_frames = n;
    pa.pa_kind = cfg.pa_kind; pa.is_mcnc = cfg.receiver_kind == MMIMO_RX_MCNC;
    pa.ibo_lin = pow(10.0, cfg.ibo_db / 10.0);
    pa.avg_symbol_power = avg_symbol_power(cfg.constel_size);
    pa.avg_sample_power = pa.avg_symbol_power * ((double)cfg.n_sub_carr / cfg.n_fft);
    pa.snr_lin = pow(10.0, rq->snr_db / 10.0);
    SweepArgs sw{c->sweep.snr_lin, c->sweep.ibo_lin, c->sweep.fpp, (unsigned)c0, sscale_ready};
    prof_begin(c, PROF_PREP);
    if (sw.fpp) k_prep_realisation<true><<<n, PREP_T, 0, c->stream>>>(pa, sw);
    else k_prep_realisation<false><<<n, PREP_T, 0, c->stream>>>(pa, sw);
    prof_end(c);
    c->launches++;
    *h_true_out = h_true;
    *h_loop_out = h_loop;
    return 0;
}

// pipelined realisations: exchange the per-realisation buffers (and their events) with the second set
static void swap_prep_set(mmimo_ctx* c) {
    mmimo_ctx::PrepSet& a = c->alt;
    std::swap(c->h_true, a.h_true); std::swap(c->h_loop, a.h_loop); std::swap(c->csi_nz, a.csi_nz);
    std::swap(c->sscale, a.sscale); std::swap(c->agc_inv, a.agc_inv); std::swap(c->hkvk_inv, a.hkvk_inv);
    std::swap(c->scal, a.scal); std::swap(c->alpha, a.alpha); std::swap(c->rx_pos, a.rx_pos);
    std::swap(c->cur_ready, a.ready); std::swap(c->cur_released, a.released); std::swap(c->cur_used, a.used);
}

static int run_link(mmimo_ctx* c, const mmimo_sim_request* rq, mmimo_sim_result* rs) {
    if (!c || !rq) return fail("mmimo_link: NULL argument");
    if (!c->configured) return fail("mmimo_link: context not configured (call mmimo_configure)");
    if (rq->n_frames <= 0) return 0;
    if (rq->n_iter < 0 || rq->n_iter > 62) return fail("mmimo_link: n_iter must be 0..62 (got %d)", rq->n_iter);
    CK(cudaSetDevice(c->device));
    const mmimo_link_config& cfg = c->cfg;
    const int K = cfg.n_ant, I = rq->n_iter;
    const bool mcnc = cfg.receiver_kind == MMIMO_RX_MCNC;
    const bool sep_loop = rq->h_loop || rq->clean_h_loop || rq->csi_eps > 0.0;
    const bool want_det = rs && rs->det, want_corr = rs && rs->corr, want_rxeq = rs && rs->rx_eq;
    const bool want_v = rs && rs->v;
    // chunking: per-frame scratch bytes
    size_t per_frame = (size_t)K * N_SC * 8 * ((sep_loop ? 2 : 1) + (want_v ? 1 : 0) + ((rq->csi_noise || rq->clean_csi_noise) ? 1 : 0)) + (size_t)N_SC * (8 * 4 + 4 + 1) + 64 + (size_t)K * 4;
    if (want_det) per_frame += (size_t)(I + 1) * N_SC;
    if (want_corr) per_frame += (size_t)(I + 1) * N_SC * 8;
    size_t chunk = std::max<size_t>(1, c->scratch_budget / per_frame);
    chunk = std::min<size_t>(chunk, (size_t)rq->n_frames);
    chunk = std::min<size_t>(chunk, 8192);
    const size_t hB = chunk * K * N_SC * 8;
    if (c->h_true.reserve(hB) || (want_v && c->v.reserve(hB)) || c->sscale.reserve(chunk * N_SC * 4) ||
        c->agc_inv.reserve(chunk * N_SC * 8) ||
        c->hkvk_inv.reserve(chunk * N_SC * 8) || c->scal.reserve(chunk * sizeof(FrameScalars)) ||
        c->alpha.reserve(chunk * K * 4) || c->rx_pos.reserve(chunk * 24) || c->tx_idx.reserve(chunk * N_SC) ||
        c->noise.reserve(chunk * N_SC * 8) || c->err.reserve(chunk * (I + 1) * 4) || c->rx_eq.reserve(chunk * N_SC * 8))
        return 1;
    if (want_det && c->det.reserve(chunk * (I + 1) * N_SC)) return 1;
    if (want_corr && c->corr.reserve(chunk * (I + 1) * N_SC * 8)) return 1;
    double cnc_alpha = 1.0;
    if (!mcnc) {
        if (c->cnc_ones.reserve(N_SC * 8) || c->cnc_v.reserve(N_SC * 4) || c->cnc_agc_inv.reserve(N_SC * 8)) return 1;
        const double alpha = cnc_alpha = (cfg.pa_kind == MMIMO_PA_TOI || cfg.pa_kind == MMIMO_PA_NONE) ? 1.0 : calc_alpha_host(cfg.ibo_db);
        k_fill_c<<<nblk(N_SC), 256, 0, c->stream>>>(c->cnc_ones.as<float2>(), make_float2(1.f, 0.f), N_SC);
        k_fill_f<<<nblk(N_SC), 256, 0, c->stream>>>(c->cnc_v.as<float>(), 1.f / 64.f, N_SC);
        k_fill_c<<<nblk(N_SC), 256, 0, c->stream>>>(c->cnc_agc_inv.as<float2>(), make_float2((float)(1.0 / (64.0 * alpha)), 0.f), N_SC);
        c->launches += 3;
    }
    unsigned long long* counters = c->counters.as<unsigned long long>();
    // Pipelined realisations (asynchronous calls with device-generated channels): see mmimo_ctx::PrepSet
    const bool pipelined = c->pipeline_prep && !rs && !rq->incl_clean_run && !rq->h_true && !rq->h_loop && !rq->csi_noise &&
                           cfg.channel_kind != MMIMO_CHAN_INJECTED;

    for (size_t c0 = 0; c0 < (size_t)rq->n_frames; c0 += chunk) {
        const int n = (int)std::min<size_t>(chunk, (size_t)rq->n_frames - c0);
        float2 *h_true = nullptr, *h_loop = nullptr;
        if (rq->incl_clean_run) {
            RealisationIn in{nullptr, rq->clean_rx_pos, rq->clean_h_true, rq->clean_h_loop, rq->clean_csi_noise, RNG_POS + 16, RNG_RAYLEIGH + 16, RNG_CSI + 16};
            if (run_realisation(c, rq, in, c0, n, &h_true, &h_loop, false)) return 1;
            CleanArgs ca{};
            ca.h_true = h_true; ca.h_fs = (long long)K * N_SC;
            ca.h_pre = h_loop; ca.sscale = c->sscale.as<float>();
            ca.hkvk_inv = c->hkvk_inv.as<float2>();
            ca.scal = c->scal.as<FrameScalars>();
            ca.n_ant = K; ca.n_frames = n;
            ca.qam.log2side = log2_side(cfg.constel_size);
            ca.tx_idx = nullptr; ca.noise = nullptr;
            if (rq->clean_tx_idx) {
                if (upload_native<uint8_t>(c, rq->clean_tx_idx + c0 * N_SC, c->tx_idx.as<uint8_t>(), n)) return 1;
                ca.tx_idx = c->tx_idx.as<uint8_t>();
            }
            if (rq->clean_noise) {
                if (upload_native<float2>(c, reinterpret_cast<const float2*>(rq->clean_noise) + c0 * N_SC, c->noise.as<float2>(), n)) return 1;
                ca.noise = c->noise.as<float2>();
            }
            ca.noise_unit = (rq->flags & MMIMO_REQ_NOISE_UNIT) ? 1 : 0;
            ca.seed = rq->seed; ca.frame_id0 = rq->frame_id0 + c0;
            if (rs && rs->clean_det && c->det.reserve((size_t)chunk * (I + 1) * N_SC)) return 1;
            ca.det_out = (rs && rs->clean_det) ? c->det.as<uint8_t>() : nullptr;
            ca.err_out = (rs && rs->clean_err_per_frame) ? c->err.as<uint32_t>() : nullptr;
            ca.err_total = counters + 64;
            k_clean_frame<<<n, 256, 0, c->stream>>>(ca);
            c->launches++;
            if (rs && rs->clean_det && download_native<uint8_t>(c, c->det.as<uint8_t>(), rs->clean_det + c0 * N_SC, n)) return 1;
            if (rs && rs->clean_err_per_frame) {
                CK(cudaMemcpyAsync(rs->clean_err_per_frame + c0, c->err.p, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
                CK(cudaStreamSynchronize(c->stream));  // c->err is rewritten by the distorted frames of this chunk
            }
        }
        if (rq->incl_clean_run == 2) continue;  // clean frames only
        // per-frame host inputs of this chunk: H2D on the copy stream into a staging slot (overlaps whatever the compute
        // stream is still running, normally the previous batch's frame kernel), permuted into place on the compute stream
        if (pipelined) {  // the buffer set the previous batch's frame kernel is NOT reading
            swap_prep_set(c);
            if (c->h_true.reserve(hB) || c->sscale.reserve(chunk * N_SC * 4) || c->agc_inv.reserve(chunk * N_SC * 8) ||
                c->hkvk_inv.reserve(chunk * N_SC * 8) || c->scal.reserve(chunk * sizeof(FrameScalars)) ||
                c->alpha.reserve(chunk * K * 4) || c->rx_pos.reserve(chunk * 24))
                return 1;
            if (c->cur_used) CK(cudaStreamWaitEvent(c->prep_stream, c->cur_released, 0));
        }
        mmimo_ctx::InSlot* sl = nullptr;
        if (rq->rx_pos || rq->tx_idx || rq->noise) {
            sl = &c->in_slot[c->slot_ctr++ & 1];
            // both slots grow together: a cudaMalloc is a device-wide synchronisation and must not land in a later batch
            for (auto& s2 : c->in_slot)
                if ((rq->rx_pos && s2.pos.reserve((size_t)chunk * 24)) || (rq->tx_idx && s2.idx.reserve((size_t)chunk * N_SC)) ||
                    (rq->noise && s2.nz.reserve((size_t)chunk * N_SC * 8)))
                    return 1;
            if (sl->in_use) CK(cudaStreamWaitEvent(c->copy_stream, sl->consumed, 0));
            if (rq->rx_pos) {
                if (sl->pos.reserve((size_t)chunk * 24)) return 1;
                CK(cudaMemcpyAsync(sl->pos.p, rq->rx_pos + c0 * 3, (size_t)n * 24, cudaMemcpyHostToDevice, c->copy_stream));
            }
            if (rq->tx_idx) {
                if (sl->idx.reserve((size_t)chunk * N_SC)) return 1;
                CK(cudaMemcpyAsync(sl->idx.p, rq->tx_idx + c0 * N_SC, (size_t)n * N_SC, cudaMemcpyHostToDevice, c->copy_stream));
            }
            if (rq->noise) {
                if (sl->nz.reserve((size_t)chunk * N_SC * 8)) return 1;
                CK(cudaMemcpyAsync(sl->nz.p, rq->noise + c0 * N_SC * 2, (size_t)n * N_SC * 8, cudaMemcpyHostToDevice, c->copy_stream));
            }
            CK(cudaEventRecord(sl->copied, c->copy_stream));
            CK(cudaStreamWaitEvent(c->stream, sl->copied, 0));
            if (pipelined) CK(cudaStreamWaitEvent(c->prep_stream, sl->copied, 0));
            sl->in_use = true;
        }
        RealisationIn in{(sl && rq->rx_pos) ? sl->pos.as<double>() : nullptr, rq->rx_pos, rq->h_true, rq->h_loop, rq->csi_noise, RNG_POS, RNG_RAYLEIGH, RNG_CSI};
        {
            cudaStream_t main_stream = c->stream;
            if (pipelined) c->stream = c->prep_stream;  // channel + realisation kernels of this batch beside the previous frame kernel
            const int rc = run_realisation(c, rq, in, c0, n, &h_true, &h_loop, want_v);
            c->stream = main_stream;
            if (rc) return 1;
            if (pipelined) {
                CK(cudaEventRecord(c->cur_ready, c->prep_stream));
                CK(cudaStreamWaitEvent(c->stream, c->cur_ready, 0));
            }
        }

        FrameArgs fa{};
        const long long hfs = (long long)K * N_SC;
        fa.tx.h_mac = h_true; fa.tx.h_mac_fs = hfs;
        fa.tx.h_pre = h_loop; fa.tx.h_pre_fs = hfs;
        fa.tx.v = nullptr; fa.tx.v_fs = 0;
        fa.tx.sscale = c->sscale.as<float>(); fa.tx.sscale_fs = N_SC;
        fa.tx.agc_inv = c->agc_inv.as<float2>(); fa.tx.agc_fs = N_SC;
        fa.tx.k = K;
        if (mcnc) {
            fa.loop = fa.tx;
            fa.loop.h_mac = h_loop;
        } else {  // single-antenna model of the transmitter, corrector.py:52-112: h = v = 1, estimate / alpha
            fa.loop.h_mac = c->cnc_ones.as<float2>(); fa.loop.h_mac_fs = 0;
            fa.loop.h_pre = c->cnc_ones.as<float2>(); fa.loop.h_pre_fs = 0;
            fa.loop.v = nullptr; fa.loop.v_fs = 0;
            fa.loop.sscale = c->cnc_v.as<float>(); fa.loop.sscale_fs = 0;
            fa.loop.agc_inv = c->cnc_agc_inv.as<float2>(); fa.loop.agc_fs = 0;
            fa.loop.k = 1;
        }
        fa.scal = c->scal.as<FrameScalars>(); fa.scal_fs = 1;
        fa.pa_p1 = (float)cfg.p_hardness;
        fa.n_iter = I;
        fa.qam.log2side = log2_side(cfg.constel_size);
        fa.n_frames = n;
        fa.stagger = c->t64_stagger;
        fa.tx_idx = nullptr; fa.noise = nullptr;
        if (rq->tx_idx) {
            k_to_native<uint8_t><<<nblk((size_t)n * N_SC), 256, 0, c->stream>>>(sl->idx.as<uint8_t>(), c->tx_idx.as<uint8_t>(), n);
            c->launches++;
            fa.tx_idx = c->tx_idx.as<uint8_t>();
        }
        if (rq->noise) {
            k_to_native<float2><<<nblk((size_t)n * N_SC), 256, 0, c->stream>>>(sl->nz.as<float2>(), c->noise.as<float2>(), n);
            c->launches++;
            fa.noise = c->noise.as<float2>();
        }
        if (sl) CK(cudaEventRecord(sl->consumed, c->stream));
        fa.noise_unit = (rq->flags & MMIMO_REQ_NOISE_UNIT) ? 1 : 0;
        fa.seed = rq->seed; fa.frame_id0 = rq->frame_id0 + c0;
        fa.det_out = want_det ? c->det.as<uint8_t>() : nullptr;
        fa.rx_eq_out = want_rxeq ? c->rx_eq.as<float2>() : nullptr;
        fa.corr_out = want_corr ? c->corr.as<float2>() : nullptr;
        fa.err_out = (rs && rs->err_per_frame) ? c->err.as<uint32_t>() : nullptr;
        fa.err_total = c->sweep.fpp ? c->sweep.counters : counters;
        fa.fpp = c->sweep.fpp; fa.frame0 = (unsigned)c0;
        if (c->rxeq_scr.reserve((size_t)8 * c->n_sm * N_SC * 8)) return 1;  // >= the largest grid launch_frame uses
        fa.rxeq_scratch = c->rxeq_scr.as<float2>();
        const int grid = n;
        // CncReceiver: the frame kernel runs the real transmission only (pass 0 -> equalised received signal), the receiver
        // loop runs one frame per 64-thread group in k_cnc_loop_t64 (cnc_loop_t64.cuh); identical decisions and counters
        const bool cnc_split = !mcnc && c->cnc_split && c->frame_kernel == 0 && I > 0;
        FrameArgs f2 = fa;
        if (cnc_split) {
            fa.n_iter = 0;
            fa.det_out = nullptr; fa.corr_out = nullptr; fa.err_out = nullptr; fa.err_total = nullptr;
            fa.rx_eq_out = c->rx_eq.as<float2>();
            f2.rx_eq_in = c->rx_eq.as<float2>();
            if (f2.err_out) CK(cudaMemsetAsync(c->err.p, 0, (size_t)n * (I + 1) * 4, c->stream));
        }
        const int grid2 = std::min((n + R64_GROUPS - 1) / R64_GROUPS, c->n_sm);
        const float2 cnc_agc_uni = make_float2((float)(1.0 / (64.0 * cnc_alpha)), 0.f);  // what k_fill_c wrote into cnc_agc_inv
        prof_begin(c, PROF_FRAME);
        switch (cfg.pa_kind) {
            case MMIMO_PA_NONE:
                launch_frame<PA_NONE>(fa, c, grid);
                if (cnc_split) k_cnc_loop_t64<PA_NONE><<<grid2, FFT_THREADS, CNC_LOOP_SMEM_BYTES, c->stream>>>(f2, c->tw_t64.as<float2>(), 1, 1.f / 64.f, cnc_agc_uni, 0, make_float2(1.f, 0.f));
                break;
            case MMIMO_PA_SOFT_LIMITER:
                launch_frame<PA_SOFT_LIMITER>(fa, c, grid);
                if (cnc_split) k_cnc_loop_t64<PA_SOFT_LIMITER><<<grid2, FFT_THREADS, CNC_LOOP_SMEM_BYTES, c->stream>>>(f2, c->tw_t64.as<float2>(), 1, 1.f / 64.f, cnc_agc_uni, 0, make_float2(1.f, 0.f));
                break;
            case MMIMO_PA_RAPP:
                launch_frame<PA_RAPP>(fa, c, grid);
                if (cnc_split) k_cnc_loop_t64<PA_RAPP><<<grid2, FFT_THREADS, CNC_LOOP_SMEM_BYTES, c->stream>>>(f2, c->tw_t64.as<float2>(), 1, 1.f / 64.f, cnc_agc_uni, 0, make_float2(1.f, 0.f));
                break;
            default:
                launch_frame<PA_TOI>(fa, c, grid);
                if (cnc_split) k_cnc_loop_t64<PA_TOI><<<grid2, FFT_THREADS, CNC_LOOP_SMEM_BYTES, c->stream>>>(f2, c->tw_t64.as<float2>(), 1, 1.f / 64.f, cnc_agc_uni, 0, make_float2(1.f, 0.f));
                break;
        }
        prof_end(c);
        c->launches += cnc_split ? 2 : 1;
        CK(cudaGetLastError());
        if (pipelined) {
            CK(cudaEventRecord(c->cur_released, c->stream));
            c->cur_used = true;
        }

        if (rs) {  // optional per-frame dumps
            if (rs->det && download_native<uint8_t>(c, c->det.as<uint8_t>(), rs->det + c0 * (I + 1) * N_SC, (size_t)n * (I + 1))) return 1;
            if (rs->corr && download_native<float2>(c, c->corr.as<float2>(), reinterpret_cast<float2*>(rs->corr) + c0 * (I + 1) * N_SC, (size_t)n * (I + 1))) return 1;
            if (rs->rx_eq && download_native<float2>(c, c->rx_eq.as<float2>(), reinterpret_cast<float2*>(rs->rx_eq) + c0 * N_SC, n)) return 1;
            if (rs->h_true && download_native<float2>(c, h_true, reinterpret_cast<float2*>(rs->h_true) + c0 * K * N_SC, (size_t)n * K)) return 1;
            if (rs->v) {
                if (download_native<float2>(c, c->v.as<float2>(), reinterpret_cast<float2*>(rs->v) + c0 * K * N_SC, (size_t)n * K)) return 1;
            }
            if (rs->agc) {
                if (c->s_a.reserve((size_t)n * N_SC * 8)) return 1;
                k_agc_from_inv<<<nblk((size_t)n * N_SC), 256, 0, c->stream>>>(c->agc_inv.as<float2>(), c->s_a.as<float2>(), (size_t)n * N_SC);
                c->launches++;
                if (download_native<float2>(c, c->s_a.as<float2>(), reinterpret_cast<float2*>(rs->agc) + c0 * N_SC, n)) return 1;
            }
            if (rs->err_per_frame) CK(cudaMemcpyAsync(rs->err_per_frame + c0 * (I + 1), c->err.p, (size_t)n * (I + 1) * 4, cudaMemcpyDeviceToHost, c->stream));
            if (rs->alpha) CK(cudaMemcpyAsync(rs->alpha + c0 * K, c->alpha.p, (size_t)n * K * 4, cudaMemcpyDeviceToHost, c->stream));
            if (rs->rx_pos && !rq->rx_pos && (cfg.channel_kind == MMIMO_CHAN_LOS || cfg.channel_kind == MMIMO_CHAN_TWO_PATH) && !rq->h_true)
                CK(cudaMemcpyAsync(rs->rx_pos + c0 * 3, c->rx_pos.p, (size_t)n * 24, cudaMemcpyDeviceToHost, c->stream));
            if (rs->pa_p0) {
                std::vector<FrameScalars> sc(n);
                CK(cudaMemcpyAsync(sc.data(), c->scal.p, (size_t)n * sizeof(FrameScalars), cudaMemcpyDeviceToHost, c->stream));
                CK(cudaStreamSynchronize(c->stream));
                for (int i = 0; i < n; ++i) rs->pa_p0[c0 + i] = sc[i].pa_p0_tx;
            }
        }
    }
    return 0;
}

extern "C" int mmimo_link_fetch(mmimo_ctx* c, int32_t n_iter, uint64_t* err, uint64_t* clean_err) {
    if (!c) return fail("NULL context");
    if (n_iter < 0 || n_iter > 62) return fail("mmimo_link_fetch: n_iter must be 0..62");
    CK(cudaSetDevice(c->device));
    unsigned long long h[65];
    CK(cudaMemcpyAsync(h, c->counters.p, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemsetAsync(c->counters.p, 0, sizeof h, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    if (err) for (int i = 0; i <= n_iter; ++i) err[i] = h[i];
    if (clean_err) clean_err[0] = h[64];
    return 0;
}

extern "C" int mmimo_link_enqueue(mmimo_ctx* c, const mmimo_sim_request* rq) { return run_link(c, rq, nullptr); }

extern "C" int mmimo_link_simulate(mmimo_ctx* c, const mmimo_sim_request* rq, mmimo_sim_result* rs) {
    if (!rs || !rs->err) return fail("mmimo_link_simulate: result.err is required");
    if (rq && rq->incl_clean_run && !rs->clean_err) return fail("mmimo_link_simulate: result.clean_err is required with incl_clean_run");
    // counters accumulated by earlier enqueue calls must not leak into this result
    if (c) { CK(cudaSetDevice(c->device)); CK(cudaMemsetAsync(c->counters.p, 0, 65 * 8, c->stream)); }
    if (run_link(c, rq, rs)) return 1;
    rs->bits_per_frame = (uint64_t)(2 * log2_side(c->cfg.constel_size)) * N_SC;
    return mmimo_link_fetch(c, rq->n_iter, rs->err, rs->clean_err);
}

// Multi-point batch: n_points * frames_per_point frames in ONE pass of the link pipeline, point p = frames
// [p * fpp, (p + 1) * fpp) with its own SNR and IBO; per-point counters.  This is the inner double loop of
// main_mp_clipping_noise_cancellation/main_mp_miso_mcnc_constant_ber_req_ebn0_vs_ibo.py:116-150 (16 IBO x 27 Eb/N0 points
// of <= 407 frames each) and of ..._ber_vs_ebn0.py / ..._ber_vs_ibo.py without a launch (and a partly filled last wave of
// CTAs) per point.
extern "C" int mmimo_link_sweep(mmimo_ctx* c, const mmimo_sim_request* base, int32_t n_points, int32_t frames_per_point,
                                const double* snr_db, const double* ibo_db, uint64_t* err) {
    if (!c || !base || !err) return fail("mmimo_link_sweep: NULL argument");
    if (!c->configured) return fail("mmimo_link_sweep: context not configured (call mmimo_configure)");
    if (n_points <= 0 || frames_per_point <= 0) return fail("mmimo_link_sweep: n_points and frames_per_point must be positive");
    if ((long long)n_points * frames_per_point > 0x7fffffffLL) return fail("mmimo_link_sweep: too many frames in one batch");
    if (base->incl_clean_run) return fail("mmimo_link_sweep: incl_clean_run is not supported in a multi-point batch");
    if (ibo_db && c->cfg.receiver_kind != MMIMO_RX_MCNC) return fail("mmimo_link_sweep: a per-point IBO needs the MCNC receiver");
    if (base->n_iter < 0 || base->n_iter > 62) return fail("mmimo_link_sweep: n_iter must be 0..62");
    CK(cudaSetDevice(c->device));
    const size_t np_ = (size_t)n_points;
    if (c->sweep_tab.reserve(2 * np_ * 8) || c->sweep_cnt.reserve(np_ * 64 * 8)) return 1;
    std::vector<double> tab(2 * np_);
    for (size_t i = 0; i < np_; ++i) {
        tab[i] = pow(10.0, (snr_db ? snr_db[i] : base->snr_db) / 10.0);
        tab[np_ + i] = pow(10.0, (ibo_db ? ibo_db[i] : c->cfg.ibo_db) / 10.0);
    }
    CK(cudaMemcpyAsync(c->sweep_tab.p, tab.data(), 2 * np_ * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));  // `tab` goes out of scope
    CK(cudaMemsetAsync(c->sweep_cnt.p, 0, np_ * 64 * 8, c->stream));
    mmimo_sim_request rq = *base;
    rq.n_frames = n_points * frames_per_point;
    c->sweep.snr_lin = c->sweep_tab.as<double>();
    c->sweep.ibo_lin = c->sweep_tab.as<double>() + np_;
    c->sweep.fpp = frames_per_point;
    c->sweep.counters = c->sweep_cnt.as<unsigned long long>();
    const int rc = run_link(c, &rq, nullptr);
    c->sweep = mmimo_ctx::Sweep();
    if (rc) return rc;
    std::vector<unsigned long long> h(np_ * 64);
    CK(cudaMemcpyAsync(h.data(), c->sweep_cnt.p, np_ * 64 * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    for (size_t p = 0; p < np_; ++p)
        for (int i = 0; i <= base->n_iter; ++i) err[p * (base->n_iter + 1) + i] = h[p * 64 + i];
    return 0;
}

extern "C" int mmimo_sdr_simulate(mmimo_ctx* c, const mmimo_sdr_request* rq, double* desired, double* distortion) {
    if (!c || !rq || !desired || !distortion) return fail("mmimo_sdr_simulate: NULL argument");
    if (!c->configured) return fail("mmimo_sdr_simulate: context not configured (call mmimo_configure)");
    if (rq->n_frames <= 0) return 0;
    CK(cudaSetDevice(c->device));
    const mmimo_link_config& cfg = c->cfg;
    const int K = cfg.n_ant;
    mmimo_sim_request lr{};
    lr.n_frames = rq->n_frames; lr.snr_db = 100.0; lr.seed = rq->seed; lr.frame_id0 = rq->frame_id0;
    const size_t chunk = std::min<size_t>({(size_t)rq->n_frames, (size_t)8192, std::max<size_t>(1, c->scratch_budget / ((size_t)K * N_SC * 8 + 65536))});
    if (c->h_true.reserve(chunk * K * N_SC * 8) || c->sscale.reserve(chunk * N_SC * 4) || c->agc_inv.reserve(chunk * N_SC * 8) ||
        c->hkvk_inv.reserve(chunk * N_SC * 8) || c->scal.reserve(chunk * sizeof(FrameScalars)) || c->alpha.reserve(chunk * K * 4) ||
        c->rx_pos.reserve(chunk * 24) || c->tx_idx.reserve(chunk * N_SC) || c->s_b.reserve(chunk * 16))
        return 1;
    for (size_t c0 = 0; c0 < (size_t)rq->n_frames; c0 += chunk) {
        const int n = (int)std::min<size_t>(chunk, (size_t)rq->n_frames - c0);
        float2 *h_true = nullptr, *h_loop = nullptr;
        RealisationIn in{nullptr, rq->rx_pos, rq->h_true, nullptr, nullptr, RNG_POS, RNG_RAYLEIGH, RNG_CSI};
        if (run_realisation(c, &lr, in, c0, n, &h_true, &h_loop, false)) return 1;
        // pass 0 of the frame kernel (array transmit -> channel, no noise) with the SDR epilogue: FrameArgs::sdr_desired
        FrameArgs fa{};
        const long long hfs = (long long)K * N_SC;
        fa.tx.h_mac = h_true; fa.tx.h_mac_fs = hfs;
        fa.tx.h_pre = h_true; fa.tx.h_pre_fs = hfs;
        fa.tx.v = nullptr; fa.tx.v_fs = 0;
        fa.tx.sscale = c->sscale.as<float>(); fa.tx.sscale_fs = N_SC;
        fa.tx.agc_inv = c->agc_inv.as<float2>(); fa.tx.agc_fs = N_SC;
        fa.tx.k = K;
        fa.loop = fa.tx;
        fa.scal = c->scal.as<FrameScalars>(); fa.scal_fs = 1;
        fa.pa_p1 = (float)cfg.p_hardness;
        fa.n_iter = 0;
        fa.qam.log2side = log2_side(cfg.constel_size);
        fa.n_frames = n;
        fa.tx_idx = nullptr;
        if (rq->tx_idx) {
            if (upload_native<uint8_t>(c, rq->tx_idx + c0 * N_SC, c->tx_idx.as<uint8_t>(), n)) return 1;
            fa.tx_idx = c->tx_idx.as<uint8_t>();
        }
        fa.seed = rq->seed; fa.frame_id0 = rq->frame_id0 + c0;
        fa.sdr_desired = c->s_b.as<double>();
        fa.sdr_distortion = c->s_b.as<double>() + n;
        const double* d_des = fa.sdr_desired;
        const double* d_dis = fa.sdr_distortion;
        prof_begin(c, PROF_FRAME);
        switch (cfg.pa_kind) {
            case MMIMO_PA_NONE: k_link_frame_t64<PA_NONE><<<std::min(n, c->n_sm), FFT_THREADS, FRAME_T64_SMEM_BYTES, c->stream>>>(fa, c->tw_t64.as<float2>()); break;
            case MMIMO_PA_SOFT_LIMITER: k_link_frame_t64<PA_SOFT_LIMITER><<<std::min(n, c->n_sm), FFT_THREADS, FRAME_T64_SMEM_BYTES, c->stream>>>(fa, c->tw_t64.as<float2>()); break;
            case MMIMO_PA_RAPP: k_link_frame_t64<PA_RAPP><<<std::min(n, c->n_sm), FFT_THREADS, FRAME_T64_SMEM_BYTES, c->stream>>>(fa, c->tw_t64.as<float2>()); break;
            default: k_link_frame_t64<PA_TOI><<<std::min(n, c->n_sm), FFT_THREADS, FRAME_T64_SMEM_BYTES, c->stream>>>(fa, c->tw_t64.as<float2>()); break;
        }
        prof_end(c);
        c->launches++;
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(desired + c0, d_des, (size_t)n * 8, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaMemcpyAsync(distortion + c0, d_dis, (size_t)n * 8, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaStreamSynchronize(c->stream));
    }
    return 0;
}
