// Host orchestration of the fused two-user frame loop (include/mmimo_b200.h: mmimo_mu_link_simulate) - included by
// mmimo_b200.cu.  Receiver instances are ordered (realisation, user): instance i = 2 * realisation + user.

template <int PA>
static void launch_frame_mu(const FrameArgs& fa, const mmimo_ctx* c, int n) {
    if (c->frame_kernel != 2) {  // default: the two-user instantiation of the scaled-tangent radix-64 kernel
        k_link_frame_t64<PA, true><<<std::min(n, c->n_sm), FFT_THREADS, FRAME_T64_SMEM_BYTES_MU, c->stream>>>(fa, c->tw_t64.as<float2>());
        return;
    }
    const int grid = std::min(n, 2 * c->n_sm);  // MMIMO_FRAME_KERNEL=r16: the round-1 three-pass kernel
    k_link_frame<PA, true><<<grid, FFT_THREADS, FRAME_SMEM_BYTES_MU, c->stream>>>(fa, c->tw4096.as<float2>(), c->tw256.as<float2>());
}

extern "C" int mmimo_mu_link_simulate(mmimo_ctx* c, const mmimo_mu_request* rq, mmimo_mu_result* rs) {
    if (!c || !rq || !rs || !rs->err) return fail("mmimo_mu_link_simulate: NULL argument (result.err is required)");
    if (!c->configured) return fail("mmimo_mu_link_simulate: context not configured (call mmimo_configure)");
    if (rq->n_iter < 0 || rq->n_iter > 31) return fail("mmimo_mu_link_simulate: n_iter must be 0..31 (got %d)", rq->n_iter);
    if (rq->receiver < 0 || rq->receiver > MMIMO_MU_RX_CNC_MU) return fail("mmimo_mu_link_simulate: unknown receiver %d", rq->receiver);
    if (rq->precoding != MMIMO_MU_MRT && rq->precoding != MMIMO_MU_ZF) return fail("mmimo_mu_link_simulate: unknown precoding %d", rq->precoding);
    if (rq->frame_id0 >> 62) return fail("mmimo_mu_link_simulate: frame_id0 too large");
    CK(cudaSetDevice(c->device));
    const mmimo_link_config& cfg = c->cfg;
    const int K = cfg.n_ant, I = rq->n_iter;
    if (K < 3 && rq->precoding == MMIMO_MU_ZF) return fail("mmimo_mu_link_simulate: zero forcing for 2 users needs n_ant > 2");
    rs->bits_per_frame = (uint64_t)(2 * log2_side(cfg.constel_size)) * N_SC;
    for (int i = 0; i < 2 * (I + 1); ++i) rs->err[i] = 0;
    if (rq->n_frames <= 0) return 0;
    CK(cudaMemsetAsync(c->counters.p, 0, 65 * 8, c->stream));

    const bool want_v = rs->v != nullptr;
    size_t per_real = 2 * ((size_t)K * N_SC * 8 * 2 + (size_t)N_SC * (8 * 3 + 4 + 1) + 64 + (size_t)K * 4);
    if (rs->det) per_real += 2 * (size_t)(I + 1) * N_SC;
    size_t chunk = std::max<size_t>(1, c->scratch_budget / per_real);
    chunk = std::min<size_t>(chunk, (size_t)rq->n_frames);
    chunk = std::min<size_t>(chunk, 4096);
    const size_t ni_max = 2 * chunk;  // instances per chunk
    if (c->h_true.reserve(ni_max * K * N_SC * 8) || c->v.reserve(ni_max * K * N_SC * 8) || c->sscale.reserve(N_SC * 4) ||
        c->agc_inv.reserve(ni_max * N_SC * 8) || c->scal.reserve(ni_max * sizeof(FrameScalars)) ||
        c->alpha.reserve(chunk * K * 4) || c->rx_pos.reserve(ni_max * 24 + 64) || c->tx_idx.reserve(ni_max * N_SC) ||
        c->noise.reserve(ni_max * N_SC * 8) || c->rx_eq.reserve(ni_max * N_SC * 8) ||
        c->rxeq_scr.reserve((size_t)8 * c->n_sm * N_SC * 8))
        return 1;
    if (rs->det && c->det.reserve(ni_max * (I + 1) * N_SC)) return 1;
    k_fill_f<<<nblk(N_SC), 256, 0, c->stream>>>(c->sscale.as<float>(), 1.f / 64.f, N_SC);
    c->launches++;
    // nominal user positions live behind the per-instance positions in c->rx_pos
    double* usr_tab = c->rx_pos.as<double>() + ni_max * 3;
    CK(cudaMemcpyAsync(usr_tab, rq->usr_pos, 48, cudaMemcpyHostToDevice, c->stream));
    unsigned long long* counters = c->counters.as<unsigned long long>();

    for (size_t c0 = 0; c0 < (size_t)rq->n_frames; c0 += chunk) {
        const int nr = (int)std::min<size_t>(chunk, (size_t)rq->n_frames - c0);
        const int ni = 2 * nr;
        const size_t i0 = 2 * c0;
        float2* h = c->h_true.as<float2>();
        // ---- channels of both users -----------------------------------------------------------------
        if (rq->h) {
            if (upload_native<float2>(c, reinterpret_cast<const float2*>(rq->h) + i0 * K * N_SC, h, (size_t)ni * K)) return 1;
        } else if (cfg.channel_kind == MMIMO_CHAN_LOS || cfg.channel_kind == MMIMO_CHAN_TWO_PATH) {
            ChanArgs ca{};
            ca.tx_pos = c->tx_pos.as<double>();
            if (rq->rx_pos) {
                CK(cudaMemcpyAsync(c->rx_pos.p, rq->rx_pos + i0 * 3, (size_t)ni * 24, cudaMemcpyHostToDevice, c->stream));
                ca.rx_pos_in = c->rx_pos.as<double>();
            }
            ca.freq_slot = c->freq_slot.as<double>();
            ca.loc_var = cfg.rx_loc_var;
            ca.gain_lin = sqrt(pow(10.0, cfg.ant_gain_db / 10.0));
            ca.kind = cfg.channel_kind == MMIMO_CHAN_TWO_PATH ? 1 : 0;
            ca.skip_attenuation = cfg.skip_attenuation;
            ca.seed = rq->seed; ca.frame_id0 = 2 * rq->frame_id0 + i0;
            ca.pos_stream = RNG_POS;
            ca.h_out = h; ca.n_ant = K;
            ca.rx0_tab = usr_tab; ca.n_usr = 2;
            k_gen_channel<<<dim3((K + CHAN_ANT_PER_CTA - 1) / CHAN_ANT_PER_CTA, ni), 256, 0, c->stream>>>(ca);
            c->launches++;
        } else if (cfg.channel_kind == MMIMO_CHAN_RAYLEIGH) {
            RayleighArgs ra{};
            ra.tx_pos = c->tx_pos.as<double>();
            ra.freq_slot = c->freq_slot.as<double>();
            ra.rx0[0] = cfg.rx_pos[0]; ra.rx0[1] = cfg.rx_pos[1]; ra.rx0[2] = cfg.rx_pos[2];
            ra.gain_lin = sqrt(pow(10.0, cfg.ant_gain_db / 10.0));
            ra.skip_attenuation = cfg.skip_attenuation;
            ra.seed = rq->seed; ra.frame_id0 = 2 * rq->frame_id0 + i0; ra.stream = RNG_RAYLEIGH;
            ra.h_out = h; ra.n_ant = K;
            k_gen_rayleigh<<<dim3((K + CHAN_ANT_PER_CTA - 1) / CHAN_ANT_PER_CTA, ni), 256, 0, c->stream>>>(ra);
            c->launches++;
        } else {
            return fail("mmimo_mu_link_simulate: channel_kind INJECTED needs h in the request");
        }
        // ---- precoders, PA re-normalisation, per-user AGC ------------------------------------------------
        MuPrepArgs mp{};
        mp.h = h; mp.v = c->v.as<float2>(); mp.agc_inv = c->agc_inv.as<float2>();
        mp.scal = c->scal.as<FrameScalars>(); mp.alpha_out = c->alpha.as<float>();
        mp.n_ant = K; mp.n_real = nr; mp.precoding = rq->precoding; mp.pa_kind = cfg.pa_kind;
        mp.ibo_lin = pow(10.0, cfg.ibo_db / 10.0);
        mp.avg_symbol_power = avg_symbol_power(cfg.constel_size);
        mp.avg_sample_power = mp.avg_symbol_power * ((double)cfg.n_sub_carr / cfg.n_fft);
        mp.snr_lin = pow(10.0, rq->snr_db / 10.0);
        mp.loop_cnc = rq->receiver == MMIMO_MU_RX_CNC || rq->receiver == MMIMO_MU_RX_CNC_MU;
        k_prep_mu<<<nr, MU_T, 0, c->stream>>>(mp);
        c->launches++;

        FrameArgs fa{};
        double cnc_alpha = 1.0;
        const long long hfs = (long long)K * N_SC;
        fa.tx.h_mac = h; fa.tx.h_mac_fs = hfs;
        fa.tx.h_pre = h; fa.tx.h_pre_fs = hfs;
        fa.tx.v = c->v.as<float2>(); fa.tx.v_fs = hfs;
        fa.tx.sscale = c->sscale.as<float>(); fa.tx.sscale_fs = 0;
        fa.tx.agc_inv = c->agc_inv.as<float2>(); fa.tx.agc_fs = N_SC;
        fa.tx.k = K;
        fa.loop = fa.tx;
        fa.loop_single_user = (rq->receiver == MMIMO_MU_RX_MCNC_SU || rq->receiver == MMIMO_MU_RX_CNC) ? 1 : 0;
        if (mp.loop_cnc) {
            // single-antenna model of the transmitter (corrector.py:52-112, :288-345): h = 1, estimate / alpha; CncMuReceiver
            // adds the other user's true symbols with equal power: explicit "precoder" rows sqrt(2)/2 for both users
            if (c->cnc_ones.reserve(N_SC * 8) || c->cnc_v.reserve(N_SC * 4) || c->cnc_agc_inv.reserve(N_SC * 8) || c->s_a.reserve(N_SC * 8)) return 1;
            const double alpha = cnc_alpha = (cfg.pa_kind == MMIMO_PA_TOI || cfg.pa_kind == MMIMO_PA_NONE) ? 1.0 : calc_alpha_host(cfg.ibo_db);
            k_fill_c<<<nblk(N_SC), 256, 0, c->stream>>>(c->cnc_ones.as<float2>(), make_float2(1.f, 0.f), N_SC);
            k_fill_f<<<nblk(N_SC), 256, 0, c->stream>>>(c->cnc_v.as<float>(), 1.f / 64.f, N_SC);
            k_fill_c<<<nblk(N_SC), 256, 0, c->stream>>>(c->cnc_agc_inv.as<float2>(), make_float2((float)(1.0 / (64.0 * alpha)), 0.f), N_SC);
            k_fill_c<<<nblk(N_SC), 256, 0, c->stream>>>(c->s_a.as<float2>(), make_float2(0.70710678118654752440f, 0.f), N_SC);
            c->launches += 4;
            fa.loop.h_mac = c->cnc_ones.as<float2>(); fa.loop.h_mac_fs = 0;
            fa.loop.h_pre = c->cnc_ones.as<float2>(); fa.loop.h_pre_fs = 0;
            fa.loop.v = rq->receiver == MMIMO_MU_RX_CNC_MU ? c->s_a.as<float2>() : nullptr; fa.loop.v_fs = 0;
            fa.loop.sscale = c->cnc_v.as<float>(); fa.loop.sscale_fs = 0;
            fa.loop.agc_inv = c->cnc_agc_inv.as<float2>(); fa.loop.agc_fs = 0;
            fa.loop.k = 1;
        } else if (rq->receiver == MMIMO_MU_RX_MCNC_SU) {
            // the script's receivers: McncReceiver over a single-user copy of the array, MRT of the user's own channel
            // (main_multiuser_mcnc_ber_vs_ebn0.py:563-572) -> the single-user realisation kernel per instance
            if (c->s_a.reserve((size_t)ni * N_SC * 4) || c->s_b.reserve((size_t)ni * N_SC * 8) || c->s_c.reserve((size_t)ni * N_SC * 8) ||
                c->s_d.reserve((size_t)ni * (sizeof(FrameScalars) + (size_t)K * 4)))
                return 1;
            PrepArgs pa{};
            pa.h_loop = h; pa.h_fs = hfs; pa.v_in = nullptr; pa.v_out = nullptr;
            pa.sscale = c->s_a.as<float>(); pa.agc_inv = c->s_b.as<float2>(); pa.hkvk_inv = c->s_c.as<float2>();
            pa.scal = c->s_d.as<FrameScalars>();
            pa.alpha_out = reinterpret_cast<float*>(c->s_d.as<FrameScalars>() + ni);
            pa.n_ant = K; pa.n_frames = ni; pa.pa_kind = cfg.pa_kind; pa.is_mcnc = 1;
            pa.ibo_lin = mp.ibo_lin; pa.avg_symbol_power = mp.avg_symbol_power; pa.avg_sample_power = mp.avg_sample_power;
            pa.snr_lin = mp.snr_lin;
            k_prep_realisation<false><<<ni, PREP_T, 0, c->stream>>>(pa, SweepArgs{});
            // With zero forcing the script's single-user arrays (n_users = 1) get their PA saturation power from the phase-only
            // precoder of antenna_array.py:175-178 (mean |v|^2 = 1: ibo * avg_sample_power) before McncReceiver.__init__
            // overwrites the precoder with MRT (corrector.py:135) - reproduced as written (tests/golden/mu_su_loop.npz)
            const bool phase_pa = rq->precoding == MMIMO_MU_ZF && (cfg.pa_kind == MMIMO_PA_SOFT_LIMITER || cfg.pa_kind == MMIMO_PA_RAPP);
            k_copy_loop_p0<<<nblk(ni), 256, 0, c->stream>>>(c->scal.as<FrameScalars>(), c->s_d.as<FrameScalars>(), ni,
                                                          phase_pa ? (float)(mp.ibo_lin * mp.avg_sample_power) : 0.f);
            c->launches += 2;
            fa.loop.v = nullptr; fa.loop.v_fs = 0;
            fa.loop.sscale = c->s_a.as<float>(); fa.loop.sscale_fs = N_SC;
            fa.loop.agc_inv = c->s_b.as<float2>(); fa.loop.agc_fs = N_SC;
        }
        fa.scal = c->scal.as<FrameScalars>(); fa.scal_fs = 1;
        fa.pa_p1 = (float)cfg.p_hardness;
        fa.n_iter = I;
        fa.qam.log2side = log2_side(cfg.constel_size);
        fa.n_frames = ni;
        if (rq->tx_idx) {
            if (upload_native<uint8_t>(c, rq->tx_idx + i0 * N_SC, c->tx_idx.as<uint8_t>(), ni)) return 1;
            fa.tx_idx = c->tx_idx.as<uint8_t>();
        }
        if (rq->noise) {
            if (upload_native<float2>(c, reinterpret_cast<const float2*>(rq->noise) + i0 * N_SC, c->noise.as<float2>(), ni)) return 1;
            fa.noise = c->noise.as<float2>();
        }
        if (rq->rx_eq_in) {
            if (upload_native<float2>(c, reinterpret_cast<const float2*>(rq->rx_eq_in) + i0 * N_SC, c->rx_eq.as<float2>(), ni)) return 1;
            fa.rx_eq_in = c->rx_eq.as<float2>();
        }
        fa.seed = rq->seed; fa.frame_id0 = 2 * rq->frame_id0 + i0;
        fa.det_out = rs->det ? c->det.as<uint8_t>() : nullptr;
        fa.rx_eq_out = (rs->rx_eq && !rq->rx_eq_in) ? c->rx_eq.as<float2>() : nullptr;
        fa.err_total = counters;
        fa.rxeq_scratch = c->rxeq_scr.as<float2>();
        // CncReceiver / CncMuReceiver per user: the two-user frame kernel runs the real transmission only, the receiver loops
        // run one instance per 64-thread group in k_cnc_loop_t64 (cnc_loop_t64.cuh); identical decisions and counters
        const bool cnc_rx = rq->receiver == MMIMO_MU_RX_CNC || rq->receiver == MMIMO_MU_RX_CNC_MU;
        const bool cnc_split = cnc_rx && c->cnc_split && c->frame_kernel == 0 && I > 0;
        FrameArgs f2 = fa;
        if (cnc_split) {
            f2.rx_eq_in = c->rx_eq.as<float2>();
            fa.n_iter = 0; fa.det_out = nullptr; fa.err_total = nullptr;
            fa.rx_eq_out = rq->rx_eq_in ? nullptr : c->rx_eq.as<float2>();
        }
        const int grid2 = std::min((ni + R64_GROUPS - 1) / R64_GROUPS, c->n_sm);
        const int mode2 = rq->receiver == MMIMO_MU_RX_CNC_MU ? 3 : 1;
        const float2 uv = make_float2(0.70710678118654752440f, 0.f), uagc = make_float2((float)(1.0 / (64.0 * cnc_alpha)), 0.f);
        prof_begin(c, PROF_FRAME);
        const bool pass0 = !(cnc_split && rq->rx_eq_in);  // a receiver-only call has its equalised signal already
        switch (cfg.pa_kind) {
            case MMIMO_PA_NONE:
                if (pass0) launch_frame_mu<PA_NONE>(fa, c, ni);
                if (cnc_split) k_cnc_loop_t64<PA_NONE><<<grid2, FFT_THREADS, CNC_LOOP_SMEM_BYTES, c->stream>>>(f2, c->tw_t64.as<float2>(), 1, 1.f / 64.f, uagc, mode2, uv);
                break;
            case MMIMO_PA_SOFT_LIMITER:
                if (pass0) launch_frame_mu<PA_SOFT_LIMITER>(fa, c, ni);
                if (cnc_split) k_cnc_loop_t64<PA_SOFT_LIMITER><<<grid2, FFT_THREADS, CNC_LOOP_SMEM_BYTES, c->stream>>>(f2, c->tw_t64.as<float2>(), 1, 1.f / 64.f, uagc, mode2, uv);
                break;
            case MMIMO_PA_RAPP:
                if (pass0) launch_frame_mu<PA_RAPP>(fa, c, ni);
                if (cnc_split) k_cnc_loop_t64<PA_RAPP><<<grid2, FFT_THREADS, CNC_LOOP_SMEM_BYTES, c->stream>>>(f2, c->tw_t64.as<float2>(), 1, 1.f / 64.f, uagc, mode2, uv);
                break;
            default:
                if (pass0) launch_frame_mu<PA_TOI>(fa, c, ni);
                if (cnc_split) k_cnc_loop_t64<PA_TOI><<<grid2, FFT_THREADS, CNC_LOOP_SMEM_BYTES, c->stream>>>(f2, c->tw_t64.as<float2>(), 1, 1.f / 64.f, uagc, mode2, uv);
                break;
        }
        prof_end(c);
        c->launches += (pass0 ? 1 : 0) + (cnc_split ? 1 : 0);
        CK(cudaGetLastError());

        if (rs->det && download_native<uint8_t>(c, c->det.as<uint8_t>(), rs->det + i0 * (I + 1) * N_SC, (size_t)ni * (I + 1))) return 1;
        if (rs->rx_eq && !rq->rx_eq_in && download_native<float2>(c, c->rx_eq.as<float2>(), reinterpret_cast<float2*>(rs->rx_eq) + i0 * N_SC, ni)) return 1;
        if (want_v && download_native<float2>(c, c->v.as<float2>(), reinterpret_cast<float2*>(rs->v) + i0 * K * N_SC, (size_t)ni * K)) return 1;
        if (rs->agc) {
            if (c->hkvk_inv.reserve((size_t)ni * N_SC * 8)) return 1;
            float2* tmp = c->hkvk_inv.as<float2>();
            k_agc_from_inv<<<nblk((size_t)ni * N_SC), 256, 0, c->stream>>>(c->agc_inv.as<float2>(), tmp, (size_t)ni * N_SC);
            c->launches++;
            if (download_native<float2>(c, tmp, reinterpret_cast<float2*>(rs->agc) + i0 * N_SC, ni)) return 1;
        }
        if (rs->pa_p0) {
            std::vector<FrameScalars> sc(ni);
            CK(cudaMemcpyAsync(sc.data(), c->scal.p, (size_t)ni * sizeof(FrameScalars), cudaMemcpyDeviceToHost, c->stream));
            CK(cudaStreamSynchronize(c->stream));
            for (int i = 0; i < nr; ++i) rs->pa_p0[c0 + i] = sc[2 * i].pa_p0_tx;
        }
    }
    unsigned long long hc[65];
    CK(cudaMemcpyAsync(hc, c->counters.p, sizeof hc, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemsetAsync(c->counters.p, 0, sizeof hc, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    for (int u = 0; u < 2; ++u)
        for (int i = 0; i <= I; ++i) rs->err[u * (I + 1) + i] = hc[u * 32 + i];
    return 0;
}
