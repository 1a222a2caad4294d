"""Generate the committed golden vectors by running the UNMODIFIED reference.

Run in the build container only (needs the read-only checkout at /root/reference):

    python tests/golden/make_golden.py

Writes tests/golden/*.npz.  Each file holds the inputs fed to the reference
(seeds, parameters, or explicit arrays) and the outputs the reference produced.
`tests/test_oracle_golden.py` pins `oracle/link_oracle.py` against them on CPU and
`tests/test_gpu_parity.py` compares the CUDA path against them on the B200.
"""
import copy
import ctypes
import multiprocessing as mp
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))

from oracle.refharness import import_reference  # noqa: E402

R = import_reference()
modulation, distortion, transceiver, antenna_array, channel, noise, corrector, mp_model, utilities = (
    R[k] for k in ("modulation", "distortion", "transceiver", "antenna_array", "channel", "noise",
                   "corrector", "mp_model", "utilities"))


def bits2idx(bits, nb):
    b = np.asarray(bits).astype(np.int64).reshape(-1, nb)
    return (b @ (1 << np.arange(nb - 1, -1, -1))).astype(np.uint16)


def build_objects(n_ant, constel=64, n_fft=4096, n_sc=2048, cp=128, pa="softlim", ibo=0.0, n_users=1,
                  rx_xy=(212.0, 212.0)):
    my_mod = modulation.OfdmQamModem(constel_size=constel, n_fft=n_fft, n_sub_carr=n_sc, cp_len=cp,
                                     n_users=n_users)
    if pa == "softlim":
        my_dist = distortion.SoftLimiter(ibo, my_mod.avg_sample_power)
    elif pa == "rapp":
        my_dist = distortion.Rapp(ibo_db=ibo, p_hardness=4.0, avg_samp_pow=my_mod.avg_sample_power)
    elif pa == "toi":
        my_dist = distortion.ThirdOrderNonLin(toi_db=ibo, avg_samp_pow=my_mod.avg_sample_power)
    my_tx = transceiver.Transceiver(modem=copy.deepcopy(my_mod), impairment=copy.deepcopy(my_dist),
                                    center_freq=int(3.5e9), carrier_spacing=int(15e3))
    my_rx = transceiver.Transceiver(modem=copy.deepcopy(my_mod), impairment=copy.deepcopy(my_dist),
                                    cord_x=rx_xy[0], cord_y=rx_xy[1], cord_z=1.5,
                                    center_freq=int(3.5e9), carrier_spacing=int(15e3))
    my_array = antenna_array.LinearArray(n_elements=n_ant, base_transceiver=my_tx, center_freq=int(3.5e9),
                                         wav_len_spacing=0.5, cord_x=0, cord_y=0, cord_z=15)
    return my_mod, my_dist, my_tx, my_rx, my_array


def make_channel(kind, my_array, my_rx):
    if kind == "los":
        ch = channel.MisoLosFd()
        ch.calc_channel_mat(tx_transceivers=my_array.array_elements, rx_transceiver=my_rx, skip_attenuation=False)
    elif kind == "two_path":
        ch = channel.MisoTwoPathFd()
        ch.calc_channel_mat(tx_transceivers=my_array.array_elements, rx_transceiver=my_rx, skip_attenuation=False)
    elif kind == "rayleigh":
        ch = channel.MisoRayleighFd(tx_transceivers=my_array.array_elements, rx_transceiver=my_rx, seed=1234)
    elif kind == "random_paths":
        ch = channel.MisoRandomPathsFd(tx_transceivers=my_array.array_elements, rx_transceiver=my_rx, seed=5, n_paths=8,
                                       max_delay_spread=1e-6)
    return ch


def golden_link(name, n_ant, chan_kind, ibo, ebn0, n_frames, iters, seeds, is_mcnc=True, incl_clean=False,
                csi_eps=None, n_fft=4096, n_sc=2048, constel=64, pa="softlim", full_frames=None, reroll=True):
    """mp_model.Link.simulate driven exactly like main_mp_miso_mcnc_ber_vs_ebn0.py:101-134 (one worker).
    full_frames: many-frame cases keep the bulky float64 arrays (noise, channel row, AGC) for the first `full_frames`
    frames only and the equalised signal as complex64; decisions, error counters, positions and PA powers of ALL
    frames are stored (the oracle regenerates everything else from the seeds)."""
    my_mod, my_dist, my_tx, my_rx, my_array = build_objects(n_ant, constel, n_fft, n_sc, 128 if n_fft >= 1024 else 16,
                                                            pa=pa, ibo=0.0)
    ch = make_channel(chan_kind, my_array, my_rx)
    nz = noise.Awgn(snr_db=10, seed=1234)
    nb = my_mod.n_bits_per_ofdm_sym
    link = mp_model.Link(mod_obj=my_mod, array_obj=my_array, std_rx_obj=my_rx, chan_obj=ch, noise_obj=nz,
                         rx_loc_var=10.0, n_err_min=int(1e12), bits_sent_max=nb * n_frames, is_mcnc=is_mcnc,
                         csi_epsylon=csi_eps)
    link.update_distortion(ibo_val_db=ibo)
    snr_db = float(utilities.ebn0_to_snr(float(ebn0), my_mod.n_sub_carr, my_mod.n_sub_carr, my_mod.constel_size))
    link.set_snr(snr_db_val=snr_db)

    rec = dict(rx_eq=[], det=[], noise=[], pos=[], h_true_row0=[], sat_pow=[], agc=[])
    orig_receive = link.my_cnc_rx.receive

    def receive(n_iters_lst, in_sig_fd, **kw):
        out = orig_receive(n_iters_lst=n_iters_lst, in_sig_fd=in_sig_fd, **kw)
        rec["rx_eq"].append(np.array(in_sig_fd))
        rec["det"].append(np.stack([bits2idx(b, my_mod.n_bits_per_symbol) for b in out]))
        rec["h_true_row0"].append(np.array(link.my_miso_chan.channel_mat_fd[0]))
        rec["sat_pow"].append(link.my_array.array_elements[0].impairment.sat_pow
                              if hasattr(link.my_array.array_elements[0].impairment, "sat_pow") else
                              link.my_array.array_elements[0].impairment.cubic_dist_coeff)
        rec["agc"].append(np.array(link.ak_hk_vk_agc_nfft))
        return out

    link.my_cnc_rx.receive = receive
    orig_noise = link.my_noise.process

    def process(in_sig, **kw):
        out = orig_noise(in_sig, **kw)
        rec["noise"].append(out - in_sig)
        return out

    link.my_noise.process = process
    orig_pos = link.my_standard_rx.set_position

    def set_position(cord_x, cord_y, cord_z):
        rec["pos"].append((cord_x, cord_y, cord_z))
        orig_pos(cord_x=cord_x, cord_y=cord_y, cord_z=cord_z)

    link.my_standard_rx.set_position = set_position

    iters = np.asarray(iters)
    n_res = len(iters) + (1 if incl_clean else 0)
    e = mp.Array(ctypes.c_double, n_res, lock=True)
    b = mp.Array(ctypes.c_double, n_res, lock=True)
    link.simulate(incl_clean, reroll, iters, list(seeds), e, b)

    n_clean = n_frames if incl_clean else 0
    bins = np.concatenate((np.arange(n_fft - n_sc // 2, n_fft), np.arange(1, n_sc // 2 + 1)))
    bit_rng = np.random.default_rng(seeds[0])
    tx_idx = np.stack([bits2idx(bit_rng.choice((0, 1), nb), my_mod.n_bits_per_symbol)
                       for _ in range(n_clean + n_frames)])
    out = dict(
        n_ant=n_ant, chan=chan_kind, ibo_db=float(ibo), snr_db=snr_db, ebn0_db=float(ebn0), n_frames=n_frames,
        iters=iters, seeds=np.asarray(seeds, dtype=np.int64), is_mcnc=int(is_mcnc), incl_clean=int(incl_clean),
        csi_eps=-1.0 if csi_eps is None else float(csi_eps), n_fft=n_fft, n_sc=n_sc, constel=constel, pa=pa, reroll=int(reroll),
        err=np.array(list(e)), bits=np.array(list(b)),
        tx_idx=tx_idx,
        noise_sc=np.stack(rec["noise"])[:, bins],
        rx_eq_sc=np.stack(rec["rx_eq"])[:, bins],
        det=np.stack(rec["det"]).astype(np.uint16),
        h_true_row0=np.stack(rec["h_true_row0"]),
        pa_p0=np.array(rec["sat_pow"], dtype=np.float64),
        agc_sc=np.stack(rec["agc"])[:, bins],
    )
    if rec["pos"]:
        out["pos"] = np.array(rec["pos"], dtype=np.float64)
    if full_frames is not None:
        for k in ("noise_sc", "h_true_row0", "agc_sc"):
            out[k] = out[k][:n_clean + full_frames] if k == "noise_sc" else out[k][:full_frames]
        out["rx_eq_sc"] = out["rx_eq_sc"].astype(np.complex64)
        out["det"] = out["det"].astype(np.uint8 if constel <= 256 else np.uint16)
        out["tx_idx"] = out["tx_idx"].astype(np.uint8 if constel <= 256 else np.uint16)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "err", list(e), "bits", list(b))


def golden_stages():
    """AntennaArray.transmit flag matrix + propagate on a K=2 full-size array."""
    my_mod, my_dist, my_tx, my_rx, my_array = build_objects(2)
    ch = make_channel("los", my_array, my_rx)
    h = ch.get_channel_mat_fd()
    my_array.set_precoding_matrix(channel_mat_fd=h, mr_precoding=True)
    my_array.update_distortion(ibo_db=0.0, avg_sample_pow=my_mod.avg_sample_power)
    bits = np.random.default_rng(7).choice((0, 1), my_mod.n_bits_per_ofdm_sym)
    x_fd, x_clean_fd = my_array.transmit(bits, out_domain_fd=True, return_both=True)
    x_td = my_array.transmit(bits, out_domain_fd=False)
    x_td_clean = my_array.transmit(bits, out_domain_fd=False, skip_dist=True)
    y = ch.propagate(in_sig_mat=x_fd)
    y_mat = ch.propagate(in_sig_mat=x_fd, sum_signals=False)
    np.savez_compressed(os.path.join(HERE, "stages_k2.npz"), bits=bits.astype(np.uint8), h=h,
                        v=my_array.get_precoding_mat(), sat_pow=my_array.array_elements[0].impairment.sat_pow,
                        x_fd=x_fd, x_clean_fd=x_clean_fd, x_td=x_td, x_td_clean=x_td_clean, y=y, y_row1=y_mat[1])
    print("stages_k2 ok", x_fd.shape, x_td.shape)


def golden_small():
    """Small-modem (n_fft=256, n_sc=128) known answers for every building block on the path."""
    out = {}
    n_fft, n_sc, n_ant = 256, 128, 4
    for m in (4, 16, 64, 256):
        out["constellation_%d" % m] = modulation.QamModem(m).constellation
    my_mod, my_dist, my_tx, my_rx, my_array = build_objects(n_ant, 64, n_fft, n_sc, 16)
    out["avg_powers"] = np.array([my_mod.avg_symbol_power, my_mod.avg_sample_power])
    ibo_vec = np.array([-9.0, -3.3, 0.0, 0.5, 3.0, 7.25, 9.0])
    out["alpha_ibo"] = ibo_vec
    out["alpha_val"] = my_mod.calc_alpha(ibo_vec)
    out["ebn0_in"] = np.arange(5, 21, 2.5)
    out["snr_out"] = utilities.ebn0_to_snr(out["ebn0_in"], n_sc, n_sc, 64)
    out["snr_out_nfft"] = utilities.ebn0_to_snr(out["ebn0_in"], n_fft, n_sc, 64)

    # modulate / detect / bits
    rng = np.random.default_rng(11)
    bits = rng.choice((0, 1), my_mod.n_bits_per_ofdm_sym)
    out["mod_bits"] = bits.astype(np.uint8)
    out["mod_symbols"] = my_mod.modulate(bits, get_symbols_only=True)
    out["mod_td"] = my_mod.modulate(bits)
    noisy = out["mod_symbols"] + 0.9 * (rng.standard_normal(n_sc) + 1j * rng.standard_normal(n_sc))
    out["det_in"] = noisy
    out["det_symbols"] = my_mod.symbol_detection(noisy)
    out["det_bits"] = my_mod.symbols_to_bits(noisy).astype(np.uint8)
    out["demod_bits"] = my_mod.demodulate(out["mod_td"]).astype(np.uint8)

    # PA models
    x = (rng.standard_normal(512) + 1j * rng.standard_normal(512)) * np.sqrt(my_mod.avg_sample_power / 2)
    x[5] = 0
    out["pa_in"] = x
    out["pa_softlim_ibo0"] = distortion.SoftLimiter(0, my_mod.avg_sample_power).process(x)
    out["pa_softlim_ibo3"] = distortion.SoftLimiter(3, my_mod.avg_sample_power).process(x)
    out["pa_rapp_ibo1_p4"] = distortion.Rapp(1, my_mod.avg_sample_power, 4.0).process(x)
    out["pa_rapp_ibo0_p2"] = distortion.Rapp(0, my_mod.avg_sample_power, 2.0).process(x)
    out["pa_toi_20"] = distortion.ThirdOrderNonLin(20, my_mod.avg_sample_power).process(x)
    rp = distortion.Rapp(1, my_mod.avg_sample_power, 4.0)
    rp.set_avg_sample_power(3.0)
    rp.set_ibo(2.0)  # quirk: uses the stale avg_samp_pow (distortion.py:157-165)
    out["rapp_stale_sat_pow"] = rp.sat_pow

    # channels
    pos = np.array([[t.cord_x, t.cord_y, t.cord_z] for t in my_array.array_elements])
    out["tx_pos"] = pos
    my_rx.set_position(215.3, 208.1, 1.5)
    out["rx_pos"] = np.array([215.3, 208.1, 1.5])
    los = make_channel("los", my_array, my_rx)
    out["h_los"] = los.get_channel_mat_fd()
    los2 = channel.MisoLosFd()
    los2.calc_channel_mat(my_array.array_elements, my_rx, skip_attenuation=True)
    out["h_los_noatt"] = los2.get_channel_mat_fd()
    out["h_two_path"] = make_channel("two_path", my_array, my_rx).get_channel_mat_fd()
    ray = make_channel("rayleigh", my_array, my_rx)
    out["h_rayleigh_0"] = np.array(ray.get_channel_mat_fd())
    ray.reroll_channel_coeffs()
    out["h_rayleigh_1"] = np.array(ray.get_channel_mat_fd())
    rp_ch = channel.MisoRandomPathsFd(tx_transceivers=my_array.array_elements, rx_transceiver=my_rx, seed=5,
                                      n_paths=8, max_delay_spread=1e-6)
    out["h_random_paths"] = np.array(rp_ch.get_channel_mat_fd())

    # precoders (single user)
    h = out["h_two_path"]
    my_array.set_precoding_matrix(channel_mat_fd=h, mr_precoding=True)
    out["v_mrt"] = my_array.get_precoding_mat()
    my_array.update_distortion(ibo_db=2.0, avg_sample_pow=my_mod.avg_sample_power)
    out["sat_pow_mrt_ibo2"] = my_array.array_elements[0].impairment.sat_pow
    out["gain_mrt"] = my_array.get_avg_precoding_gain()
    my_array.set_precoding_matrix(channel_mat_fd=h)
    out["v_phase"] = my_array.get_precoding_mat()

    # McncReceiver AGC (corrector.py:126-163)
    my_array.set_precoding_matrix(channel_mat_fd=h, mr_precoding=True)
    ch2 = make_channel("two_path", my_array, my_rx)
    mc = corrector.McncReceiver(my_array, ch2)
    out["mcnc_agc_nfft"] = mc.agc_corr_vec

    # standalone CNC receiver, K=1 AWGN (config 1, driven like main_miso_cnc_ber_vs_ebn0.py:242)
    mod1, dist1, tx1, rx1, arr1 = build_objects(1, 64, n_fft, n_sc, 16)
    tx1.impairment.set_ibo(1.0)
    cnc = corrector.CncReceiver(copy.deepcopy(mod1), copy.deepcopy(tx1.impairment))
    bits1 = rng.choice((0, 1), mod1.n_bits_per_ofdm_sym)
    sig = tx1.transmit(bits1, out_domain_fd=True)
    awgn = noise.Awgn(snr_db=22.0, seed=99)
    sig_n = awgn.process(sig, avg_sample_pow=mod1.avg_symbol_power * cnc.modem.alpha ** 2)
    sig_eq = sig_n / cnc.modem.alpha
    res = cnc.receive(n_iters_lst=[0, 1, 2, 5], in_sig_fd=sig_eq)
    out["cnc_bits"] = bits1.astype(np.uint8)
    out["cnc_in_fd"] = sig_eq
    out["cnc_alpha"] = cnc.modem.alpha
    out["cnc_sat_pow"] = cnc.impairment.sat_pow
    out["cnc_det"] = np.stack([bits2idx(b, 6) for b in res])
    res_sym = cnc.receive(n_iters_lst=[2], in_sig_fd=sig_eq, return_bits=False)
    out["cnc_corr_it2"] = res_sym[0]

    # multi-user precoders / receivers (2 users)
    mod2, dist2, tx2, rx2, arr2 = build_objects(n_ant, 64, n_fft, n_sc, 16, n_users=2)
    rxa = copy.deepcopy(rx2)
    rxa.set_position(100.0, 50.0, 1.5)
    rxb = copy.deepcopy(rx2)
    rxb.set_position(-80.0, 120.0, 1.5)
    cha = make_channel("los", arr2, rxa)
    chb = make_channel("two_path", arr2, rxb)
    hl = [cha.get_channel_mat_fd(), chb.get_channel_mat_fd()]
    out["mu_h0"], out["mu_h1"] = hl
    arr2.set_precoding_matrix(channel_mat_fd=hl, mr_precoding=True)
    out["mu_v_mrt"] = arr2.get_precoding_mat()  # [K, U, N_sc]
    arr2.update_distortion(ibo_db=1.0, avg_sample_pow=mod2.avg_sample_power)
    out["mu_sat_pow_mrt"] = arr2.array_elements[0].impairment.sat_pow
    bits_mu = rng.choice((0, 1), (2, mod2.n_bits_per_ofdm_sym))
    out["mu_bits"] = bits_mu.astype(np.uint8)
    out["mu_x_fd_mrt"] = arr2.transmit(bits_mu, out_domain_fd=True)
    # McncMuReceiver for user 1 on user-1's channel
    mcmu = corrector.McncMuReceiver(arr2, copy.deepcopy(chb), usr_idx=1)
    arr2.set_precoding_matrix(channel_mat_fd=hl, mr_precoding=True)  # ctor overwrote it with SU MRT
    arr2.update_distortion(ibo_db=1.0, avg_sample_pow=mod2.avg_sample_power)
    mcmu.update_agc()
    out["mu_agc_u1"] = mcmu.agc_corr_vec
    rx_u1 = np.sum(out["mu_x_fd_mrt"] * hl[1], axis=0)
    rx_u1 = rx_u1 + noise.Awgn(snr_db=25.0, seed=3).process(np.zeros(n_fft, dtype=complex),
                                                            avg_sample_pow=float(np.mean(np.abs(rx_u1) ** 2)))
    rx_u1 = rx_u1 / mcmu.agc_corr_vec
    out["mu_rx_u1_eq"] = rx_u1
    res = mcmu.receive(n_iters_lst=[0, 1, 2, 3], in_sig_fd=rx_u1, other_usr_bits=bits_mu[0])
    out["mu_mcnc_det_u1"] = np.stack([bits2idx(b, 6) for b in res])
    # ZF
    arr2.set_precoding_matrix(channel_mat_fd=hl, zf_precoding=True)
    out["mu_v_zf"] = arr2.get_precoding_mat()
    arr2.update_distortion(ibo_db=1.0, avg_sample_pow=mod2.avg_sample_power)
    out["mu_sat_pow_zf"] = arr2.array_elements[0].impairment.sat_pow
    out["mu_x_fd_zf"] = arr2.transmit(bits_mu, out_domain_fd=True)
    # separate sub-carrier sets per user
    arr2.set_precoding_matrix(channel_mat_fd=hl, mr_precoding=True, sep_carr_per_usr=True)
    out["mu_v_sep"] = np.stack([t.modem.precoding_mat for t in arr2.array_elements])
    # CncMuReceiver
    cmu = corrector.CncMuReceiver(copy.deepcopy(mod2), copy.deepcopy(dist2))
    cmu.update_distortion(ibo_db=1.0)
    sym_mu = np.stack([mod2.modulate(bits_mu, get_symbols_only=True)[u] for u in range(2)])
    comb = np.sqrt(2) / 2 * sym_mu[0] + np.sqrt(2) / 2 * sym_mu[1]
    grid = np.zeros(n_fft, dtype=complex)
    grid[-(n_sc // 2):] = comb[:n_sc // 2]
    grid[1:n_sc // 2 + 1] = comb[n_sc // 2:]
    td = utilities.to_time_domain(grid)
    fd = utilities.to_freq_domain(cmu.impairment.process(td), remove_cp=False)
    fd = fd / cmu.modem.alpha
    # user 0's view: remove the other user's (known) contribution and its own scaling
    rx_u0 = (fd - np.where(np.abs(grid) > 0, 0, 0))  # full band
    sc = np.concatenate((rx_u0[-n_sc // 2:], rx_u0[1:n_sc // 2 + 1]))
    sc_u0 = (sc - np.sqrt(2) / 2 * sym_mu[1]) / (np.sqrt(2) / 2)
    in_fd = np.ones(n_fft, dtype=complex)
    in_fd[-(n_sc // 2):] = sc_u0[:n_sc // 2]
    in_fd[1:n_sc // 2 + 1] = sc_u0[n_sc // 2:]
    out["mu_cnc_in_fd"] = in_fd
    out["mu_cnc_other_sym"] = sym_mu[1]
    out["mu_cnc_alpha"] = cmu.modem.alpha
    out["mu_cnc_sat_pow"] = cmu.impairment.sat_pow
    res = cmu.receive(n_iters_lst=[0, 1, 3], in_sig_fd=in_fd, other_usr_symbols=sym_mu[1])
    out["mu_cnc_det"] = np.stack([bits2idx(b, 6) for b in res])

    np.savez_compressed(os.path.join(HERE, "small_blocks.npz"), **out)
    print("small_blocks ok:", len(out), "arrays")


if __name__ == "__main__":
    golden_full_size_extras()
    golden_small()
    golden_stages()
    golden_link("link_k64_los_mcnc", 64, "los", 0.0, 15.0, 3, [0, 1, 2, 3, 4, 5], [1, 2, 3])
    golden_link("link_k8_two_path_mcnc_clean", 8, "two_path", 3.0, 12.0, 3, [0, 1, 2, 3], [11, 12, 13],
                incl_clean=True)
    golden_link("link_k4_rayleigh_cnc", 4, "rayleigh", 1.0, 14.0, 3, [0, 1, 2, 5], [21, 22, 23], is_mcnc=False)
    golden_link("link_k8_los_csi", 8, "los", 0.0, 15.0, 2, [0, 1, 2], [31, 32, 33, 34], csi_eps=0.2)
    golden_link("link_k8_los_rapp", 8, "los", 1.0, 15.0, 2, [0, 1, 2], [41, 42, 43], pa="rapp")
    golden_link("link_k16_los_small", 16, "los", 0.0, 15.0, 4, [0, 1, 2, 3, 4, 5, 6, 7, 8], [51, 52, 53],
                n_fft=1024, n_sc=512, constel=16)
    golden_random_paths_fixed()


def golden_random_paths_fixed():
    """Link.simulate(reroll_chan=False) over a MisoRandomPathsFd channel (BASELINE config 3; the reference cannot re-roll this
    channel through Link - it has no calc_channel_mat - but a fixed one works, mp_model.py:190-200)"""
    golden_link("link_k8_random_paths_fixed", 8, "random_paths", 1.0, 16.0, 3, [0, 1, 2, 3], [61, 62, 63], reroll=False)


def golden_full_size_extras():
    """Full-size (n_fft 4096 / 2048 sub-carriers) cases for BASELINE configs 4 and 5: SDR snapshots as
    main_beampatterns_plotting/sdr_mp.py:124-151 computes them, and the two-user MRT / ZF frame of
    main_multiuser/main_multiuser_mcnc_ber_vs_ebn0.py:514-642 (per-user McncMuReceiver that knows the other user's bits)."""
    out = {}
    n_ant = 8
    # ---- SDR (config 4) ----
    my_mod, my_dist, my_tx, my_rx, my_array = build_objects(n_ant)
    rng = np.random.default_rng(17)
    des, dis, poss, bits_l = [], [], [], []
    for ibo in (0.0, 3.0):
        my_rx.set_position(212.0 + rng.uniform(-5, 5), 212.0 + rng.uniform(-5, 5), 1.5)
        ch = make_channel("los", my_array, my_rx)
        my_array.set_precoding_matrix(channel_mat_fd=ch.channel_mat_fd, mr_precoding=True)
        my_array.update_distortion(ibo_db=ibo, avg_sample_pow=my_mod.avg_sample_power)
        vk = my_array.get_precoding_mat()
        vk_pow = np.sum(np.power(np.abs(vk), 2), axis=1)
        ibo_vec = 10 * np.log10(10 ** (ibo / 10) * my_mod.n_sub_carr / (vk_pow * n_ant))
        ak = np.expand_dims(my_mod.calc_alpha(ibo_db=ibo_vec), axis=1)
        bits = rng.choice((0, 1), my_mod.n_bits_per_ofdm_sym)
        x_fd, x_clean = my_array.transmit(bits, out_domain_fd=True, skip_dist=False, return_both=True)
        rx = ch.propagate(in_sig_mat=x_fd, sum_signals=False)
        rx_sc = np.concatenate((rx[:, -1024:], rx[:, 1:1025]), axis=1)
        cl = ch.propagate(in_sig_mat=x_clean, sum_signals=False)
        cl_sc = np.concatenate((cl[:, -1024:], cl[:, 1:1025]), axis=1)
        dist_sig = rx_sc - ak * cl_sc
        des.append(np.sum(np.abs(np.sum(ak * cl_sc, axis=0)) ** 2))
        dis.append(np.sum(np.abs(np.sum(dist_sig, axis=0)) ** 2))
        poss.append((my_rx.cord_x, my_rx.cord_y, my_rx.cord_z))
        bits_l.append(bits.astype(np.uint8))
    out["sdr_ibo"] = np.array([0.0, 3.0])
    out["sdr_pos"] = np.array(poss)
    out["sdr_bits"] = np.stack(bits_l)
    out["sdr_desired"] = np.array(des)
    out["sdr_distortion"] = np.array(dis)

    # ---- two users (config 5) ----
    mod2, dist2, tx2, rx2, arr2 = build_objects(n_ant, n_users=2)
    rxs = [copy.deepcopy(rx2), copy.deepcopy(rx2)]
    rxs[0].set_position(np.cos(np.deg2rad(-30)) * 100.0, np.sin(np.deg2rad(-30)) * 100.0, 1.5)
    rxs[1].set_position(np.cos(np.deg2rad(30)) * 316.3, np.sin(np.deg2rad(30)) * 316.3, 1.5)
    chans = [make_channel("los", arr2, r) for r in rxs]
    hl = [c.get_channel_mat_fd() for c in chans]
    out["mu_h"] = np.stack(hl)
    bits_mu = rng.choice((0, 1), (2, mod2.n_bits_per_ofdm_sym))
    out["mu_bits"] = bits_mu.astype(np.uint8)
    for tag, kw in (("mrt", dict(mr_precoding=True)), ("zf", dict(zf_precoding=True))):
        rcv = []
        for u in range(2):
            arr_u = copy.deepcopy(arr2)
            rcv.append(corrector.McncMuReceiver(arr_u, copy.deepcopy(chans[u]), usr_idx=u))
        arr2.set_precoding_matrix(channel_mat_fd=hl, **kw)
        arr2.update_distortion(ibo_db=2.0, avg_sample_pow=mod2.avg_sample_power)
        out["mu_v_" + tag] = arr2.get_precoding_mat()
        out["mu_sat_pow_" + tag] = arr2.array_elements[0].impairment.sat_pow
        x = arr2.transmit(bits_mu, out_domain_fd=True)
        out["mu_x_fd_rows07_" + tag] = x[[0, 7]]
        for u in range(2):
            rcv[u].antenna_array.set_precoding_matrix(channel_mat_fd=hl, **kw)
            rcv[u].antenna_array.update_distortion(ibo_db=2.0, avg_sample_pow=mod2.avg_sample_power)
            rcv[u].update_agc()
            y = chans[u].propagate(in_sig_mat=x)
            nz = noise.Awgn(snr_db=28.0, seed=40 + u)
            y = nz.process(y, avg_sample_pow=float(np.mean(np.abs(y[np.r_[1:1025, 3072:4096]]) ** 2)))
            y = y / rcv[u].agc_corr_vec
            res = rcv[u].receive(n_iters_lst=[0, 1, 2], in_sig_fd=y, other_usr_bits=bits_mu[1 - u])
            out["mu_rx_eq_%s_u%d" % (tag, u)] = y
            out["mu_agc_%s_u%d" % (tag, u)] = rcv[u].agc_corr_vec
            out["mu_det_%s_u%d" % (tag, u)] = np.stack([bits2idx(b, 6) for b in res])
            out["mu_err_%s_u%d" % (tag, u)] = np.array([int(np.sum(b != bits_mu[u])) for b in res])
    np.savez_compressed(os.path.join(HERE, "full_size_extras.npz"), **out)
    print("full_size_extras ok: sdr", out["sdr_desired"] / out["sdr_distortion"],
          "mu err", {k: v.tolist() for k, v in out.items() if k.startswith("mu_err")})
