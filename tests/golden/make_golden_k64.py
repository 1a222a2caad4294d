"""Many-frame / 64-antenna golden vectors from the UNMODIFIED reference (build container only, ~5 min of CPU):

    python tests/golden/make_golden_k64.py

  link_k64_los_mcnc32.npz   the headline shape (BASELINE configs[1]): K=64 LOS MRT, soft limiter IBO 0 dB, Eb/N0 15 dB,
                            MCNC iterations 0..5, 32 frames (same seeds as link_k64_los_mcnc: its 3 frames are the first 3)
  link_k64_los_cnc16.npz    the same link with the CNC receiver (configs[1] names CNC and MCNC), 16 frames, iterations 0,1,2,5
  link_k64_two_path_mcnc8.npz, link_k64_rayleigh_mcnc8.npz   BASELINE configs[2] at its named size: K=64 two-path and Rayleigh
                            channels, MCNC iterations 0..5, 8 frames each (python tests/golden/make_golden_k64.py chan)
  mu_k64.npz                two users, K=64, MU-MRT and ZF, McncMuReceiver per user (configs[4], the frame of
                            main_multiuser/main_multiuser_mcnc_ber_vs_ebn0.py:514-642); channels are NOT stored (2 x 4 MiB):
                            the users' positions are, and channel.MisoLosFd is pinned by link_oracle.channel_los elsewhere
"""
import copy
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402  (imports the reference through oracle/refharness.py)


def golden_mu_k64(n_ant=64, ibo=0.0, ebn0=15.0, iters=(0, 1, 2, 3)):
    antenna_array, corrector, noise, utilities = mg.antenna_array, mg.corrector, mg.noise, mg.utilities
    out = {}
    mod2, dist2, tx2, rx2, arr2 = mg.build_objects(n_ant, n_users=2)
    rxs = [copy.deepcopy(rx2), copy.deepcopy(rx2)]
    rxs[0].set_position(np.cos(np.deg2rad(-30)) * 100.0, np.sin(np.deg2rad(-30)) * 100.0, 1.5)
    rxs[1].set_position(np.cos(np.deg2rad(30)) * 316.3, np.sin(np.deg2rad(30)) * 316.3, 1.5)
    out["usr_pos"] = np.array([[r.cord_x, r.cord_y, r.cord_z] for r in rxs])
    chans = [mg.make_channel("los", arr2, r) for r in rxs]
    hl = [c.get_channel_mat_fd() for c in chans]
    out["h_rows"] = np.stack([h[[0, n_ant - 1]] for h in hl])  # [U, 2, n_fft]: first and last antenna, as a check
    rng = np.random.default_rng(23)
    bits_mu = rng.choice((0, 1), (2, mod2.n_bits_per_ofdm_sym))
    out["bits"] = bits_mu.astype(np.uint8)
    snr_db = float(utilities.ebn0_to_snr(ebn0, mod2.n_sub_carr, mod2.n_sub_carr, mod2.constel_size))
    out["snr_db"], out["ibo_db"], out["n_ant"], out["iters"] = snr_db, float(ibo), n_ant, np.asarray(iters)
    bins = np.r_[3072:4096, 1:1025]
    for tag, kw in (("mrt", dict(mr_precoding=True)), ("zf", dict(zf_precoding=True))):
        rcv = []
        for u in range(2):
            arr_u = copy.deepcopy(arr2)
            rcv.append(corrector.McncMuReceiver(arr_u, copy.deepcopy(chans[u]), usr_idx=u))
        arr2.set_precoding_matrix(channel_mat_fd=hl, **kw)
        arr2.update_distortion(ibo_db=ibo, avg_sample_pow=mod2.avg_sample_power)
        v = arr2.get_precoding_mat()  # [K, U, N_sc]
        out["v_rows_" + tag] = v[[0, n_ant - 1]]
        out["sat_pow_" + tag] = arr2.array_elements[0].impairment.sat_pow
        x = arr2.transmit(bits_mu, out_domain_fd=True)
        out["x_fd_rows_" + tag] = x[[0, n_ant - 1]]
        for u in range(2):
            rcv[u].antenna_array.set_precoding_matrix(channel_mat_fd=hl, **kw)
            rcv[u].antenna_array.update_distortion(ibo_db=ibo, avg_sample_pow=mod2.avg_sample_power)
            rcv[u].update_agc()
            y0 = chans[u].propagate(in_sig_mat=x)
            nz = noise.Awgn(snr_db=snr_db, seed=70 + u)
            y = nz.process(y0, avg_sample_pow=float(np.mean(np.abs(y0[bins]) ** 2)))
            out["noise_sc_%s_u%d" % (tag, u)] = (y - y0)[bins]
            y = y / rcv[u].agc_corr_vec
            res = rcv[u].receive(n_iters_lst=list(iters), in_sig_fd=y, other_usr_bits=bits_mu[1 - u])
            out["rx_eq_sc_%s_u%d" % (tag, u)] = y[bins]
            out["agc_sc_%s_u%d" % (tag, u)] = rcv[u].agc_corr_vec[bins]
            out["det_%s_u%d" % (tag, u)] = np.stack([mg.bits2idx(b, 6) for b in res]).astype(np.uint8)
            out["err_%s_u%d" % (tag, u)] = np.array([int(np.sum(b != bits_mu[u])) for b in res])
    np.savez_compressed(os.path.join(HERE, "mu_k64.npz"), **out)
    print("mu_k64 ok:", {k: v.tolist() for k, v in out.items() if k.startswith("err_")})


def golden_sdr_k64(n_ant=64):
    """SDR snapshots at the size BASELINE config 4 names (main_beampatterns_plotting/sdr_mp.py:124-151 on the reference classes):
    K = 64 LOS MRT, IBO 0 and 3 dB, one snapshot each; receiver positions and bits are stored, the channel is rebuilt from them."""
    my_mod, my_dist, my_tx, my_rx, my_array = mg.build_objects(n_ant)
    rng = np.random.default_rng(29)
    out = dict(sdr_ibo=np.array([0.0, 3.0]))
    des, dis, poss, bits_l = [], [], [], []
    for ibo in out["sdr_ibo"]:
        my_rx.set_position(212.0 + rng.uniform(-5, 5), 212.0 + rng.uniform(-5, 5), 1.5)
        ch = mg.make_channel("los", my_array, my_rx)
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
        des.append(np.sum(np.abs(np.sum(ak * cl_sc, axis=0)) ** 2))
        dis.append(np.sum(np.abs(np.sum(rx_sc - ak * cl_sc, axis=0)) ** 2))
        poss.append((my_rx.cord_x, my_rx.cord_y, my_rx.cord_z))
        bits_l.append(bits.astype(np.uint8))
    out.update(sdr_pos=np.array(poss), sdr_bits=np.stack(bits_l), sdr_desired=np.array(des), sdr_distortion=np.array(dis), n_ant=n_ant)
    np.savez_compressed(os.path.join(HERE, "sdr_k64.npz"), **out)
    print("sdr_k64 ok: SDR [dB]", 10 * np.log10(out["sdr_desired"] / out["sdr_distortion"]))


if __name__ == "__main__":
    what = sys.argv[1:] or ["mcnc", "cnc", "mu", "sdr"]
    if "sdr" in what:
        golden_sdr_k64()
    if "mu" in what:
        golden_mu_k64()
    if "cnc" in what:
        mg.golden_link("link_k64_los_cnc16", 64, "los", 0.0, 15.0, 16, [0, 1, 2, 5], [1, 2, 3], is_mcnc=False, full_frames=2)
    if "chan" in what:
        mg.golden_link("link_k64_two_path_mcnc8", 64, "two_path", 0.0, 15.0, 8, [0, 1, 2, 3, 4, 5], [11, 12, 13], full_frames=2)
        mg.golden_link("link_k64_rayleigh_mcnc8", 64, "rayleigh", 0.0, 15.0, 8, [0, 1, 2, 3, 4, 5], [14, 15, 16], full_frames=2)
    if "mcnc" in what:
        mg.golden_link("link_k64_los_mcnc32", 64, "los", 0.0, 15.0, 32, [0, 1, 2, 3, 4, 5], [1, 2, 3], full_frames=3)
