#!/usr/bin/env python
"""Small end-to-end case for compute-sanitizer (memcheck / racecheck): fused link (device RNG + injected inputs, clean
run, CSI error) and every stage entry point once.  Usage: compute-sanitizer --tool memcheck python tools/sanitize_case.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine, dropin  # noqa: E402

eng = LinkEngine(0)
for chan, rx in (("los", "mcnc"), ("rayleigh", "cnc"), ("two_path", "mcnc")):
    eng.configure(n_ant=3, channel=chan, receiver=rx, ibo_db=1.0)
    out = eng.simulate(3, 2, 20.0, seed=5, incl_clean_run=True, csi_eps=0.1 if chan == "los" else 0.0,
                       dumps=("det", "err_per_frame", "rx_eq", "corr", "h_true", "v", "agc", "pa_p0", "alpha", "clean_det"))
    print(chan, rx, out["err"].tolist(), out["clean_err"].tolist())
# two-user fused path (both precoders, both receiver loops, all dumps), multi-point batch, SDR snapshots
eng.configure(n_ant=3, channel="los", receiver="mcnc", ibo_db=1.0)
for tag in ("mrt", "zf"):
    for su in (False, True):
        out = eng.mu_simulate(3, 2, 20.0, precoding=tag, loop_single_user=su, seed=9, dumps=("det", "rx_eq", "v", "agc", "pa_p0"))
        out = eng.mu_simulate(3, 2, 20.0, precoding=tag, receiver="cnc_mu" if su else "cnc", seed=9)
        print("mu", tag, su, out["err"].tolist())
eng.configure(n_ant=3, channel="two_path", receiver="mcnc", ibo_db=1.0)
print("sweep", eng.sweep([15.0, 20.0, 25.0], [0.0, 2.0, 4.0], 3, 2, seed=4).tolist())
print("sdr", [v.tolist() for v in eng.sdr(3, seed=2)])
eng.close()

dropin.install()
import copy  # noqa: E402

import antenna_array, channel, corrector, distortion, modulation, noise, transceiver, utilities  # noqa: E402,E401

mod = modulation.OfdmQamModem(64, 512, 256, 32, n_users=2)
pa = distortion.Rapp(1.0, mod.avg_sample_power, 3.0)
tx = transceiver.Transceiver(copy.deepcopy(mod), int(3.5e9), int(15e3), copy.deepcopy(pa))
rxa = transceiver.Transceiver(copy.deepcopy(mod), int(3.5e9), int(15e3), copy.deepcopy(pa), cord_x=100, cord_y=40, cord_z=1.5)
rxb = transceiver.Transceiver(copy.deepcopy(mod), int(3.5e9), int(15e3), copy.deepcopy(pa), cord_x=-60, cord_y=90, cord_z=1.5)
arr = antenna_array.LinearArray(5, tx, int(3.5e9), 0.5, cord_z=15)
cha, chb = channel.MisoLosFd(), channel.MisoTwoPathFd()
cha.calc_channel_mat(arr.array_elements, rxa)
chb.calc_channel_mat(arr.array_elements, rxb)
ray = channel.MisoRayleighFd(arr.array_elements, rxa, seed=1)
rp = channel.MisoRandomPathsFd(arr.array_elements, rxa, seed=2, n_paths=4)
for kw in (dict(mr_precoding=True), dict(zf_precoding=True), dict(), dict(mr_precoding=True, sep_carr_per_usr=True)):
    arr.set_precoding_matrix([cha.channel_mat_fd, chb.channel_mat_fd], **kw)
arr.set_precoding_matrix([cha.channel_mat_fd, chb.channel_mat_fd], mr_precoding=True)
arr.update_distortion(1.0, mod.avg_sample_power)
bits = np.random.default_rng(0).choice((0, 1), (2, mod.n_bits_per_ofdm_sym))
x, xc = arr.transmit(bits, return_both=True)
xt = arr.transmit(bits, out_domain_fd=False, sum_usr_signals=False)
y = noise.Awgn(snr_db=20, seed=3).process(chb.propagate(x), avg_sample_pow=1e-9)
rcv = corrector.McncMuReceiver(copy.deepcopy(arr), copy.deepcopy(chb), usr_idx=1)
rcv.antenna_array.set_precoding_matrix([cha.channel_mat_fd, chb.channel_mat_fd], mr_precoding=True)
rcv.update_agc()
res = rcv.receive([0, 1, 2], y / rcv.agc_corr_vec, other_usr_bits=bits[0])
cmu = corrector.CncMuReceiver(copy.deepcopy(mod), copy.deepcopy(pa))
res2 = cmu.receive([0, 2], y, other_usr_symbols=mod.modulate(bits, get_symbols_only=True)[0])
llr = mod.soft_detection_llr(mod.modulate(bits, get_symbols_only=True)[0] + 0.1, noise_var=0.5)
td = utilities.to_time_domain(x)
fd = utilities.to_freq_domain(td, remove_cp=False)
f_psd, psd = utilities.welch(td.ravel(), fs=512, nfft=512, nperseg=128)
print("stage ops ok", float(psd.sum()), utilities.count_mismatched_bits(res[0], bits[1]), len(res2), float(np.abs(fd - x).max()))
