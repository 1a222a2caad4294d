#!/usr/bin/env python
"""Time of one Monte-Carlo frame driven block by block through the drop-in classes, the way the single-process reference
scripts do it (main_clipping_noise_cancellation/main_miso_mcnc_ber_vs_ebn0.py:184-244): re-drawn position -> channel ->
MRT precoding -> update_distortion -> update_agc -> AGC vectors -> transmit -> propagate -> noise -> receive (MCNC 0..5)."""
import copy
import cProfile
import os
import pstats
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import dropin  # noqa: E402

dropin.install()
import antenna_array, channel, corrector, distortion, modulation, noise, transceiver, utilities  # noqa: E402,E401

K = int(sys.argv[1]) if len(sys.argv) > 1 else 64
mod = modulation.OfdmQamModem(constel_size=64, n_fft=4096, n_sub_carr=2048, cp_len=128)
dist = distortion.SoftLimiter(0, mod.avg_sample_power)
tx = transceiver.Transceiver(modem=copy.deepcopy(mod), impairment=copy.deepcopy(dist), center_freq=int(3.5e9), carrier_spacing=int(15e3))
rx = transceiver.Transceiver(modem=copy.deepcopy(mod), impairment=copy.deepcopy(dist), cord_x=212, cord_y=212, cord_z=1.5,
                             center_freq=int(3.5e9), carrier_spacing=int(15e3))
arr = antenna_array.LinearArray(n_elements=K, base_transceiver=tx, center_freq=int(3.5e9), wav_len_spacing=0.5, cord_x=0, cord_y=0, cord_z=15)
ch = channel.MisoLosFd()
ch.calc_channel_mat(tx_transceivers=arr.array_elements, rx_transceiver=rx, skip_attenuation=False)
nz = noise.Awgn(snr_db=22.8, seed=1)
rcv = corrector.McncReceiver(copy.deepcopy(arr), copy.deepcopy(ch))
rng = np.random.default_rng(2)
iters = [0, 1, 2, 3, 4, 5]


def frame():
    rx.set_position(cord_x=212 + rng.uniform(-5, 5), cord_y=212 + rng.uniform(-5, 5), cord_z=1.5)
    ch.calc_channel_mat(tx_transceivers=arr.array_elements, rx_transceiver=rx, skip_attenuation=False)
    arr.set_precoding_matrix(channel_mat_fd=ch.channel_mat_fd, mr_precoding=True)
    arr.update_distortion(ibo_db=0, avg_sample_pow=mod.avg_sample_power)
    rcv.channel.set_channel_mat_fd(ch.channel_mat_fd) if hasattr(rcv.channel, "set_channel_mat_fd") else None
    rcv.antenna_array.set_precoding_matrix(channel_mat_fd=ch.channel_mat_fd, mr_precoding=True)
    rcv.antenna_array.update_distortion(ibo_db=0, avg_sample_pow=mod.avg_sample_power)
    rcv.update_agc()
    bits = rng.choice((0, 1), mod.n_bits_per_ofdm_sym)
    x = arr.transmit(bits, out_domain_fd=True)
    y = ch.propagate(in_sig_mat=x)
    y = nz.process(y, avg_sample_pow=mod.avg_symbol_power * 1e-9)
    y = y / rcv.agc_corr_vec
    out = rcv.receive(n_iters_lst=iters, in_sig_fd=y)
    return [utilities.count_mismatched_bits(bits, b) for b in out]


print("errors", frame())
t0 = time.perf_counter()
for _ in range(20):
    frame()
dt = (time.perf_counter() - t0) / 20
print("K = %d: %.2f ms per frame through the building blocks (%.0f frames/s per process)" % (K, dt * 1e3, 1 / dt))
pr = cProfile.Profile()
pr.enable()
for _ in range(10):
    frame()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
