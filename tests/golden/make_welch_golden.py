"""Welch PSD golden vectors (build container only): time-domain signals produced by the UNMODIFIED reference the way
main_multiuser/main_multiuser_mrt_precoding_radiation_pattern.py:355-381 collects them (clean and distortion components of the
received signal of a K = 8 LOS MRT array over several OFDM symbols, utilities.to_time_domain, concatenated) and
scipy.signal.welch of them with the three (nfft, nperseg) pairs the reference's scripts use.  The signals are stored as
complex64 (what the device path receives); scipy runs on exactly those values.

    python tests/golden/make_welch_golden.py
"""
import os
import sys

import numpy as np
from scipy.signal import welch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402

mod, dist, tx, rx, arr = mg.build_objects(8, ibo=0.0)
ch = mg.make_channel("los", arr, rx)
h = ch.get_channel_mat_fd()
arr.set_precoding_matrix(channel_mat_fd=h, mr_precoding=True)
arr.update_distortion(ibo_db=0.0, avg_sample_pow=mod.avg_sample_power)
rng = np.random.default_rng(77)
cln, dst = [], []
for _ in range(6):
    bits = rng.choice((0, 1), mod.n_bits_per_ofdm_sym)
    x_fd, x_clean = arr.transmit(bits, out_domain_fd=True, return_both=True)
    rx_fd = ch.propagate(in_sig_mat=x_fd)
    rx_clean = ch.propagate(in_sig_mat=x_clean)
    cln.append(mg.utilities.to_time_domain(rx_clean))
    dst.append(mg.utilities.to_time_domain(rx_fd - rx_clean))
out = {"clean": np.concatenate(cln).ravel().astype(np.complex64), "distortion": np.concatenate(dst).ravel().astype(np.complex64)}
for name in ("clean", "distortion"):
    x = out[name].astype(np.complex128)
    for nfft, nperseg in ((4096, 1024), (4096, 2048), (1024, 256), (128, 64)):
        f, p = welch(x, fs=nfft, nfft=nfft, nperseg=nperseg, return_onesided=False)
        out["psd_%s_%d_%d" % (name, nfft, nperseg)] = p
        out["freq_%d" % nfft] = f
np.savez_compressed(os.path.join(HERE, "welch.npz"), **out)
print({k: v.shape for k, v in out.items()})
