"""Collect a few of the reference's SHIPPED Monte-Carlo result curves (figs/csv_results/*.csv: data, not code) into
tests/golden/ref_curves.npz as statistical parity targets.  Row layout (main_mp_miso_mcnc_ber_vs_ebn0.py:175-179):
row 0 = swept axis, row 1 = no distortion, row 2 = standard RX (0 iterations), rows 3.. = iterations 1..8.
Generated with bits_sent_max = 1e7, n_err_min = 1e6 (scripts' literals).  Run in the build container only."""
import csv
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/figs/csv_results"
FILES = {
    "mcnc_los": "ber_vs_ebn0_mcnc_los_nant64_ibo0_ebn0_min5_max20_step0.50_niter1_2_3_4_5_6_7_8.csv",
    "mcnc_two_path": "ber_vs_ebn0_mcnc_two_path_nant64_ibo0_ebn0_min5_max20_step0.50_niter1_2_3_4_5_6_7_8.csv",
    "mcnc_rayleigh": "ber_vs_ebn0_mcnc_rayleigh_nant64_ibo0_ebn0_min5_max20_step0.50_niter1_2_3_4_5_6_7_8.csv",
    "cnc_los": "ber_vs_ebn0_cnc_los_nant64_ibo0_ebn0_min5_max20_step0.50_niter1_2_3_4_5_6_7_8.csv",
    "mcnc_los_vs_ibo_ebn0_15": "ber_vs_ibo_mcnc_los_nant64_ebn0_15_ibo_min-9_max9_step0.50_niter1_2_3_4_5_6_7_8.csv",
}
# ragged files: fixed-BER grid (row 0 = 16 IBO values, then one 9-value row per (IBO, Eb/N0) pair, IBO-major,
# main_mp_miso_mcnc_constant_ber_req_ebn0_vs_ibo.py:210-215) -> stored as axis + [n_ibo, n_ebn0, n_iter] grid
RAGGED = {
    "fixed_ber_mcnc_los": ("fixed_ber1.0e-02_mcnc_los_nant64_ebn0_min10_max22_step0.50_ibo_min0_max7_step0.50_niter1_2_3_4_5_6_7_8.csv",
                           np.arange(10.0, 22.1, 0.5)),
}
for _eps in (0.1, 0.2, 0.4):  # main_mp_miso_mcnc_csi_err_ber_vs_ebn0.py (BASELINE config 5, CSI-error sweep)
    FILES["mcnc_los_csi_eps%.1f" % _eps] = ("ber_vs_ebn0_mcnc_los_csi_eps%.3f_nant64_ibo0_ebn0_min5_max20_step0.50_niter1_2_3_4_5_6_7_8.csv" % _eps)
for _tag in ("mr", "zf"):  # main_multiuser/main_multiuser_cnc_ber_vs_ebn0.py: two users, CncReceiver; axis + 2 x (clean + 9 iterations)
    FILES["mu_%s_cnc" % _tag] = ("ber_vs_ebn0_mu_%s_cnc_los_nant64_ibo0_ebn0_min5_max20_step1.00_niter1_2_3_4_5_6_7_8_angles-30_30_"
                                 "distances100_316.3.csv" % _tag)
FILES["ber_vs_nant_mcnc"] = "ber_vs_nant_mcnc_nant1_2_4_8_16_32_64_128_ebn0_15_ibo0_niter1_2_3_4_5_6_7_8.csv"
out = {}
for key, (fn, ebn0) in RAGGED.items():
    with open(os.path.join(SRC, fn), newline="") as fh:
        rows = [[float(v) for v in r] for r in csv.reader(fh) if r]
    ibo = np.array(rows[0])
    grid = np.array(rows[1:]).reshape(len(ibo), len(ebn0), -1)
    out[key + "_ibo"], out[key + "_ebn0"], out[key + "_grid"] = ibo, ebn0, grid
    print(key, grid.shape)
for key, fn in FILES.items():
    with open(os.path.join(SRC, fn), newline="") as fh:
        rows = [[float(v) for v in r] for r in csv.reader(fh) if r]
    out[key] = np.array(rows, dtype=np.float64)
    print(key, out[key].shape)
# two-user MU-MRT run of main_multiuser/main_mp_mulituser_mcnc_ber_vs_ebn0.py (12 worker processes of
# multiuser_mp_prototype.multiuser_mcnc_parallel): every worker wrote its raw ERROR and BIT counts - row 0 the Eb/N0 axis, then
# per Eb/N0 point: errors user 0, bits user 0, errors user 1, bits user 1 (10 slots: no distortion, iterations 0..8;
# multiuser_mp_prototype.py:680-682).  Summed over the workers -> exact counts behind the reference's curve.
err = bits = None
for proc in range(12):
    fn = ("proc_%d_ber_vs_ebn0_mu_mr_mcnc_los_nant64_ibo0_ebn0_min5_max20_step1.00_niter1_2_3_4_5_6_7_8_angles-30_30_"
          "distances100_316.3.csv" % proc)
    with open(os.path.join(SRC, fn), newline="") as fh:
        rows = [[float(v) for v in r] for r in csv.reader(fh) if r]
    ax = np.array(rows[0])
    body = np.array(rows[1:]).reshape(len(ax), 2, 2, 10)      # [ebn0][user][err|bits][slot]
    err = body[:, :, 0] if err is None else err + body[:, :, 0]
    bits = body[:, :, 1] if bits is None else bits + body[:, :, 1]
out["mu_mr_mcnc_ebn0"], out["mu_mr_mcnc_err"], out["mu_mr_mcnc_bits"] = ax, err, bits
print("mu_mr_mcnc", err.shape, (err / np.maximum(bits, 1))[10])
np.savez_compressed(os.path.join(HERE, "ref_curves.npz"), **out)
