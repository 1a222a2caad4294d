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
np.savez_compressed(os.path.join(HERE, "ref_curves.npz"), **out)
