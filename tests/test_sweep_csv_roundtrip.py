"""The CSV files tools/ber_sweep.py writes must be readable by the reference's plotting scripts: written with
ber_sweep.write_csv (= utilities.save_to_csv, utilities.py:342-352), read back with the reader logic of
utilities.read_from_csv (:355-365) and indexed exactly like final_plots/ber_vs_ebn0.py:38-40,55-58 (row 0 the axis, row 1 no
distortion, row idx + 1 for iteration cnc_n_iter_lst[idx]) and final_plots/fixed_ber_ebno_vs_ibo.py:46-57 (IBO-major grid)."""
import os
import sys

import numpy as np
from scipy import interpolate

from conftest import ROOT, load_golden

sys.path.insert(0, os.path.join(ROOT, "tools"))
import ber_sweep  # noqa: E402


def test_ber_vs_ebn0_layout_round_trip(tmp_path):
    ref = load_golden("ref_curves")["mcnc_los"]          # a file the reference shipped: axis + clean + iterations 0..8
    path = str(tmp_path / "ber_vs_ebn0.csv")
    ber_sweep.write_csv(path, ber_sweep.csv_rows("ebn0", ref[0], ref[1:]))
    data_lst = ber_sweep.read_csv(path)
    # final_plots/ber_vs_ebn0.py:38-40
    ebn0_arr, ber_per_dist = data_lst[0], data_lst[1:]
    assert np.array_equal(ebn0_arr, ref[0])
    cnc_n_iter_lst = [0, 1, 2, 3, 4, 5, 6, 7, 8]
    assert len(ber_per_dist) == len(cnc_n_iter_lst) + 1
    for idx, it in enumerate(cnc_n_iter_lst):            # :55-58: ber_per_dist[idx + 1] is iteration cnc_n_iter_lst[idx]
        assert np.array_equal(ber_per_dist[idx + 1], ref[2 + it])
    assert np.array_equal(ber_per_dist[0], ref[1])       # no distortion
    assert all(isinstance(v, float) for v in ebn0_arr)   # QUOTE_NONNUMERIC: floats, not strings


def test_fixed_ber_layout_round_trip(tmp_path):
    g = load_golden("ref_curves")
    ibo, ebn0, grid = g["fixed_ber_mcnc_los_ibo"], g["fixed_ber_mcnc_los_ebn0"], g["fixed_ber_mcnc_los_grid"]
    path = str(tmp_path / "fixed_ber.csv")
    ber_sweep.write_csv(path, ber_sweep.csv_rows("fixed-ber", ibo, None, grid))
    data_lst = ber_sweep.read_csv(path)
    # final_plots/fixed_ber_ebno_vs_ibo.py:46-57
    ibo_arr, tmp_data = data_lst[0], data_lst[1:]
    n_it = grid.shape[2]
    ber = np.zeros((len(ibo_arr), len(ebn0), n_it))
    for ibo_idx in range(len(ibo_arr)):
        for snr_idx in range(len(ebn0)):
            for ite_idx in range(n_it):
                ber[ibo_idx, snr_idx, ite_idx] = tmp_data[ibo_idx * len(ebn0) + snr_idx][ite_idx]
    assert np.array_equal(ber, grid) and np.array_equal(ibo_arr, ibo)
    # :63-92 required Eb/N0 by interpolation: finite where the target is crossed
    req = float(interpolate.interp1d(ber[0, :, 5], ebn0, kind="linear")(1e-2))
    assert 14.0 < req < 16.0


def test_nant_chan_layout_round_trip(tmp_path):
    ref = load_golden("ref_curves")["ber_vs_nant_mcnc"]  # 31 rows: axis + {los, two_path, rayleigh} x (clean + 9 iterations)
    assert ref.shape == (31, 8)
    path = str(tmp_path / "ber_vs_nant.csv")
    ber_sweep.write_csv(path, ber_sweep.csv_rows("nant-chan", ref[0], ref[1:]))
    data_lst = ber_sweep.read_csv(path)
    assert np.array_equal(data_lst[0], [1, 2, 4, 8, 16, 32, 64, 128])
    assert np.array_equal(np.array(data_lst[1:]), ref[1:])
