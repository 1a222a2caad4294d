"""GPU, device RNG (Philox): BER curves of the fused kernels against the Monte-Carlo curves the reference SHIPS
(figs/csv_results, collected in tests/golden/ref_curves.npz by make_ref_curves.py; generated there with 1e7 bits per
point).  Statistical comparison: both sides are Monte-Carlo estimates with per-frame (receiver-position) variation, so
rows are compared with a relative + absolute band, only where the reference BER is large enough to be resolved."""
import argparse
import os
import sys

import numpy as np
import pytest

from conftest import ROOT, load_golden

pytestmark = pytest.mark.gpu
sys.path.insert(0, os.path.join(ROOT, "tools"))


def sweep(**kw):
    import ber_sweep
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine
    a = argparse.Namespace(sweep="ebn0", ebn0="15", ibo="0", n_ant=64, constel=64, channel="los", receiver="mcnc", pa="softlim",
                           n_iter=8, csi_eps=0.0, bits_max=3e7, err_min=1e9, frames_per_step=1024, seed=2137)
    for k, v in kw.items():
        setattr(a, k, v)
    eng = LinkEngine(0)
    eng.configure(constel_size=64, n_ant=64, pa="softlim", ibo_db=ber_sweep.axis(a.ibo)[0], channel=a.channel, receiver=a.receiver)
    try:
        return ber_sweep.run_sweep(eng, a)
    finally:
        eng.close()


def check(rows, ref_rows, rel=0.12, floor=1e-3):
    n = 0
    for got, want in zip(rows, ref_rows):
        if want < floor:
            assert got < 2.5 * floor, (got, want)
            continue
        assert abs(got - want) <= rel * want + 2e-4, (got, want, rows, ref_rows)
        n += 1
    return n


@pytest.mark.parametrize("key,channel,receiver", [("mcnc_los", "los", "mcnc"), ("mcnc_two_path", "two_path", "mcnc"),
                                                  ("mcnc_rayleigh", "rayleigh", "mcnc"), ("cnc_los", "los", "cnc")])
def test_ber_vs_ebn0_matches_reference_curves(key, channel, receiver):
    ref = load_golden("ref_curves")[key]          # rows: axis, clean, I = 0..8
    ax, rows, frames = sweep(ebn0="10:18:4", channel=channel, receiver=receiver)
    checked = 0
    for c, e in enumerate(ax):
        j = int(np.argmin(np.abs(ref[0] - e)))
        assert ref[0][j] == e
        checked += check(rows[:, c], ref[1:, j])
    assert checked >= 15


def test_ber_vs_ibo_matches_reference_curve():
    ref = load_golden("ref_curves")["mcnc_los_vs_ibo_ebn0_15"]   # rows: axis, clean, I = 0..7
    ax, rows, frames = sweep(sweep="ibo", ibo="-3:3:3", ebn0="15", n_iter=7)
    for c, ibo in enumerate(ax):
        j = int(np.argmin(np.abs(ref[0] - ibo)))
        check(rows[:, c], ref[1:, j])
