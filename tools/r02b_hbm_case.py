#!/usr/bin/env python
"""One 4144-frame batch per channel kind + one pass over the unfused stage kernels: the launches `ncu` reads DRAM bytes and
durations from for the HBM-bound kernels (profiles/r02b_hbm_kernels.csv)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine  # noqa: E402

eng = LinkEngine(0)
for chan, eps in (("los", 0.0), ("two_path", 0.0), ("rayleigh", 0.0), ("los", 0.2)):
    eng.configure(n_ant=64, channel=chan, receiver="mcnc", pa="softlim", ibo_db=0.0)
    eng.simulate(4144, 0, 22.8, seed=1, csi_eps=eps)
print(eng.measure_stage_kernels(16384, 1))
eng.close()
