#!/usr/bin/env python
"""One two-user batch of 2072 realisations (4144 user-frames), K = 64, I = 5, MU-MRT then ZF: launch list for ncu."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine  # noqa: E402

eng = LinkEngine(0)
eng.configure(n_ant=64, channel="los", receiver="mcnc", pa="softlim", ibo_db=0.0)
for tag in ("mrt", "zf"):
    print(eng.mu_simulate(2072, 5, 22.8, precoding=tag, seed=11)["err"][:, -1])
eng.close()
