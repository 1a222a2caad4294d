#!/usr/bin/env python
"""One two-user batch (K = 64, I = 5, MU-MRT, 1184 user-frames): the launch `ncu --set full` captures for k_link_frame_t64<PA, MU=true>."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine  # noqa: E402

eng = LinkEngine(0)
eng.configure(n_ant=64, channel="los", receiver="mcnc", pa="softlim", ibo_db=0.0)
print(eng.mu_simulate(592, 5, 22.8, precoding="mrt", seed=11)["err"])
eng.close()
