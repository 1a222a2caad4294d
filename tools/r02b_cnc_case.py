#!/usr/bin/env python
"""One 4144-frame CncReceiver batch (K = 64, I = 5): the launch `ncu --set full` captures for k_cnc_loop_t64."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine  # noqa: E402

eng = LinkEngine(0)
eng.configure(n_ant=64, channel="los", receiver="cnc", pa="softlim", ibo_db=0.0)
print(eng.simulate(4144, 5, 22.8, seed=1)["err"])
eng.close()
