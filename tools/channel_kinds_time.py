#!/usr/bin/env python
"""Per-kernel time of one 4144-frame batch for every channel kind the fused path generates (BASELINE configs[2]):
channel kernel / realisation kernel / frame kernel, CUDA events on the library's stream (LinkEngine.profile)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine  # noqa: E402

eng = LinkEngine(0)
F, I = 4144, 5
for chan, eps in (("los", 0.0), ("two_path", 0.0), ("rayleigh", 0.0), ("los", 0.2)):
    eng.configure(n_ant=64, channel=chan, receiver="mcnc", pa="softlim", ibo_db=0.0)
    eng.simulate(F, I, 22.8, seed=1, csi_eps=eps)
    eng.profile(True)
    t0 = time.perf_counter()
    for r in range(3):
        out = eng.simulate(F, I, 22.8, seed=2 + r, csi_eps=eps)
    dt = (time.perf_counter() - t0) / 3
    ms, n = eng.profile_read()
    eng.profile(False)
    print("%-9s eps %.1f  channel %.3f ms  prep %.3f ms  frame %.3f ms  call %.3f ms  %.1f k frames/s  ber5 %.5f"
          % (chan, eps, ms[0] / 3, ms[1] / 3, ms[2] / 3, dt * 1e3, F / dt / 1e3, out["err"][-1] / (F * 12288.0)), flush=True)
eng.close()
