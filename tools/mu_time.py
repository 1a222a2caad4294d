#!/usr/bin/env python
"""Two-user frame kernel: ms per launch and user-frames/s (K = 64, I = 5, 4144 user-frames), MU-MRT and ZF, all four receivers."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine  # noqa: E402

eng = LinkEngine(0)
eng.configure(n_ant=64, channel="los", receiver="mcnc", pa="softlim", ibo_db=0.0)
for tag, kw in (("mrt", {}), ("zf", {}), ("mrt", dict(loop_single_user=True)), ("mrt", dict(receiver="cnc")), ("zf", dict(receiver="cnc_mu"))):
    eng.mu_simulate(2072, 5, 22.8, precoding=tag, seed=11, **kw)
    eng.profile(True)
    t0 = time.perf_counter()
    for r in range(3):
        out = eng.mu_simulate(2072, 5, 22.8, precoding=tag, seed=11 + r, **kw)
    dt = (time.perf_counter() - t0) / 3
    ms, n = eng.profile_read()
    eng.profile(False)
    print(os.environ.get("MMIMO_LIB", "default"), tag, kw, "kernel %.3f ms  %.1f k user-frames/s  err %s" % (ms[2] / max(int(n[2]), 1), 4144 / dt / 1e3, out["err"].ravel()[:6].tolist()), flush=True)
eng.close()
