"""Throughput of the fused two-user path (BASELINE config 5: 64 antennas, two users, MU-MRT / ZF, MCNC 5 iterations)
with device-generated channels, bits and noise.  Prints user-frames/s (one realisation = one frame per user)."""
import sys
import time

import numpy as np

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine  # noqa: E402

eng = LinkEngine(0)
eng.configure(n_ant=64, channel="los", pa="softlim", ibo_db=0.0, receiver="mcnc")
snr = 15.0 + 10 * np.log10(6.0)
n = 2048
for tag in ("mrt", "zf"):
    for su in (False, True):
        eng.mu_simulate(n, 5, snr, precoding=tag, loop_single_user=su, seed=1)
        t0 = time.perf_counter()
        reps = 3
        for r in range(reps):
            out = eng.mu_simulate(n, 5, snr, precoding=tag, loop_single_user=su, seed=2 + r)
        dt = (time.perf_counter() - t0) / reps
        ber = out["err"].astype(float) / out["bits"][:, None]
        print("%s %-22s %8.0f user-frames/s   BER user0 %s" % (tag, "single-user loop" if su else "McncMuReceiver loop", 2 * n / dt,
                                                              np.array2string(ber[0], precision=4)))
eng.close()
