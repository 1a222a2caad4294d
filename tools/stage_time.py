import os, sys, json
sys.path.insert(0, "/root/repo")
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine
eng = LinkEngine(0)
for r in range(2):
    ms = eng.measure_stage_kernels(16384, 5)
    print(os.environ.get("MMIMO_LIB", "default"), {k: round(b * 16384 / (ms[k] * 1e-3) * 1e-9) for k, b in zip(eng.STAGE_KERNELS, eng.STAGE_BYTES_PER_ROW)}, flush=True)
