"""Golden vectors for the soft-output demodulator: runs the UNMODIFIED reference (modulation.soft_decoding through
OfdmQamModem.soft_detection_llr, modulation.py:29-59, :407-417) and stores inputs + LLRs in soft_llr.npz.
  python tests/golden/make_soft_llr.py        (needs /root/reference)"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle.refharness import import_reference  # noqa: E402

R = import_reference()
modulation = R["modulation"]
rng = np.random.default_rng(20261104)
out = {}
for M in (4, 16, 64):
    mod = modulation.OfdmQamModem(constel_size=M, n_fft=64, n_sub_carr=32, cp_len=4)
    n = 96
    bits = rng.choice((0, 1), n * int(np.log2(M)))
    sym = modulation.modulate(mod.constellation, mod.n_bits_per_symbol, bits)
    for tag, nv in (("lo", 0.05 * mod.avg_symbol_power), ("hi", 0.8 * mod.avg_symbol_power), ("tiny", 2e-4 * mod.avg_symbol_power)):
        y = sym + (rng.standard_normal(n) + 1j * rng.standard_normal(n)) * np.sqrt(nv / 2)
        y = y.astype(np.complex64).astype(np.complex128)   # exactly representable in the c64 ABI
        out["y_%d_%s" % (M, tag)] = y
        out["nv_%d_%s" % (M, tag)] = np.float64(nv)
        out["llr_%d_%s" % (M, tag)] = mod.soft_detection_llr(y, noise_var=nv)
    # per-symbol noise variance through the module-level function
    nvv = rng.uniform(0.05, 1.0, n) * mod.avg_symbol_power
    out["nvv_%d" % M] = nvv
    out["llrv_%d" % M] = modulation.soft_decoding(mod.constellation, mod.n_bits_per_symbol, out["y_%d_hi" % M], nvv)
np.savez_compressed(os.path.join(HERE, "soft_llr.npz"), **out)
print({k: (v.shape, int(np.isinf(v).sum())) for k, v in out.items() if k.startswith("llr")})
