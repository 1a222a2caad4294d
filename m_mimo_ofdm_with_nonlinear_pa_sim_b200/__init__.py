"""B200-native (sm_100a) Monte-Carlo link hot path of m-mimo-ofdm-with-nonlinear-pa-sim.

`LinkEngine` drives the fused CUDA kernels through the C ABI in include/mmimo_b200.h;
`dropin/` holds modules with the reference's names (modulation, transceiver, ...) for unchanged scripts.
There is no CPU fallback: importing the engine without libmmimo_b200.so raises."""
from ._lib import MmimoError, lib_path  # noqa: F401
from .engine import LinkEngine, subcarrier_freqs, ula_positions  # noqa: F401

__all__ = ["LinkEngine", "MmimoError", "lib_path", "subcarrier_freqs", "ula_positions"]
