"""Drop-in for the reference's noise.py: additive white Gaussian noise.

The draw comes from the object's public numpy Generator `rng_gen` (mp_model re-assigns it, and the reference's
results depend on its stream), exactly as noise.py:66 does; scaling and adding run on the GPU."""
from abc import ABC

import numpy as np
from numpy import ndarray

from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import _stage_abi as _abi


class Noise(ABC):
    def __init__(self, snr_db: float, noise_p_dbm: float, seed: int):
        self.snr_db = snr_db
        self.noise_p_dbm = noise_p_dbm
        if seed is not None:
            self.seed = seed
            self.rng_gen = np.random.default_rng(seed)
        else:
            self.rng_gen = np.random.default_rng(1234)


class Awgn(Noise):
    """AWGN source (noise.py:34-83)."""

    def __init__(self, snr_db=None, noise_p_dbm=None, seed=None):
        super().__init__(snr_db, noise_p_dbm, seed)

    def process(self, in_sig: ndarray, avg_sample_pow: float = 1, fixed_noise_power: bool = False,
                disp_data: bool = False) -> ndarray:
        """in_sig + n, E|n|^2 = avg_sample_pow / SNR (or the fixed noise power in dBm)."""
        in_sig = np.asarray(in_sig)
        n_sampl = len(in_sig)
        if fixed_noise_power:
            noise_std = np.sqrt(2 * 0.001 * 10 ** (self.noise_p_dbm / 10))
        else:
            noise_std = np.sqrt(2 * avg_sample_pow / (10 ** (self.snr_db / 10)))
        unit = self.rng_gen.standard_normal((n_sampl, 2)).view(np.complex128)[:, 0]
        scale = float(np.real(noise_std)) * 0.5
        if disp_data:
            n_nsc = 2048
            sig_nsc = np.concatenate((in_sig[-n_nsc // 2:], in_sig[1:(n_nsc // 2) + 1]))
            nz_nsc = np.concatenate((unit[-n_nsc // 2:], unit[1:(n_nsc // 2) + 1])) * scale
            p_sig, p_nz = np.sum(np.abs(sig_nsc) ** 2), np.sum(np.abs(nz_nsc) ** 2)
            print("Signal power:[dBm]", 10 * np.log10(p_sig + 30))
            print("Noise power:[dBm]", 10 * np.log10(p_nz) + 30)
            print("SNR: ", 10 * np.log10(p_sig / p_nz))
        return _abi.get_ops().awgn_add(in_sig, unit, scale)
