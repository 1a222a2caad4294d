"""Drop-in for the reference's distortion.py: memoryless power-amplifier models.

Classes keep the reference's names (control flow elsewhere uses isinstance on them), constructor
arguments, attributes (`ibo_db`, `sat_pow`, `avg_samp_pow`, `p_hardness`, `toi_db`, `cubic_dist_coeff`),
setters and `__str__` tags; `process` runs the PA kernel of libmmimo_b200.so."""
import numpy as np
from numpy import ndarray

from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import _stage_abi as _abi


def _run_pa(desc, in_sig):
    x = np.asarray(in_sig)
    out = _abi.get_ops().pa_process(desc, x.astype(np.complex128, copy=False))
    out = out.reshape(x.shape)
    if not np.iscomplexobj(x):  # real-valued amplitude sweeps (plot_characteristics) stay real
        return out.real
    return out


class SoftLimiter:
    """Ideal soft limiter (distortion.py:22-98): unit gain up to |x|^2 = sat_pow, constant envelope above."""

    def __init__(self, ibo_db: float, avg_samp_pow: float):
        self.ibo_db = ibo_db
        self.avg_samp_pow = avg_samp_pow
        self.sat_pow = np.power(10, ibo_db / 10) * avg_samp_pow

    def __str__(self):
        return "softlim"

    def set_ibo(self, ibo_db: float) -> None:
        self.ibo_db = ibo_db
        self.sat_pow = np.power(10, ibo_db / 10) * self.avg_samp_pow

    def set_avg_sample_power(self, avg_samp_pow: float) -> None:
        self.avg_samp_pow = avg_samp_pow
        self.sat_pow = np.power(10, self.ibo_db / 10) * avg_samp_pow

    def pa_desc(self):
        return _abi.pa_desc(_abi.PA_SOFT_LIMITER, self.sat_pow)

    def process(self, in_sig: ndarray) -> ndarray:
        return _run_pa(self.pa_desc(), in_sig)

    def plot_characteristics(self, in_ampl_min: float = -10, in_ampl_max: float = 10, step: float = 0.1) -> None:
        _plot_transfer(self, "Soft limiter transfer characteristic", "IBO [dB]", self.ibo_db, in_ampl_min, in_ampl_max, step)


class Rapp:
    """Rapp solid-state PA model (distortion.py:116-198)."""

    def __init__(self, ibo_db: float, avg_samp_pow: float, p_hardness: float):
        self.p_hardness = p_hardness
        self.ibo_db = ibo_db
        self.avg_samp_pow = avg_samp_pow
        self.sat_pow = np.power(10, ibo_db / 10) * avg_samp_pow

    def __str__(self):
        return "rapp"

    def set_hardness(self, p_hardness: float) -> None:
        self.p_hardness = p_hardness

    def set_ibo(self, ibo_db: float) -> None:
        self.ibo_db = ibo_db
        self.sat_pow = np.power(10, ibo_db / 10) * self.avg_samp_pow

    def set_avg_sample_power(self, avg_samp_pow: float) -> None:
        # like the reference (distortion.py:157-165) this does NOT store avg_samp_pow: a later set_ibo
        # keeps using the value given at construction
        self.sat_pow = np.power(10, self.ibo_db / 10) * avg_samp_pow

    def pa_desc(self):
        return _abi.pa_desc(_abi.PA_RAPP, self.sat_pow, self.p_hardness)

    def process(self, in_sig: ndarray) -> ndarray:
        return _run_pa(self.pa_desc(), in_sig)

    def plot_characteristics(self, in_ampl_min: float = -10, in_ampl_max: float = 10, step: float = 0.1) -> None:
        _plot_transfer(self, "Rapp model transfer characteristic", "IBO [dB]", self.ibo_db, in_ampl_min, in_ampl_max, step)


class ThirdOrderNonLin:
    """Third-order polynomial PA model (distortion.py:215-282): y = x - c3 x |x|^2."""

    def __init__(self, toi_db: float, avg_samp_pow: float):
        self.avg_samp_pow = avg_samp_pow
        self.toi_db = toi_db
        self.cubic_dist_coeff = 1 / (np.power(10, (toi_db / 10))) / avg_samp_pow

    def __str__(self):
        return "toi"

    def set_toi(self, toi_db: float) -> None:
        self.toi_db = toi_db
        self.cubic_dist_coeff = 1 / (np.power(10, (toi_db / 10))) / self.avg_samp_pow

    def set_avg_sample_power(self, avg_samp_pow: float) -> None:
        self.avg_samp_pow = avg_samp_pow
        self.cubic_dist_coeff = 1 / (np.power(10, (self.toi_db / 10))) / avg_samp_pow

    def pa_desc(self):
        return _abi.pa_desc(_abi.PA_TOI, self.cubic_dist_coeff)

    def process(self, in_sig):
        return _run_pa(self.pa_desc(), in_sig)

    def plot_characteristics(self, in_ampl_min: float = -10, in_ampl_max: float = 10, step: float = 0.1) -> None:
        _plot_transfer(self, "Third order distortion transfer characteristic", "TOI [dB]", self.toi_db, in_ampl_min,
                       in_ampl_max, step)


def pa_desc_of(impairment):
    """mmimo_pa_desc of any impairment object (None -> no PA)."""
    if impairment is None:
        return None
    return impairment.pa_desc()


def _plot_transfer(model, title, legend_title, label, lo, hi, step):
    import matplotlib.pyplot as plt
    amp = np.arange(lo, hi + step, step)
    fig, ax = plt.subplots(1, 1)
    ax.plot(amp, model.process(amp), label=label)
    ax.set_title(title)
    ax.set_xlabel("Input signal amplitude [V]")
    ax.set_ylabel("Output signal amplitude [V]")
    ax.legend(title=legend_title)
    ax.grid()
    plt.tight_layout()
    plt.show()
