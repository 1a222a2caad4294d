"""Drop-in for the reference's channel.py: frequency-domain MISO channels.

Same class names, `__str__` tags (used in result file names), the public `channel_mat_fd` attribute (scripts and
mp_model assign it directly) and the `calc_channel_mat / reroll_channel_coeffs / propagate / get_channel_mat_fd /
set_channel_mat_fd` methods.  Channel matrices [K, n_fft] and the propagation multiply-and-sum are computed by
the CUDA stage kernels; random draws stay on the object's numpy Generator (`rng_gen`) so the streams - and with
them every later draw of a script - are the reference's."""
import numpy as np
from numpy import ndarray

from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import _stage_abi as _abi


def _freqs(rx_transceiver) -> ndarray:
    """Sub-carrier frequencies as channel.py:51-52 computes them: torch.fft.fftfreq is float32 and the product
    with the (integer) spacing and centre frequency stays float32, so the values sit on the float32 grid."""
    n_fft = rx_transceiver.modem.n_fft
    k = np.fft.fftfreq(n_fft, d=1.0 / n_fft).astype(np.float32)
    return (k * np.float32(rx_transceiver.carrier_spacing) + np.float32(rx_transceiver.center_freq)).astype(np.float64)


def _geometry(tx_transceivers, rx_transceiver):
    pos = np.array([[t.cord_x, t.cord_y, t.cord_z] for t in tx_transceivers], dtype=np.float64)
    gains = np.array([t.tx_ant_gain_db for t in tx_transceivers], dtype=np.float64)
    rx = np.array([rx_transceiver.cord_x, rx_transceiver.cord_y, rx_transceiver.cord_z], dtype=np.float64)
    return pos, gains, rx


class _MisoFdBase:
    def get_channel_mat_fd(self) -> ndarray:
        return self.channel_mat_fd

    def propagate(self, in_sig_mat: ndarray, sum_signals: bool = True) -> ndarray:
        """Y = X .* H, summed over antennas unless sum_signals=False (channel.py:74-89)."""
        return _abi.get_ops().propagate(np.asarray(in_sig_mat), self.channel_mat_fd, sum_signals=sum_signals)


class MisoLosFd(_MisoFdBase):
    """Line-of-sight channel (channel.py:11-89)."""

    def __init__(self):
        self.channel_mat_fd = None

    def __str__(self):
        return "los"

    def calc_channel_mat(self, tx_transceivers: list, rx_transceiver, skip_attenuation: bool = False) -> None:
        pos, gains, rx = _geometry(tx_transceivers, rx_transceiver)
        self.channel_mat_fd = _abi.get_ops().channel_geometric(_abi.CHAN_LOS, pos, gains, rx_transceiver.rx_ant_gain_db, rx,
                                                               _freqs(rx_transceiver), skip_attenuation)


class MisoTwoPathFd(_MisoFdBase):
    """LOS + ground reflection with coefficient -1 (channel.py:92-184)."""

    def __init__(self):
        self.channel_mat_fd = None

    def __str__(self):
        return "two_path"

    def calc_channel_mat(self, tx_transceivers: list, rx_transceiver, skip_attenuation: bool = False) -> None:
        pos, gains, rx = _geometry(tx_transceivers, rx_transceiver)
        self.channel_mat_fd = _abi.get_ops().channel_geometric(_abi.CHAN_TWO_PATH, pos, gains, rx_transceiver.rx_ant_gain_db,
                                                               rx, _freqs(rx_transceiver), skip_attenuation)


class MisoRayleighFd(_MisoFdBase):
    """IID Rayleigh coefficients times the free-space loss at construction-time geometry (channel.py:187-292)."""

    def __init__(self, tx_transceivers: list, rx_transceiver, seed: int = None):
        self.n_inputs = len(tx_transceivers)
        self.fd_samp_size = tx_transceivers[0].modem.n_fft
        self.fd_att_mat = None
        self.los_fd_att_mat = None
        if seed is not None:
            self.seed = seed
            self.rng_gen = np.random.default_rng(seed)
        else:
            self.rng_gen = np.random.default_rng(1234)
        self._geom = _geometry(tx_transceivers, rx_transceiver) + (rx_transceiver.rx_ant_gain_db, _freqs(rx_transceiver))
        pos, gains, rx, rx_gain, freqs = self._geom
        self.los_fd_att_mat = _abi.get_ops().channel_fspl(pos, gains, rx_gain, rx, freqs).real
        self.set_channel_mat_fd()

    def __str__(self):
        return "rayleigh"

    def _draw(self, skip_attenuation):
        coeffs = self.rng_gen.standard_normal(size=(self.n_inputs, self.fd_samp_size * 2)).view(dtype=np.complex128) / np.sqrt(2.0)
        if skip_attenuation:
            return coeffs
        pos, gains, rx, rx_gain, freqs = self._geom
        return _abi.get_ops().channel_fspl(pos, gains, rx_gain, rx, freqs, coeffs=coeffs)

    def set_channel_mat_fd(self, channel_mat_fd: ndarray = None, skip_attenuation: bool = False) -> None:
        self.channel_mat_fd = self._draw(skip_attenuation) if channel_mat_fd is None else channel_mat_fd

    def reroll_channel_coeffs(self, skip_attenuation: bool = False) -> None:
        self.channel_mat_fd = self._draw(skip_attenuation)


class MisoRandomPathsFd(_MisoFdBase):
    """Sum of n_paths random plane waves (channel.py:295-401), evaluated exactly as written there."""

    def __init__(self, tx_transceivers: list, rx_transceiver, seed: int = None, n_paths: int = 10,
                 max_delay_spread: float = 1000e-9):
        self.n_inputs = len(tx_transceivers)
        self.fd_samp_size = tx_transceivers[0].modem.n_fft
        self.n_paths = n_paths
        self.max_delay_spread = max_delay_spread
        if seed is not None:
            self.seed = seed
            self.rng_gen = np.random.default_rng(seed)
        else:
            self.rng_gen = np.random.default_rng(1234)
        self._roll(tx_transceivers, _freqs(rx_transceiver))

    def __str__(self):
        return "random_paths"

    def _roll(self, tx_transceivers, freqs):
        angles = self.rng_gen.uniform(low=-np.pi / 2.0, high=np.pi / 2, size=self.n_paths)
        taus = self.rng_gen.uniform(low=0.0, high=self.max_delay_spread, size=self.n_paths)
        pos = np.array([[t.cord_x, t.cord_y, t.cord_z] for t in tx_transceivers], dtype=np.float64)
        self.channel_mat_fd = _abi.get_ops().channel_random_paths(pos, freqs, angles, taus)

    def reroll_channel_coeffs(self, tx_transceivers: list) -> None:
        self._roll(tx_transceivers, _freqs(tx_transceivers[0]))


class MisoQuadrigaFd(_MisoFdBase):
    """QuaDRiGa channel through the MATLAB engine (channel.py:404-494).  The channel coefficients come from MATLAB;
    only `propagate` is numeric work of this package.  Needs `matlab.engine` and QuaDRiGa, like the reference."""

    def __init__(self, tx_transceivers: list, rx_transceiver, channel_model_str: str, start_matlab_eng: bool = True):
        self.tx_transceivers = tx_transceivers
        self.rx_transceiver = rx_transceiver
        self.n_ant_val = len(tx_transceivers)
        self.n_fft = tx_transceivers[0].modem.n_fft
        self.subcarr_spacing = tx_transceivers[0].carrier_spacing
        self.center_freq = tx_transceivers[0].center_freq
        self.channel_model_str = channel_model_str
        self.distance = np.sqrt(rx_transceiver.cord_x ** 2 + rx_transceiver.cord_y ** 2)
        if start_matlab_eng:
            import matlab.engine
            self.meng = matlab.engine.start_matlab()
            self.meng.addpath(r"main_quadriga_channel")
            self.meng.rng(5)
            self.meng.qd_channel_env_setup(self.meng.double(self.n_ant_val), self.meng.double(self.n_fft),
                                           self.meng.double(self.subcarr_spacing), self.meng.double(self.center_freq),
                                           self.meng.double(self.distance), channel_model_str, nargout=0)
            self.channel_mat_fd = np.array(
                self.meng.qd_get_channel_mat(rx_transceiver.cord_x, rx_transceiver.cord_y, rx_transceiver.cord_z))

    def __str__(self):
        return "quadriga"

    def calc_channel_mat(self, tx_transceivers: list, rx_transceiver) -> None:
        self.channel_mat_fd = np.array(
            self.meng.qd_get_channel_mat(rx_transceiver.cord_x, rx_transceiver.cord_y, rx_transceiver.cord_z))

    def reroll_channel_coeffs(self, tx_transceivers: list, rx_transceiver) -> None:
        self.calc_channel_mat(tx_transceivers, rx_transceiver)
