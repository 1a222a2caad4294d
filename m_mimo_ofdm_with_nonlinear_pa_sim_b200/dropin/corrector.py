"""Drop-in for the reference's corrector.py: iterative clipping-noise-cancellation receivers.

`CncReceiver`, `McncReceiver`, `CncMuReceiver`, `McncMuReceiver` keep their constructor side effects (the
objects handed in are kept BY REFERENCE and mutated, e.g. McncReceiver.__init__ sets MRT precoding on the
caller's array, corrector.py:135), the externally assigned `agc_corr_vec` and the `receive` signatures.
The whole iteration loop of one `receive` call runs on the GPU (mmimo_cnc_receive / mmimo_mcnc_receive);
the reference's regeneration after the last iteration, whose result it discards, is not computed."""
import numpy as np
from numpy import ndarray

import distortion
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import _stage_abi as _abi


def _iters(n_iters_lst):
    lst = [int(v) for v in np.asarray(n_iters_lst).reshape(-1)]
    return lst, int(max(lst))


def _collect(n_iters_lst, n_max, per_iter):
    """The reference appends once per loop index that is `in n_iters_lst` (corrector.py:80-84): ascending order."""
    wanted = set(n_iters_lst)
    return [per_iter[i] for i in range(n_max + 1) if i in wanted]


class _CncBase:
    def __init__(self, modem, impairment):
        self.modem = modem
        self.impairment = impairment
        if isinstance(self.impairment, distortion.SoftLimiter):
            self.modem.alpha = self.modem.calc_alpha(self.impairment.ibo_db)
        else:
            self.modem.alpha = 1.0
        self.upsample_factor = self.modem.n_fft / self.modem.n_sub_carr
        self.impairment.set_avg_sample_power(avg_samp_pow=self.modem.avg_symbol_power * (1 / self.upsample_factor))

    def update_distortion(self, ibo_db: float, alpha_val: float = 1.0) -> None:
        if isinstance(self.impairment, distortion.ThirdOrderNonLin):
            self.impairment.set_toi(ibo_db)
            self.modem.alpha = alpha_val
        else:
            self.impairment.set_ibo(ibo_db)
            self.modem.alpha = self.modem.calc_alpha(ibo_db)

    def _run(self, n_iters_lst, in_sig_fd, alpha_estimate, other_usr_symbols, return_bits):
        lst, n_max = _iters(n_iters_lst)
        alpha = self.modem.alpha if alpha_estimate is None else alpha_estimate
        bits, corr = _abi.get_ops().cnc_receive(self.modem._modem_desc(), self.impairment.pa_desc(), float(alpha), n_max,
                                                in_sig_fd, other_usr_symbols=other_usr_symbols, want_bits=return_bits,
                                                want_corr=not return_bits)
        return _collect(lst, n_max, bits if return_bits else corr)


class CncReceiver(_CncBase):
    """Clipping noise cancellation with a single-antenna model of the transmitter (corrector.py:12-112)."""

    def receive(self, n_iters_lst: list, in_sig_fd: ndarray, alpha_estimate: float = None, return_bits: bool = True) -> list:
        return self._run(n_iters_lst, in_sig_fd, alpha_estimate, None, return_bits)


class CncMuReceiver(_CncBase):
    """Two-user CNC that knows the other user's symbols (corrector.py:248-345)."""

    def receive(self, n_iters_lst: list, in_sig_fd: ndarray, other_usr_symbols: ndarray, alpha_estimate: float = None) -> list:
        return self._run(n_iters_lst, in_sig_fd, alpha_estimate, np.asarray(other_usr_symbols), True)


class _McncBase:
    def _agc_vec(self, alpha_estimate, use_estimate, usr_idx):
        """sum_k alpha_k h_k v_k on the n_fft grid, ones off the data sub-carriers (corrector.py:136-163, :366-403)."""
        arr = self.antenna_array
        e0 = arr.array_elements[0]
        v = arr.get_precoding_mat()
        res = _abi.get_ops().agc(self.channel.get_channel_mat_fd(), v, usr_idx if v.ndim == 3 else 0,
                                 getattr(e0.impairment, "ibo_db", 0.0), e0.modem.n_sub_carr,
                                 alpha_fixed=alpha_estimate if use_estimate else None)
        return res["agc_nfft"]

    def _receive(self, n_iters_lst, in_sig_fd, usr_idx, other_usr_bits, return_bits):
        lst, n_max = _iters(n_iters_lst)
        arr = self.antenna_array
        uni = arr._uniform()
        if uni is None:
            raise NotImplementedError("MCNC needs array elements that share modem and PA settings")
        md, pa = uni
        v = arr._precoder_or_none()
        if v is None:
            v = np.ones((arr.n_elements, arr.array_elements[0].modem.n_users, md.n_sub_carr), dtype=np.complex128)
        bits, corr = _abi.get_ops().mcnc_receive(md, pa, self.channel.get_channel_mat_fd(), v, self.agc_corr_vec, n_max, in_sig_fd,
                                                 usr_idx=usr_idx, other_usr_bits=other_usr_bits, want_bits=return_bits,
                                                 want_corr=not return_bits)
        return _collect(lst, n_max, bits if return_bits else corr)


class McncReceiver(_McncBase):
    """Multi-antenna CNC: regenerates the full K-antenna transmission and channel each iteration (corrector.py:115-245)."""

    def __init__(self, antenna_array, channel, alpha_estimate: float = None):
        self.antenna_array = antenna_array
        self.channel = channel
        self.antenna_array.set_precoding_matrix(channel_mat_fd=self.channel.get_channel_mat_fd(), mr_precoding=True)
        self.n_sub_carr = self.antenna_array.array_elements[0].modem.n_sub_carr
        is_toi = isinstance(self.antenna_array.array_elements[0].impairment, distortion.ThirdOrderNonLin)
        self.agc_corr_vec = self._agc_vec(alpha_estimate, is_toi, 0)

    def receive(self, n_iters_lst: list, in_sig_fd: ndarray, return_bits: bool = True) -> list:
        return self._receive(n_iters_lst, in_sig_fd, 0, None, return_bits)

    def update_agc(self, alpha_estimate: float = None) -> None:
        self.agc_corr_vec = self._agc_vec(alpha_estimate, alpha_estimate is not None, 0)


class McncMuReceiver(_McncBase):
    """MCNC for one user of a multi-user transmission, given the other user's bits (corrector.py:348-489)."""

    def __init__(self, antenna_array, channel, usr_idx: int, alpha_estimate: float = None):
        self.antenna_array = antenna_array
        self.channel = channel
        self.usr_idx = usr_idx
        self.antenna_array.set_precoding_matrix(channel_mat_fd=self.channel.get_channel_mat_fd(), mr_precoding=True)
        self.n_sub_carr = self.antenna_array.array_elements[0].modem.n_sub_carr
        is_toi = isinstance(self.antenna_array.array_elements[0].impairment, distortion.ThirdOrderNonLin)
        self.agc_corr_vec = self._agc_vec(alpha_estimate, is_toi, usr_idx)

    def receive(self, n_iters_lst: list, in_sig_fd: ndarray, other_usr_bits: ndarray) -> list:
        return self._receive(n_iters_lst, in_sig_fd, self.usr_idx, np.asarray(other_usr_bits), True)

    def update_agc(self, alpha_estimate: float = None) -> None:
        self.agc_corr_vec = self._agc_vec(alpha_estimate, alpha_estimate is not None, self.usr_idx)
