"""Drop-in for the reference's transceiver.py: one antenna front-end (modem + PA model + position)."""
import numpy as np
from numpy import ndarray

import distortion
from modulation import OfdmQamModem
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import _stage_abi as _abi


class Transceiver:
    """Radio transceiver (transceiver.py:9-184): same constructor, attributes and `transmit` flag matrix."""

    def __init__(self, modem: OfdmQamModem, center_freq: int, carrier_spacing: int, impairment, cord_x: float = 0,
                 cord_y: float = 0, cord_z: float = 0):
        self.modem = modem
        self.impairment = impairment
        self.center_freq = center_freq
        self.carrier_spacing = carrier_spacing
        if isinstance(impairment, distortion.SoftLimiter):
            self.modem.update_alpha(ibo_db=impairment.ibo_db)
        self.tx_ant_gain_db = 0
        self.rx_ant_gain_db = 0
        self.tx_power_dbm = 10 * np.log10(1000 * self.modem.avg_sample_power)
        self.cord_x = cord_x
        self.cord_y = cord_y
        self.cord_z = cord_z

    def set_position(self, cord_x: float, cord_y: float, cord_z: float) -> None:
        self.cord_x = cord_x
        self.cord_y = cord_y
        self.cord_z = cord_z

    def set_ant_gains(self, tx_ant_gain_db: float, rx_ant_gain_db: float) -> None:
        self.tx_ant_gain_db = tx_ant_gain_db
        self.rx_ant_gain_db = rx_ant_gain_db

    def set_tx_power_dbm(self, tx_power_dbm: float) -> None:
        self.tx_power_dbm = tx_power_dbm

    def correct_constellation(self) -> None:
        self.modem.correct_constellation(ibo_db=self.impairment.ibo_db)

    def update_distortion(self, ibo_db: float) -> None:
        self.impairment.set_ibo(ibo_db=ibo_db)
        self.modem.update_alpha(ibo_db=ibo_db)

    def transmit(self, in_bits: ndarray, out_domain_fd: bool = True, skip_dist: bool = False, return_both: bool = False,
                 sum_usr_signals: bool = True):
        """modulate -> precode -> IFFT (+CP) -> PA -> (FFT) for this antenna (transceiver.py:98-174).

        Return shapes follow the reference: an array, a (distorted, clean) tuple with `return_both`, or per-user
        lists with `sum_usr_signals=False`."""
        m = self.modem
        bits = m._check_bits(in_bits)
        v = None
        if m.precoding_mat is not None:
            v = np.asarray(m.precoding_mat, dtype=np.complex128).reshape(1, m.n_users, m.n_sub_carr)
        clean_only = skip_dist or self.impairment is None
        both = return_both and not clean_only
        per_user = (m.n_users > 1) and not sum_usr_signals
        res = _abi.get_ops().array_transmit(m._modem_desc(), distortion.pa_desc_of(self.impairment), 1, m.n_users, bits, v,
                                            skip_dist=clean_only, time_domain=not out_domain_fd, per_user=per_user,
                                            return_both=both)
        if both:
            out, clean = res
            if sum_usr_signals or m.n_users == 1:
                return out[0, 0], clean[0, 0]
            return [[out[u, 0], clean[u, 0]] for u in range(m.n_users)]
        if sum_usr_signals or m.n_users == 1:
            return res[0, 0]
        return [res[u, 0] for u in range(m.n_users)]

    def receive(self, in_symb: ndarray):
        """Demodulate and hard-decide a time-domain OFDM symbol (transceiver.py:176-184)."""
        return self.modem.demodulate(in_symb)
