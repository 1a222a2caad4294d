"""Drop-in for the reference's antenna_array.py: K transceivers + geometry, array transmit and precoding.

`array_elements` stays a list of K independent `Transceiver` objects (scripts read and write
`array_elements[i].modem.precoding_mat`, `.impairment.ibo_db`, `.cord_x` ...).  Per-element precoding rows are
views into ONE [K, (U,) N_sc] matrix, and `transmit` hands the whole array to a single batched CUDA call when the
elements share modem and PA settings (they do unless a script edits one element by hand; then it falls back to
one call per element, still on the GPU)."""
import copy
from abc import ABC

import numpy as np
import scipy.constants
from numpy import ndarray

import distortion
import utilities
from transceiver import Transceiver
from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import _stage_abi as _abi


def _pa_key(imp):
    if imp is None:
        return None
    d = imp.pa_desc()
    return (d.kind, d.p0, d.p1)


class AntennaArray(ABC):
    """Base antenna array (antenna_array.py:13-411)."""

    def __init__(self, n_elements: int, base_transceiver: Transceiver, center_freq: int,
                 wav_len_spacing: float = 0.5, cord_x: float = 0, cord_y: float = 0, cord_z: float = 0):
        self.n_elements = n_elements
        self.n_users = base_transceiver.modem.n_users
        self.base_transceiver = base_transceiver
        self.center_freq = center_freq
        self.wav_len_spacing = wav_len_spacing
        self.array_elements = []
        self.cord_x = cord_x
        self.cord_y = cord_y
        self.cord_z = cord_z

    def _add_element(self, x, y, z):
        tx = copy.deepcopy(self.base_transceiver)
        tx.cord_x, tx.cord_y, tx.cord_z = x, y, z
        self.array_elements.append(tx)

    def set_tx_power_lvls(self, tx_power_dbm: float, total: bool = False) -> None:
        for tx in self.array_elements:
            if total:
                tx.set_tx_power_dbm(10 * (np.log10(10 ** (tx_power_dbm / 10) / len(self.array_elements))))
            else:
                tx.set_tx_power_dbm(tx_power_dbm)

    # ------------------------------------------------------------------------------------------
    def _uniform(self):
        """(modem desc, pa desc) when every element shares them, else None."""
        e0 = self.array_elements[0]
        m0 = e0.modem
        key = (m0.constel_size, m0.n_fft, m0.n_sub_carr, m0.cp_len, m0.n_users, _pa_key(e0.impairment))
        scale0 = None
        for e in self.array_elements:
            m = e.modem
            if (m.constel_size, m.n_fft, m.n_sub_carr, m.cp_len, m.n_users, _pa_key(e.impairment)) != key:
                return None
            s = m._constellation[0]
            if scale0 is None:
                scale0 = s
            elif s != scale0:
                return None
        return m0._modem_desc(), distortion.pa_desc_of(e0.impairment)

    def _precoder_or_none(self):
        if all(e.modem.precoding_mat is None for e in self.array_elements):
            return None
        m = self.base_transceiver.modem
        n_usr = self.array_elements[0].modem.n_users
        v = np.ones((self.n_elements, n_usr, m.n_sub_carr), dtype=np.complex128)
        for k, e in enumerate(self.array_elements):
            if e.modem.precoding_mat is not None:
                v[k] = np.asarray(e.modem.precoding_mat)  # [n_sc] rows broadcast over users like np.multiply does
        return v

    def transmit(self, in_bits: ndarray, out_domain_fd: bool = True, return_both: bool = False, skip_dist: bool = False,
                 sum_usr_signals: bool = True):
        """Array transmit (antenna_array.py:58-140): every element modulates the same bits with its own precoding row.

        Returns [K, n] (squeezed like the reference, so K = 1 gives a vector), [U, K, n] with sum_usr_signals=False,
        and a (distorted, clean) pair with return_both."""
        if sum_usr_signals is False and self.n_users == 1:
            sum_usr_signals = True
        uni = self._uniform()
        if uni is None:
            return self._transmit_per_element(in_bits, out_domain_fd, return_both, skip_dist, sum_usr_signals)
        md, pa = uni
        n_usr = self.array_elements[0].modem.n_users
        bits = self.array_elements[0].modem._check_bits(in_bits)
        clean_only = skip_dist or pa is None
        both = return_both and not clean_only
        res = _abi.get_ops().array_transmit(md, pa, self.n_elements, n_usr, bits, self._precoder_or_none(),
                                            skip_dist=clean_only, time_domain=not out_domain_fd,
                                            per_user=not sum_usr_signals, return_both=both)
        if return_both:
            if not both:
                raise ValueError("too many values to unpack: return_both needs an impairment and skip_dist=False")
            out, clean = res
            if sum_usr_signals:
                return np.squeeze(out[0]), np.squeeze(clean[0])
            return out, clean
        if sum_usr_signals:
            return np.squeeze(res[0])
        return res

    def _transmit_per_element(self, in_bits, out_domain_fd, return_both, skip_dist, sum_usr_signals):
        outs, cleans = [], []
        for tx in self.array_elements:
            r = tx.transmit(in_bits, out_domain_fd=out_domain_fd, return_both=return_both, skip_dist=skip_dist,
                            sum_usr_signals=sum_usr_signals)
            if return_both:
                if sum_usr_signals:
                    outs.append(r[0]); cleans.append(r[1])
                else:
                    outs.append([p[0] for p in r]); cleans.append([p[1] for p in r])
            else:
                outs.append(r)
        out = np.asarray(outs)
        if not sum_usr_signals:
            out = np.swapaxes(out, 0, 1)
        if return_both:
            clean = np.asarray(cleans)
            if not sum_usr_signals:
                return out, np.swapaxes(clean, 0, 1)
            return np.squeeze(out), np.squeeze(clean)
        return np.squeeze(out) if sum_usr_signals else out

    # ------------------------------------------------------------------------------------------
    def set_precoding_matrix(self, channel_mat_fd=None, mr_precoding: bool = False, zf_precoding: bool = False,
                             update_distortion: bool = False, sep_carr_per_usr: bool = False) -> None:
        """Precoder from the channel matrix / list of per-user matrices (antenna_array.py:142-311):
        MRT, phase-only, multi-user MRT / ZF / phase-only, or separate sub-carrier blocks per user.
        Row k of the result becomes `array_elements[k].modem.precoding_mat` (a view, not a copy)."""
        n_sc = self.base_transceiver.modem.n_sub_carr
        ops = _abi.get_ops()
        if not isinstance(channel_mat_fd, list):
            mode = _abi.PREC_MRT if mr_precoding is True else _abi.PREC_PHASE
            v = ops.precoding(mode, np.asarray(channel_mat_fd)[None], n_sc)[:, 0, :]  # [K, n_sc]
            for k, tx in enumerate(self.array_elements):
                tx.modem.set_precoding(v[k, :])
        elif not sep_carr_per_usr:
            mode = _abi.PREC_MRT if mr_precoding else (_abi.PREC_ZF if zf_precoding else _abi.PREC_PHASE)
            v = ops.precoding(mode, np.stack([np.asarray(h) for h in channel_mat_fd]), n_sc)  # [K, U, n_sc]
            for k, tx in enumerate(self.array_elements):
                tx.modem.set_precoding(v[k, :, :])
        else:
            mode = _abi.PREC_MRT if mr_precoding is True else _abi.PREC_PHASE
            v = ops.precoding(mode, np.stack([np.asarray(h) for h in channel_mat_fd]), n_sc, sep_carriers=True)
            for k, tx in enumerate(self.array_elements):
                tx.modem.set_precoding(v[k, :])
        if update_distortion:
            self.update_distortion(ibo_db=self.array_elements[0].impairment.ibo_db,
                                   avg_sample_pow=self.array_elements[0].modem.avg_sample_power)

    def _precoding_power_per_antenna(self, default_one: bool):
        """[K, n_sc] |v|^2 summed over users; elements without a precoder count as 1 (update_distortion) ."""
        n_sc = self.base_transceiver.modem.n_sub_carr
        pw = np.ones((self.n_elements, n_sc), dtype=np.float64)
        for k, tx in enumerate(self.array_elements):
            p = tx.modem.precoding_mat
            if p is None:
                if not default_one:
                    raise TypeError("precoding matrix is not set")
                continue
            p = np.asarray(p)
            a = p.real * p.real + p.imag * p.imag  # |v|^2 without the square root of np.abs
            pw[k, :] = a if a.ndim == 1 else np.sum(a, axis=0)
        return pw

    def update_distortion(self, ibo_db: float, avg_sample_pow: float, alpha_val: float = None) -> None:
        """Set IBO / TOI, Bussgang alpha and the precoding-aware average sample power of every element
        (antenna_array.py:313-360)."""
        avg_precoding_gain = np.average(self._precoding_power_per_antenna(default_one=True))
        alpha = None  # calc_alpha depends on ibo_db only: evaluated once, not once per antenna
        for tx in self.array_elements:
            if isinstance(tx.impairment, distortion.ThirdOrderNonLin):
                tx.modem.alpha = alpha_val
                tx.impairment.set_toi(ibo_db)
            else:
                if alpha is None:
                    alpha = tx.modem.calc_alpha(ibo_db=ibo_db)
                tx.modem.alpha = alpha
                tx.impairment.set_ibo(ibo_db)
            tx.impairment.set_avg_sample_power(avg_sample_pow * avg_precoding_gain)

    def get_avg_precoding_gain(self) -> float:
        return np.average(self._precoding_power_per_antenna(default_one=False))

    def get_precoding_mat(self) -> ndarray:
        """[K, n_sc], or [K, U, n_sc] with several users (antenna_array.py:384-401)."""
        n_sc = self.base_transceiver.modem.n_sub_carr
        if self.n_users == 1:
            out = np.empty((self.n_elements, n_sc), dtype=np.complex128)
        else:
            out = np.empty((self.n_elements, self.n_users, n_sc), dtype=np.complex128)
        for k, tx in enumerate(self.array_elements):
            out[k] = tx.modem.precoding_mat
        return out

    def update_n_users(self, n_users) -> None:
        self.n_users = n_users
        for tx in self.array_elements:
            tx.modem.n_users = n_users


class LinearArray(AntennaArray):
    """Uniform linear array along X centred on the origin (antenna_array.py:415-445)."""

    def __init__(self, n_elements: int, base_transceiver: Transceiver, center_freq: int, wav_len_spacing: float,
                 cord_x: float = 0, cord_y: float = 0, cord_z: float = 0):
        super().__init__(n_elements, base_transceiver, center_freq, wav_len_spacing, cord_x, cord_y, cord_z)
        lam = scipy.constants.c / self.center_freq
        half = (self.n_elements - 1) * self.wav_len_spacing * lam / 2
        for x in np.linspace(-half, half, self.n_elements):
            self._add_element(x, 0, self.cord_z)


class CircularArray(AntennaArray):
    """Elements on a semi-circle (antenna_array.py:448-479).  The reference constructor calls
    `utilities.pts_on_semicircum(r=..., n=...)`, keyword names that function does not have, so it raises TypeError
    there; this version passes the arguments positionally."""

    def __init__(self, n_elements: int, base_transceiver: Transceiver, center_freq: int, wav_len_spacing: float,
                 cord_x: float = 0, cord_y: float = 0, cord_z: float = 0):
        super().__init__(n_elements, base_transceiver, center_freq, wav_len_spacing, cord_x, cord_y, cord_z)
        lam = scipy.constants.c / self.center_freq
        radius = lam * (self.n_elements - 1) / (2 * np.pi)
        pts = utilities.pts_on_semicircum(radius, self.n_elements)
        for idx in range(self.n_elements):
            self._add_element(pts[idx][0], pts[idx][1], self.cord_z)


class PlanarRectangularArray(AntennaArray):
    """Uniform rectangular array in the X-Z plane (antenna_array.py:482-520)."""

    def __init__(self, n_elements_per_row: int, n_elements_per_col: int, base_transceiver: Transceiver, center_freq: int,
                 wav_len_spacing: float, cord_x: float = 0, cord_y: float = 0, cord_z: float = 0):
        super().__init__(n_elements_per_row * n_elements_per_col, base_transceiver, center_freq, wav_len_spacing, cord_x,
                         cord_y, cord_z)
        lam = scipy.constants.c / self.center_freq
        hc = (n_elements_per_col - 1) * self.wav_len_spacing * lam / 2
        hr = (n_elements_per_row - 1) * self.wav_len_spacing * lam / 2
        for col in np.linspace(-hc, hc, n_elements_per_col):
            for row in np.linspace(-hr, hr, n_elements_per_row):
                self._add_element(col, 0, self.cord_z + row)
