"""Drop-in for the reference's modulation.py: QAM mapper / slicer and the OFDM modem.

Same class names (`Modem`, `QamModem`, `OfdmQamModem`), attributes and method signatures as the reference
(modulation.py:91-424).  Bit mapping, precoding, IFFT/FFT, cyclic prefix and hard decisions run in the CUDA
stage kernels of libmmimo_b200.so; the 2^m-entry constellation table itself is built with numpy at
construction (it is an attribute scripts read).  Only Gray-mapped square QAM has kernels - exactly what
`QamModem` builds - so a `Modem` with a hand-made constellation raises on numeric calls."""
import numpy as np
from numpy import ndarray

from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import _stage_abi as _abi


def _square_qam_table(constel_size: int) -> ndarray:
    """Gray-reordered square QAM table (modulation.py:230-244 base table, :110-114 reorder)."""
    side = int(round(np.sqrt(constel_size)))
    pam = np.arange(-side + 1, side, 2, dtype=np.float64)
    n = np.arange(constel_size)
    col, row = n // side, n % side
    imag = np.where(col % 2 == 0, pam[row], pam[side - 1 - row])
    base = pam[col] + 1j * imag
    return base[np.argsort(n ^ (n >> 1), kind="stable")]


def _table_scale(constellation) -> float:
    """Scale s with constellation == s * (canonical Gray square QAM); raises if it is not such a table."""
    c = np.asarray(constellation).reshape(-1)
    size = c.size
    side = int(round(np.sqrt(size)))
    if side * side != size or size not in (4, 16, 64, 256):
        raise NotImplementedError("libmmimo_b200 has kernels for Gray-mapped square 4/16/64/256-QAM only")
    ref = _square_qam_table(size)
    s = (c[0] / ref[0])
    if abs(s.imag) > 1e-12 * abs(s) or s.real <= 0 or not np.allclose(c, s.real * ref, rtol=1e-9, atol=0):
        raise NotImplementedError("libmmimo_b200 has kernels for Gray-mapped square QAM tables only")
    return float(s.real)


def _desc(constellation, n_fft=8, n_sub_carr=2, cp_len=0):
    c = np.asarray(constellation)
    return _abi.modem_desc(c.size, n_fft, n_sub_carr, cp_len, _table_scale(c))


# ---- module-level functions of the reference (modulation.py:13-88) -------------------------------
def modulate(constellation: ndarray, n_bits_per_symbol: int, input_bits: ndarray) -> ndarray:
    """Bits -> complex baseband symbols (modulation.py:13-25)."""
    bits = np.asarray(input_bits).reshape(-1)
    if bits.size % n_bits_per_symbol:
        raise ValueError("number of bits must be a multiple of n_bits_per_symbol")
    return _abi.get_ops().qam_map(_desc(constellation), bits)


def demodulate(constellation: ndarray, n_bits_per_symbol: int, input_symbols: ndarray, soft: bool = False,
               noise_var: float = 0.0) -> ndarray:
    """Symbols -> hard-decision bits (modulation.py:63-88)."""
    if soft:
        return soft_decoding(constellation, n_bits_per_symbol, np.asarray(input_symbols),
                             noise_var if isinstance(noise_var, np.ndarray) else np.repeat(noise_var, len(input_symbols)))
    scalar = np.ndim(input_symbols) == 0
    _, bits = _abi.get_ops().qam_detect(_desc(constellation), np.atleast_1d(input_symbols), want_symbols=False)
    return bits if not scalar else bits[:n_bits_per_symbol]


def soft_decoding(constellation: ndarray, n_bits_per_symbol: int, input_symbols: ndarray, noise_var_vec: ndarray) -> ndarray:
    """Exact LLR demodulation (modulation.py:29-59) on the device: mmimo_qam_soft_llr."""
    return _abi.get_ops().qam_soft_llr(_desc(constellation), np.atleast_1d(input_symbols), noise_var_vec)


# ---- classes --------------------------------------------------------------------------------------
class Modem:
    """Digital modem (modulation.py:91-220)."""

    def __init__(self, constellation: list, reorder_as_gray: bool = True):
        self.alpha = 1
        self.constel_size = len(constellation)
        if reorder_as_gray:
            gray = np.asarray([x ^ (x >> 1) for x in range(self.constel_size)])
            self.constellation = np.array(constellation)[gray.argsort()]
        else:
            self.constellation = constellation

    # -- constellation property: same side effects as the reference setter (:205-220) --
    @property
    def constellation(self):
        return self._constellation

    @constellation.setter
    def constellation(self, constelation_symb_lst: list):
        n_bits = np.log2(len(constelation_symb_lst))
        if n_bits != int(n_bits):
            raise ValueError('Constellation length must be a power of 2.')
        self._constellation = np.array(constelation_symb_lst)
        self.avg_symbol_power = np.mean(np.abs(self._constellation) ** 2)
        self.constellation_size = self._constellation.size
        self.n_bits_per_symbol = int(n_bits)

    def _modem_desc(self, n_fft=8, n_sub_carr=2, cp_len=0):
        return _desc(self._constellation, n_fft, n_sub_carr, cp_len)

    def modulate(self, input_bits: ndarray) -> ndarray:
        return modulate(self._constellation, self.n_bits_per_symbol, input_bits)

    def demodulate(self, input_symbols: ndarray) -> ndarray:
        return demodulate(self._constellation, self.n_bits_per_symbol, input_symbols)

    def symbol_detection(self, input_symbols: ndarray) -> ndarray:
        """Nearest constellation point of every input symbol (modulation.py:138-146)."""
        x = np.asarray(input_symbols)
        sym, _ = _abi.get_ops().qam_detect(self._modem_desc(), x.reshape(-1), want_bits=False)
        return sym.reshape(x.shape)

    def plot_constellation(self) -> None:
        import matplotlib.pyplot as plt
        fig, ax = plt.subplots()
        ax.scatter(self.constellation.real, self.constellation.imag)
        ax.set_title('Constellation')
        ax.set_xlabel("In-phase")
        ax.set_ylabel("Quadrature")
        ax.grid()
        plt.show()

    def correct_constellation(self, ibo_db: float) -> None:
        """Shrink the reference constellation by the Bussgang alpha (modulation.py:167-176)."""
        self.alpha = self.calc_alpha(ibo_db)
        self._constellation = self.alpha * self._constellation

    def calc_alpha(self, ibo_db):
        """alpha = 1 - exp(-g^2) + sqrt(pi) g / 2 erfc(g), g = 10^(ibo/20) (modulation.py:178-189).  A closed-form
        scalar/K-vector parameter computation; the per-realisation, per-antenna evaluation of the hot path is done
        inside k_prep_realisation / k_vk_pow."""
        from scipy import special
        gamma = np.power(10, np.asarray(ibo_db, dtype=np.float64) / 20)
        alpha = 1 - np.exp(-np.power(gamma, 2)) + (np.sqrt(np.pi) * gamma / 2) * special.erfc(gamma)
        return alpha if np.ndim(ibo_db) else float(alpha)

    def update_alpha(self, ibo_db: float) -> None:
        self.alpha = self.calc_alpha(ibo_db)


class QamModem(Modem):
    """Square QAM modem (modulation.py:223-244)."""

    def __init__(self, constel_size):
        n_symb = np.sqrt(constel_size)
        if n_symb != int(n_symb):
            raise ValueError('Constellation size must be a power of some number, only square QAM supported.')
        pam = np.arange(-n_symb + 1, n_symb, 2)
        base = np.tile(np.hstack((pam, pam[::-1])), int(n_symb) // 2) * 1j + pam.repeat(n_symb)
        super().__init__(base)


def _tx_ofdm_symbol(mod_symbols: ndarray, n_fft: int, n_sub_carr: int, cp_length: int) -> ndarray:
    """Sub-carrier placement + orthonormal IFFT + cyclic prefix (modulation.py:248-273)."""
    if len(mod_symbols) != n_sub_carr:
        raise ValueError('mod_symbols length must match n_sub_carr value')
    grid = np.zeros(n_fft, dtype=np.complex128)
    grid[-(n_sub_carr // 2):] = mod_symbols[0:n_sub_carr // 2]
    grid[1:(n_sub_carr // 2) + 1] = mod_symbols[n_sub_carr // 2:]
    return _abi.get_ops().fft(grid, inverse=True, cp=cp_length)


def _rx_ofdm_symbol(ofdm_symbol: ndarray, n_fft: int, n_sub_carr: int, cp_length: int) -> ndarray:
    """Cyclic-prefix removal + orthonormal FFT + data sub-carrier extraction (modulation.py:277-293)."""
    fd = _abi.get_ops().fft(np.asarray(ofdm_symbol), inverse=False, skip=cp_length)
    return np.concatenate((fd[-n_sub_carr // 2:], fd[1:(n_sub_carr // 2) + 1]))


class OfdmQamModem(QamModem):
    """QAM OFDM modem (modulation.py:296-424)."""

    def __init__(self, constel_size: int, n_fft: int, n_sub_carr: int, cp_len: int, n_users: int = 1):
        super().__init__(constel_size)
        self.n_fft = n_fft
        self.n_sub_carr = n_sub_carr
        self.cp_len = cp_len
        self.n_bits_per_ofdm_sym = int(np.log2(constel_size) * n_sub_carr)
        self.avg_sample_power = self.ofdm_avg_sample_pow()
        self.precoding_mat = None
        self.n_users = n_users

    def _modem_desc(self, n_fft=None, n_sub_carr=None, cp_len=None):
        return _desc(self._constellation, self.n_fft, self.n_sub_carr, self.cp_len)

    def set_precoding(self, precoding_mat: ndarray) -> None:
        self.precoding_mat = precoding_mat

    def precode_symbols(self, in_symbols: ndarray, precoding_mat: ndarray = None) -> ndarray:
        """Element-wise precoding multiply (modulation.py:332-344)."""
        if precoding_mat is None:
            return in_symbols
        a, b = np.broadcast_arrays(np.asarray(in_symbols, dtype=np.complex128), np.asarray(precoding_mat, dtype=np.complex128))
        return _abi.get_ops().propagate(a, b, sum_signals=False)

    def _check_bits(self, bits):
        bits = np.asarray(bits)
        want = (self.n_bits_per_ofdm_sym,) if self.n_users == 1 else (self.n_users, self.n_bits_per_ofdm_sym)
        if self.n_users == 1:
            bits = bits.reshape(-1)
        if bits.shape != want:
            raise ValueError('mod_symbols length must match n_sub_carr value')
        return bits

    def modulate(self, input_bits: ndarray, get_symbols_only: bool = False, sum_usr_signals: bool = True):
        """Bits -> symbols -> precode -> OFDM symbol in the time domain with cyclic prefix (modulation.py:346-382)."""
        ops = _abi.get_ops()
        md = self._modem_desc()
        if get_symbols_only:
            bits = np.asarray(input_bits)
            if self.n_users == 1:
                return ops.qam_map(md, bits.reshape(-1))
            return np.stack([ops.qam_map(md, bits[u]) for u in range(self.n_users)])
        bits = self._check_bits(input_bits)
        v = None
        if self.precoding_mat is not None:
            v = np.asarray(self.precoding_mat, dtype=np.complex128).reshape(1, self.n_users, self.n_sub_carr)
        per_user = (self.n_users > 1) and not sum_usr_signals
        out = ops.array_transmit(md, None, 1, self.n_users, bits, v, skip_dist=True, time_domain=True, per_user=per_user)
        if per_user:
            return [out[u, 0] for u in range(self.n_users)]
        return out[0, 0]

    def demodulate(self, ofdm_symbol: ndarray, get_symbols_only: bool = False) -> ndarray:
        """Time-domain OFDM symbol -> data symbols or hard-decision bits (modulation.py:384-396)."""
        sym, bits = _abi.get_ops().ofdm_demodulate(self._modem_desc(), ofdm_symbol, want_bits=not get_symbols_only)
        return sym if get_symbols_only else bits

    def symbols_to_bits(self, input_symbols: ndarray) -> ndarray:
        return demodulate(self._constellation, self.n_bits_per_symbol, input_symbols)

    def soft_detection_llr(self, baseband_symbols, noise_var: float) -> ndarray:
        return demodulate(self._constellation, self.n_bits_per_symbol, baseband_symbols, soft=True, noise_var=noise_var)

    def ofdm_avg_sample_pow(self) -> float:
        return self.avg_symbol_power * (self.n_sub_carr / self.n_fft)
