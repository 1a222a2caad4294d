"""CPU oracle for the Monte-Carlo link hot path  --  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A vectorised numpy (complex128 / float64) restatement of the reference's link
arithmetic.  Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s
`cpu_baseline` / `--impl reference` legs may import this module; the product
package (`m_mimo_ofdm_with_nonlinear_pa_sim_b200`) never does and has no CPU fallback.

Parity status: the reference ships no golden vectors or known-answer tests
(SURVEY.md section 4), and its FFT is third-party `torch.fft` (unpinned).  This
oracle is therefore pinned against OUTPUTS OF THE REFERENCE ITSELF RUN IN THE
BUILD CONTAINER: `tests/golden/make_golden.py` imports the unmodified reference
(via oracle/refharness.py) and stores its inputs/outputs under `tests/golden/`;
`tests/test_oracle_golden.py` checks this file against those fixtures and
`tests/test_oracle_vs_reference.py` re-checks live when `/root/reference` exists.

Every function cites the reference file:line it restates (paths relative to the
reference root).  Layout conventions are the reference's: frequency-domain
vectors in FFT-bin order (DC at index 0); data-subcarrier vectors in "negative
half then positive half" order (modulation.py:266-267).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np
from scipy import special as _sp

C_LIGHT = 299792458.0  # scipy.constants.c, used by channel.py:61,66


# ----------------------------------------------------------------------------------------------
# constellation / bit mapping
# ----------------------------------------------------------------------------------------------
def qam_constellation(constel_size: int) -> np.ndarray:
    """Gray-reordered square-QAM table, index = integer value of the bit group (MSB first).

    modulation.py:230-244 (base table: pam.repeat + zig-zag imaginary part) and
    modulation.py:110-114 (constellation = base[argsort(gray)]).
    """
    side = int(round(math.sqrt(constel_size)))
    if side * side != constel_size:
        raise ValueError('Constellation size must be a power of some number, only square QAM supported.')
    nb = math.log2(constel_size)
    if nb != int(nb):
        raise ValueError('Constellation length must be a power of 2.')
    pam = np.arange(-side + 1, side, 2, dtype=np.float64)
    n = np.arange(constel_size)
    col = n // side  # real level index
    row = n % side
    # np.tile(hstack(pam, pam[::-1]), side//2): even columns ascend, odd columns descend
    imag = np.where(col % 2 == 0, pam[row], pam[side - 1 - row])
    base = pam[col] + 1j * imag
    gray = n ^ (n >> 1)
    return base[np.argsort(gray, kind="stable")]


def bits_to_indices(bits: np.ndarray, n_bits_per_symbol: int) -> np.ndarray:
    """Groups of bits, MSB first, to integers.  utilities.py:54-67 (bitarray2dec), modulation.py:22-24."""
    b = np.asarray(bits).astype(np.int64).reshape(-1, n_bits_per_symbol)
    w = 1 << np.arange(n_bits_per_symbol - 1, -1, -1, dtype=np.int64)
    return b @ w


def indices_to_bits(idx: np.ndarray, n_bits_per_symbol: int) -> np.ndarray:
    """Integers to int8 bit array, MSB first.  utilities.py:18-51 (dec2bitarray)."""
    idx = np.asarray(idx).astype(np.int64)
    sh = np.arange(n_bits_per_symbol - 1, -1, -1, dtype=np.int64)
    return ((idx[:, None] >> sh[None, :]) & 1).astype(np.int8).reshape(-1)


def detect_indices_argmin(symbols: np.ndarray, constellation: np.ndarray) -> np.ndarray:
    """Literal nearest-point search.  modulation.py:145 / :76 (first index wins ties)."""
    return np.abs(symbols[None, :] - constellation[:, None]).argmin(0)


def detect_indices(symbols: np.ndarray, constel_size: int, scale: float = 1.0) -> np.ndarray:
    """Per-axis slicing equivalent of `detect_indices_argmin` for the (optionally alpha-scaled,
    modulation.py:176) square-QAM table.  Identical away from exact decision boundaries."""
    side = int(round(math.sqrt(constel_size)))
    z = np.asarray(symbols) / scale
    li = np.clip(np.floor((z.real + side) / 2.0), 0, side - 1).astype(np.int64)
    qi = np.clip(np.floor((z.imag + side) / 2.0), 0, side - 1).astype(np.int64)
    row = np.where(li % 2 == 0, qi, side - 1 - qi)
    n = li * side + row
    return n ^ (n >> 1)


def detection_margin(symbols: np.ndarray, constel_size: int) -> np.ndarray:
    """Distance of each sample (per axis, min of both) to the nearest decision boundary."""
    side = int(round(math.sqrt(constel_size)))
    z = np.asarray(symbols)

    def axis_margin(x):
        xc = np.clip(x, -(side - 2), side - 2)  # outside the outer boundaries there is none
        frac = np.abs(xc / 2.0 - np.round(xc / 2.0)) * 2.0
        out = np.abs(x) > (side - 2)
        return np.where(out, np.abs(np.abs(x) - (side - 2)), frac)

    return np.minimum(axis_margin(z.real), axis_margin(z.imag))


def soft_llr(symbols: np.ndarray, constellation: np.ndarray, noise_var) -> np.ndarray:
    """modulation.soft_decoding, modulation.py:29-59: exact LLR of every bit (MSB of each symbol first),
    log(sum_{bit set} e^{-|y-c|^2/nv} / sum_{bit clear} e^{...}); +inf where the denominator underflows to zero."""
    y = np.asarray(symbols, dtype=np.complex128).reshape(-1)
    nv = np.broadcast_to(np.asarray(noise_var, dtype=np.float64), y.shape)
    n_bits = int(math.log2(len(constellation)))
    e = np.exp(-np.abs(y[:, None] - constellation[None, :]) ** 2 / nv[:, None])  # [n, M]
    labels = np.arange(len(constellation))
    out = np.empty((y.size, n_bits))
    with np.errstate(divide="ignore", invalid="ignore"):
        for b in range(n_bits):
            m = ((labels >> b) & 1).astype(bool)
            num, den = e[:, m].sum(axis=1), e[:, ~m].sum(axis=1)
            out[:, n_bits - 1 - b] = np.where(den == 0, np.inf, np.log(num / den))
    return out.reshape(-1)


def calc_alpha(ibo_db):
    """Bussgang shrinking coefficient of the soft limiter.  modulation.py:178-189."""
    gamma = np.power(10.0, np.asarray(ibo_db, dtype=np.float64) / 20.0)
    return 1.0 - np.exp(-np.power(gamma, 2)) + (np.sqrt(np.pi) * gamma / 2.0) * _sp.erfc(gamma)


def ebn0_to_snr(eb_per_n0, n_fft: int, n_sub_carr: int, constel_size: int):
    """utilities.py:107-118.  NB the scripts pass n_fft := n_sub_carr."""
    return 10 * np.log10(10 ** (np.asarray(eb_per_n0, dtype=np.float64) / 10) * n_sub_carr
                         * np.log2(constel_size) / n_fft)


# ----------------------------------------------------------------------------------------------
# OFDM grid / transforms
# ----------------------------------------------------------------------------------------------
def sc_to_bins(n_fft: int, n_sub_carr: int) -> np.ndarray:
    """FFT-bin index of every data sub-carrier, in data-vector order.  modulation.py:266-267."""
    h = n_sub_carr // 2
    return np.concatenate((np.arange(n_fft - h, n_fft), np.arange(1, h + 1)))


def take_data_sc(vec_fd: np.ndarray, n_sub_carr: int) -> np.ndarray:
    """corrector.py:66,176 / modulation.py:293 (works on the last axis)."""
    h = n_sub_carr // 2
    return np.concatenate((vec_fd[..., -h:], vec_fd[..., 1:h + 1]), axis=-1)


def place_on_grid(data_sc: np.ndarray, n_fft: int, fill=0.0) -> np.ndarray:
    """modulation.py:264-267 (fill=0) and mp_model.py:307-309 (fill=1)."""
    n_sc = data_sc.shape[-1]
    h = n_sc // 2
    grid = np.full(data_sc.shape[:-1] + (n_fft,), fill, dtype=np.complex128)
    grid[..., n_fft - h:] = data_sc[..., :h]
    grid[..., 1:h + 1] = data_sc[..., h:]
    return grid


def ifft_ortho(x):
    """torch.fft.ifft(norm="ortho") call sites modulation.py:270, utilities.py:339, corrector.py:93."""
    return np.fft.ifft(x, axis=-1, norm="ortho")


def fft_ortho(x):
    """torch.fft.fft(norm="ortho") call sites modulation.py:290, utilities.py:329, corrector.py:98."""
    return np.fft.fft(x, axis=-1, norm="ortho")


# ----------------------------------------------------------------------------------------------
# power-amplifier models
# ----------------------------------------------------------------------------------------------
PA_NONE, PA_SOFT_LIMITER, PA_RAPP, PA_TOI = 0, 1, 2, 3


def pa_soft_limiter(x, sat_pow):
    """distortion.py:9-19."""
    p = np.abs(x) ** 2
    safe = np.where(x != 0, x, 1)
    return np.where(p <= sat_pow, x, x * np.sqrt(sat_pow / np.abs(safe) ** 2))


def pa_rapp(x, sat_pow, p_hardness):
    """distortion.py:102-113."""
    return x / np.power(1 + np.power(np.abs(x) / np.sqrt(sat_pow), 2 * p_hardness), 1 / (2 * p_hardness))


def pa_toi(x, cubic_coeff):
    """distortion.py:202-211."""
    return x - cubic_coeff * x * np.abs(x) ** 2


def pa_apply(x, kind, p0, p1=0.0):
    if kind == PA_NONE:
        return x
    if kind == PA_SOFT_LIMITER:
        return pa_soft_limiter(x, p0)
    if kind == PA_RAPP:
        return pa_rapp(x, p0, p1)
    if kind == PA_TOI:
        return pa_toi(x, p0)
    raise ValueError(kind)


# ----------------------------------------------------------------------------------------------
# geometry / channels
# ----------------------------------------------------------------------------------------------
def ula_positions(n_elements: int, center_freq: float, wav_len_spacing: float, cord_z: float) -> np.ndarray:
    """[K,3] element coordinates of a ULA along X.  antenna_array.py:436-445."""
    lam = C_LIGHT / center_freq
    half = (n_elements - 1) * wav_len_spacing * lam / 2
    x = np.linspace(-half, half, n_elements)
    return np.stack((x, np.zeros(n_elements), np.full(n_elements, float(cord_z))), axis=1)


def subcarrier_freqs(n_fft: int, carrier_spacing: float, center_freq: float) -> np.ndarray:
    """channel.py:51-52: `torch.fft.fftfreq(n_fft, d=1/n_fft).numpy() * spacing + fc`.

    QUIRK reproduced on purpose: torch.fft.fftfreq returns float32 and numpy keeps float32 when a
    Python int is multiplied/added, so the reference's sub-carrier frequencies are ROUNDED TO FLOAT32
    (a 256 Hz grid around 3.5 GHz).  They are then used in float64 arithmetic (np.outer with float64
    distances), which is what the returned float64 array holding the float32 values does."""
    k = np.fft.fftfreq(n_fft, d=1.0 / n_fft).astype(np.float32)
    f32 = k * np.float32(carrier_spacing) + np.float32(center_freq)
    assert f32.dtype == np.float32
    return f32.astype(np.float64)


def _distances(tx_pos: np.ndarray, rx_pos) -> np.ndarray:
    rx = np.asarray(rx_pos, dtype=np.float64)
    return np.sqrt(np.power(tx_pos[:, 0] - rx[0], 2) + np.power(tx_pos[:, 1] - rx[1], 2)
                   + np.power(tx_pos[:, 2] - rx[2], 2))


def fspl_matrix(dist, freqs, tx_gain_db=0.0, rx_gain_db=0.0):
    """channel.py:65-66."""
    g = np.sqrt(np.power(10, (np.asarray(tx_gain_db, dtype=np.float64) + rx_gain_db) / 10))
    g = np.broadcast_to(g, dist.shape)[:, None]
    return g * (C_LIGHT / (4 * np.pi * np.outer(dist, freqs)))


def channel_los(tx_pos, rx_pos, freqs, skip_attenuation=False):
    """MisoLosFd.calc_channel_mat, channel.py:35-72."""
    d = _distances(tx_pos, rx_pos)
    ph = np.exp(2j * np.pi * np.outer(d, freqs) / C_LIGHT)
    if skip_attenuation:
        return ph
    return np.multiply(ph, fspl_matrix(d, freqs))


def channel_two_path(tx_pos, rx_pos, freqs, skip_attenuation=False):
    """MisoTwoPathFd.calc_channel_mat, channel.py:116-167 (ground reflection coefficient -1)."""
    rx = np.asarray(rx_pos, dtype=np.float64)
    d1 = _distances(tx_pos, rx)
    dxy = np.sqrt(np.power(tx_pos[:, 0] - rx[0], 2) + np.power(tx_pos[:, 1] - rx[1], 2))
    elev = np.arctan((tx_pos[:, 2] + rx[2]) / dxy)
    d2 = tx_pos[:, 2] / np.sin(elev) + rx[2] / np.sin(elev)
    p1 = np.exp(2j * np.pi * np.outer(d1, freqs) / C_LIGHT)
    p2 = -1.0 * np.exp(2j * np.pi * np.outer(d2, freqs) / C_LIGHT)
    if not skip_attenuation:
        p1 = np.multiply(p1, fspl_matrix(d1, freqs))
        p2 = np.multiply(p2, fspl_matrix(d2, freqs))
    return np.add(p1, p2)


def rayleigh_coeffs(rng: np.random.Generator, n_ant: int, n_fft: int) -> np.ndarray:
    """channel.py:244-245 / :270-271: CN(0,1) IID coefficients from a numpy Generator."""
    return rng.standard_normal(size=(n_ant, n_fft * 2)).view(dtype=np.complex128) / np.sqrt(2.0)


def channel_rayleigh(coeffs, tx_pos, rx_pos, freqs, skip_attenuation=False):
    """MisoRayleighFd, channel.py:198-229,262-275: IID coefficients times the LOS free-space loss
    evaluated ONCE at construction time (the attenuation does not follow later receiver moves)."""
    if skip_attenuation:
        return coeffs
    return np.multiply(coeffs, fspl_matrix(_distances(tx_pos, rx_pos), freqs))


def channel_random_paths(tx_pos, freqs, angles_of_dep, tau_delays):
    """MisoRandomPathsFd, channel.py:330-344 -- literally as written there (no pi in the exponent,
    sin(angle / c)); reference antenna = element 0."""
    delta = np.sqrt(np.sum(np.power(tx_pos - tx_pos[0][None, :], 2), axis=1))  # [K]
    arg = tau_delays[None, None, :] + delta[:, None, None] * np.sin(angles_of_dep / C_LIGHT)[None, None, :]
    coeff = np.exp(-2j * freqs[None, :, None] * arg)
    return (1 / np.sqrt(len(tau_delays))) * np.sum(coeff, axis=2)


def csi_error_channel(h_fd, n_sub_carr, eps, rng: np.random.Generator):
    """mp_model.py:264-284: per antenna, data sub-carriers only,
    h_hat = sqrt(1-eps^2) h + eps * CN(0, mean|h|^2), noise drawn row by row from `rng`."""
    n_fft = h_fd.shape[1]
    bins = sc_to_bins(n_fft, n_sub_carr)
    out = np.copy(h_fd)
    for k in range(h_fd.shape[0]):
        sc = h_fd[k, bins]
        p = np.sum(np.abs(sc) ** 2) / len(sc)
        nz = rng.standard_normal((len(sc), 2)).view(np.complex128)[:, 0] * 0.5 * np.sqrt(2 * p) * eps
        out[k, bins] = np.sqrt(1 - np.power(eps, 2)) * sc + nz
    return out


# ----------------------------------------------------------------------------------------------
# precoding
# ----------------------------------------------------------------------------------------------
def precoding_mrt(h_fd, n_sub_carr):
    """antenna_array.py:163-171: V = conj(H_sc) / sqrt(sum_k |H_sc|^2), shape [K, N_sc]."""
    hsc = take_data_sc(h_fd, n_sub_carr)
    return np.divide(np.conjugate(hsc), np.sqrt(np.sum(np.power(np.abs(hsc), 2), axis=0)))


def precoding_phase_only(h_fd, n_sub_carr):
    """antenna_array.py:175-178."""
    return np.exp(1j * np.angle(np.conjugate(take_data_sc(h_fd, n_sub_carr))))


def precoding_mu_mrt(h_list, n_sub_carr):
    """antenna_array.py:202-220: [U, K, N_sc], common per-sub-carrier normalisation over users."""
    hs = [take_data_sc(h, n_sub_carr) for h in h_list]
    norm = np.sqrt(np.sum([np.sum(np.power(np.abs(h), 2), axis=0) for h in hs], axis=0))
    return np.stack([np.divide(np.conjugate(h), norm) for h in hs], axis=0)


def precoding_mu_zf(h_list, n_sub_carr):
    """antenna_array.py:222-257: per sub-carrier T=[h_1..h_U] (K x U),
    V = sqrt(K-U) conj(T) inv(T^T conj(T)), then unit Frobenius norm per sub-carrier.  [U, K, N_sc]."""
    hs = np.stack([take_data_sc(h, n_sub_carr) for h in h_list], axis=0)  # [U,K,S]
    n_usr, n_ant, n_sc = hs.shape
    t = np.transpose(hs, (2, 1, 0))  # [S,K,U]
    gram = np.matmul(np.transpose(t, (0, 2, 1)), np.conjugate(t))  # [S,U,U]
    v = np.sqrt(n_ant - n_usr) * np.matmul(np.conjugate(t), np.linalg.inv(gram))  # [S,K,U]
    v = np.transpose(v, (2, 1, 0))  # [U,K,S]
    norm = np.sqrt(np.sum(np.power(np.abs(v), 2), axis=(0, 1)))
    return v / norm


def precoding_sep_carriers(h_list, n_sub_carr, mr_precoding=True):
    """antenna_array.py:275-305: user u owns the u-th contiguous block of data sub-carriers."""
    n_usr = len(h_list)
    blocks = [np.hsplit(take_data_sc(h, n_sub_carr), n_usr)[u] for u, h in enumerate(h_list)]
    comp = np.concatenate(blocks, axis=1)
    if mr_precoding:
        return np.divide(np.conjugate(comp), np.sqrt(np.sum(np.power(np.abs(comp), 2), axis=0)))
    return np.exp(1j * np.angle(np.conjugate(comp)))


def avg_precoding_gain(v):
    """antenna_array.py:328-341.  v is [K,N_sc] or [U,K,N_sc] (users summed per antenna)."""
    if v.ndim == 2:
        return np.average(np.power(np.abs(v), 2))
    return np.average(np.sum(np.power(np.abs(v), 2), axis=0))


# ----------------------------------------------------------------------------------------------
# configuration + array transmit
# ----------------------------------------------------------------------------------------------
@dataclass
class LinkConfig:
    """Static parameters of one link (main_mp_miso_mcnc_ber_vs_ebn0.py:41-97)."""
    constel_size: int = 64
    n_fft: int = 4096
    n_sub_carr: int = 2048
    cp_len: int = 128
    n_ant: int = 64
    center_freq: float = 3.5e9
    carrier_spacing: float = 15e3
    wav_len_spacing: float = 0.5
    array_z: float = 15.0
    rx_pos: tuple = (212.0, 212.0, 1.5)
    rx_loc_var: float = 10.0
    pa_kind: int = PA_SOFT_LIMITER
    ibo_db: float = 0.0          # soft limiter / Rapp: IBO; TOI: the value handed to set_toi
    p_hardness: float = 4.0
    channel: str = "los"         # los | two_path | rayleigh
    constellation: np.ndarray = field(default=None, repr=False)

    def __post_init__(self):
        self.constellation = qam_constellation(self.constel_size)
        self.n_bits_per_symbol = int(math.log2(self.constel_size))
        self.n_bits_per_ofdm_sym = self.n_bits_per_symbol * self.n_sub_carr
        self.avg_symbol_power = float(np.mean(np.abs(self.constellation) ** 2))  # modulation.py:218
        self.avg_sample_power = self.avg_symbol_power * (self.n_sub_carr / self.n_fft)  # modulation.py:424
        self.tx_pos = ula_positions(self.n_ant, self.center_freq, self.wav_len_spacing, self.array_z)
        self.freqs = subcarrier_freqs(self.n_fft, self.carrier_spacing, self.center_freq)
        self.bins = sc_to_bins(self.n_fft, self.n_sub_carr)


def pa_params(cfg: LinkConfig, v) -> tuple:
    """(kind, p0, p1) the array's PAs end up with after AntennaArray.update_distortion
    (antenna_array.py:313-360): ONE saturation power for every antenna,
    sat_pow = 10^(ibo/10) * avg_sample_power * mean|V|^2  (distortion.py:37,51,61)."""
    gain = 1.0 if v is None else avg_precoding_gain(v)
    avg = cfg.avg_sample_power * gain
    if cfg.pa_kind == PA_NONE:
        return PA_NONE, 0.0, 0.0
    if cfg.pa_kind == PA_TOI:
        return PA_TOI, 1 / (np.power(10, (cfg.ibo_db / 10))) / avg, 0.0  # distortion.py:228,251
    return cfg.pa_kind, np.power(10, cfg.ibo_db / 10) * avg, cfg.p_hardness


def array_transmit(cfg: LinkConfig, sym_idx, v, pa, skip_dist=False, return_stages=False):
    """AntennaArray.transmit(out_domain_fd=True), antenna_array.py:58-140 -> Transceiver.transmit
    (transceiver.py:98-174) -> OfdmQamModem.modulate (modulation.py:346-382) for all antennas at once.

    sym_idx: [N_sc] (single user) or [U, N_sc] integer symbol values; v: [K,N_sc] or [U,K,N_sc] or None.
    Returns X [K, n_fft] in the frequency domain.  The cyclic prefix (modulation.py:273) is added and
    removed again (utilities.py:323) around a memoryless PA, so it never reaches the output and is
    skipped here.
    """
    sym = cfg.constellation[np.asarray(sym_idx)]
    if sym.ndim == 1:
        pre = sym[None, :] * (v if v is not None else np.ones((cfg.n_ant, 1)))
    else:
        pre = np.sum(sym[:, None, :] * v, axis=0)  # modulation.py:373
    grid = place_on_grid(pre, cfg.n_fft)
    td = ifft_ortho(grid)
    td_pa = td if skip_dist else pa_apply(td, *pa)
    out = fft_ortho(td_pa)
    if return_stages:
        return out, dict(symbols=sym, precoded=pre, grid=grid, td=td, td_pa=td_pa)
    return out


def propagate(x_fd, h_fd, sum_signals=True):
    """channel.py:74-89."""
    y = np.multiply(x_fd, h_fd)
    return np.sum(y, axis=0) if sum_signals else y


def awgn_noise(rng: np.random.Generator, n_sampl: int, avg_sample_pow: float, snr_db: float) -> np.ndarray:
    """noise.py:64-66: the additive term only."""
    noise_std = np.complex128(np.sqrt(2 * avg_sample_pow / (10 ** (snr_db / 10))))
    return rng.standard_normal((n_sampl, 2)).view(np.complex128)[:, 0] * noise_std * 0.5


# ----------------------------------------------------------------------------------------------
# per-realisation AGC / alpha bookkeeping
# ----------------------------------------------------------------------------------------------
@dataclass
class Realisation:
    """Everything Link.set_precoding_and_recalculate_agc (mp_model.py:253-329) derives from one
    channel realisation."""
    h_true: np.ndarray      # [K, n_fft] propagates the real transmission
    h_loop: np.ndarray      # [K, n_fft] channel knowledge used for precoding/AGC/MCNC (== h_true w/o CSI error)
    v: np.ndarray           # [K, N_sc]
    pa: tuple
    alpha_k: np.ndarray     # [K]
    hk_vk_agc: np.ndarray   # [N_sc]  sum_k h v
    ak_hk_vk_agc: np.ndarray  # [N_sc]  sum_k alpha_k h v
    hk_vk_noise_scaler: float
    ak_hk_vk_noise_scaler: float


def per_antenna_alpha(cfg: LinkConfig, v, usr_axis=False):
    """mp_model.py:302,315-317 / corrector.py:143,149-152."""
    if v.ndim == 3:
        vk_pow = np.sum(np.sum(np.power(np.abs(v), 2), axis=2), axis=0)  # v is [U,K,S] here
    else:
        vk_pow = np.sum(np.power(np.abs(v), 2), axis=1)
    ibo_vec = 10 * np.log10(10 ** (cfg.ibo_db / 10) * cfg.n_sub_carr / (vk_pow * cfg.n_ant))
    return calc_alpha(ibo_vec)


def make_realisation(cfg: LinkConfig, h_true, h_loop=None) -> Realisation:
    """MRT precoding from `h_loop` + Link.recalculate_agc (mp_model.py:290-329)."""
    if h_loop is None:
        h_loop = h_true
    v = precoding_mrt(h_loop, cfg.n_sub_carr)
    hk = take_data_sc(h_loop, cfg.n_sub_carr)
    hv = np.multiply(hk, v)
    hk_vk = np.sum(hv, axis=0)
    ak = per_antenna_alpha(cfg, v)
    ak_hk_vk = np.sum(ak[:, None] * hv, axis=0)
    return Realisation(h_true=h_true, h_loop=h_loop, v=v, pa=pa_params(cfg, v), alpha_k=ak,
                       hk_vk_agc=hk_vk, ak_hk_vk_agc=ak_hk_vk,
                       hk_vk_noise_scaler=float(np.mean(np.power(np.abs(hk_vk), 2))),
                       ak_hk_vk_noise_scaler=float(np.mean(np.power(np.abs(ak_hk_vk), 2))))


# ----------------------------------------------------------------------------------------------
# receivers
# ----------------------------------------------------------------------------------------------
def mcnc_receive(cfg: LinkConfig, rx_sc, h_loop, v, pa, agc_sc, n_iter_max, trace=None,
                 other_usr_idx=None, usr_idx=0, skip_last_regen=True):
    """McncReceiver.receive, corrector.py:165-207 (MU variant :405-451 when `other_usr_idx` is given).

    rx_sc: equalised data sub-carrier vector [N_sc]; agc_sc: AGC on data sub-carriers [N_sc].
    Returns [n_iter_max+1, N_sc] detected symbol indices (iteration 0 = standard receiver).
    The reference also regenerates after the last iteration and discards it (corrector.py:198-205);
    `skip_last_regen` drops that dead work without changing any output.
    """
    h_sc = take_data_sc(h_loop, cfg.n_sub_carr)
    out = np.empty((n_iter_max + 1, cfg.n_sub_carr), dtype=np.int64)
    dist_est = 0.0
    for it in range(n_iter_max + 1):
        corr = rx_sc - dist_est if it != 0 else rx_sc
        idx = detect_indices(corr, cfg.constel_size)
        out[it] = idx
        if trace is not None:
            trace.setdefault("corr", []).append(corr)
        if it == n_iter_max and skip_last_regen:
            break
        if other_usr_idx is None:
            tx_idx = idx
        else:
            tx_idx = np.vstack((idx, other_usr_idx)) if usr_idx == 0 else np.vstack((other_usr_idx, idx))
        x = array_transmit(cfg, tx_idx, v, pa)
        z = np.sum(x[:, cfg.bins] * h_sc, axis=0) / agc_sc  # channel.py:84-87, corrector.py:200-203
        dist_est = z - cfg.constellation[idx]
        if trace is not None:
            trace.setdefault("est", []).append(z)
    return out


def cnc_receive(cfg: LinkConfig, rx_sc, sat_pow_or_pa, alpha, n_iter_max, trace=None,
                other_usr_symbols=None, skip_last_regen=True):
    """CncReceiver.receive, corrector.py:52-112 (CncMuReceiver.receive :288-345 with `other_usr_symbols`):
    single-antenna model of the transmitter, PA average sample power = avg_symbol_power / upsampling
    (corrector.py:34-35), estimate divided by alpha."""
    pa = sat_pow_or_pa if isinstance(sat_pow_or_pa, tuple) else (PA_SOFT_LIMITER, sat_pow_or_pa, 0.0)
    out = np.empty((n_iter_max + 1, cfg.n_sub_carr), dtype=np.int64)
    dist_est = 0.0
    for it in range(n_iter_max + 1):
        corr = rx_sc - dist_est if it != 0 else rx_sc
        idx = detect_indices(corr, cfg.constel_size)
        out[it] = idx
        if trace is not None:
            trace.setdefault("corr", []).append(corr)
        if it == n_iter_max and skip_last_regen:
            break
        sym = cfg.constellation[idx]
        tx = sym if other_usr_symbols is None else np.sqrt(2) / 2.0 * sym + np.sqrt(2) / 2.0 * other_usr_symbols
        td = ifft_ortho(place_on_grid(tx, cfg.n_fft))
        fd = fft_ortho(pa_apply(td, *pa))
        est = np.divide(fd[cfg.bins], alpha)
        dist_est = est - sym
    return out


def cnc_pa_params(cfg: LinkConfig) -> tuple:
    """PA of the CNC loop: avg_samp_pow = avg_symbol_power * n_sub_carr / n_fft (corrector.py:34-35),
    IBO from Link.update_distortion -> CncReceiver.update_distortion (corrector.py:37-50)."""
    avg = cfg.avg_symbol_power * (1 / (cfg.n_fft / cfg.n_sub_carr))
    if cfg.pa_kind == PA_TOI:
        return PA_TOI, 1 / (np.power(10, (cfg.ibo_db / 10))) / avg, 0.0
    if cfg.pa_kind == PA_NONE:
        return PA_NONE, 0.0, 0.0
    return cfg.pa_kind, np.power(10, cfg.ibo_db / 10) * avg, cfg.p_hardness


# ----------------------------------------------------------------------------------------------
# one frame and the Monte-Carlo loop
# ----------------------------------------------------------------------------------------------
def frame_distorted(cfg: LinkConfig, real: Realisation, tx_idx, noise_fd, n_iter_max, is_mcnc=True,
                    trace=None):
    """One pass of the distorted loop body, mp_model.py:208-222.

    tx_idx: [N_sc] transmitted symbol values; noise_fd: [n_fft] additive noise (all bins, noise.py:83).
    Returns (detected indices [I+1, N_sc], bit errors [I+1]).
    """
    x = array_transmit(cfg, tx_idx, real.v, real.pa)
    rx = propagate(x, real.h_true) + noise_fd
    agc_nfft = place_on_grid(real.ak_hk_vk_agc, cfg.n_fft, fill=1.0)
    rx = np.divide(rx, agc_nfft)  # mp_model.py:214
    rx_sc = rx[cfg.bins]
    if trace is not None:
        trace["x_fd"] = x
        trace["rx_eq_sc"] = rx_sc
    if is_mcnc:
        det = mcnc_receive(cfg, rx_sc, real.h_loop, real.v, real.pa, real.ak_hk_vk_agc, n_iter_max, trace=trace)
    else:
        alpha = float(calc_alpha(cfg.ibo_db)) if cfg.pa_kind in (PA_SOFT_LIMITER, PA_RAPP) else 1.0
        det = cnc_receive(cfg, rx_sc, cnc_pa_params(cfg), alpha, n_iter_max, trace=trace)
    err = count_bit_errors(tx_idx, det)
    return det, err


def mu_realisation(cfg: LinkConfig, h_list, precoding="mrt"):
    """Two-user realisation: multi-user precoder (antenna_array.py:202-257), array PA parameters (:313-360) and the
    per-user AGC / noise scalers of main_multiuser/main_multiuser_mcnc_ber_vs_ebn0.py:580-622 (= corrector.py:374-400).
    Returns dict(v [U,K,N_sc], pa, alpha_k [K], agc [U,N_sc], noise_scaler [U])."""
    v = precoding_mu_mrt(h_list, cfg.n_sub_carr) if precoding == "mrt" else precoding_mu_zf(h_list, cfg.n_sub_carr)
    ak = per_antenna_alpha(cfg, v)
    agc = np.stack([np.sum(ak[:, None] * take_data_sc(h_list[u], cfg.n_sub_carr) * v[u], axis=0) for u in range(len(h_list))])
    return dict(v=v, pa=pa_params(cfg, v), alpha_k=ak, agc=agc,
                noise_scaler=np.mean(np.power(np.abs(agc), 2), axis=1))


def mu_frame(cfg: LinkConfig, h_list, tx_idx, noise_fd, n_iter_max, precoding="mrt", loop_single_user=False, real=None,
             receiver=None, rx_eq_in=None):
    """One two-user frame (script :624-642): one array transmission carrying both users, per user propagate over its
    channel, add noise (noise_fd [U, n_fft], already scaled; None = noise free), AGC, McncMuReceiver.receive
    (corrector.py:404-446) or - loop_single_user - a single-user McncReceiver on the user's own channel (:563-572).
    receiver: "mcnc_mu" (default) | "mcnc_su" (= loop_single_user) | "cnc" (single-user CncReceiver of
    main_multiuser_cnc_ber_vs_ebn0.py:124,600) | "cnc_mu" (CncMuReceiver with the other user's symbols, corrector.py:288-345).
    rx_eq_in [U, N_sc]: receiver-only use, the equalised signal is given instead of being computed.
    Returns dict(rx_eq [U,N_sc], det [U, n_iter_max+1, N_sc], err [U, n_iter_max+1], corr [U, n_iter_max+1, N_sc] = the
    slicer inputs, used by the parity tests for the decision margins)."""
    receiver = receiver or ("mcnc_su" if loop_single_user else "mcnc_mu")
    real = real or mu_realisation(cfg, h_list, precoding)
    tx_idx = np.asarray(tx_idx)
    x = array_transmit(cfg, tx_idx, real["v"], real["pa"])
    n_usr = len(h_list)
    rx_eq, det, err, corr = [], [], [], []
    for u in range(n_usr):
        if rx_eq_in is not None:
            y = np.asarray(rx_eq_in[u])
        else:
            rx = np.sum(x * h_list[u], axis=0)
            if noise_fd is not None:
                rx = rx + noise_fd[u]
            y = rx[cfg.bins] / real["agc"][u]
        tr = {}
        if receiver in ("cnc", "cnc_mu"):
            alpha = float(calc_alpha(cfg.ibo_db)) if cfg.pa_kind in (PA_SOFT_LIMITER, PA_RAPP) else 1.0
            other = cfg.constellation[tx_idx[1 - u]] if receiver == "cnc_mu" else None
            d = cnc_receive(cfg, y, cnc_pa_params(cfg), alpha, n_iter_max, other_usr_symbols=other, trace=tr)
        elif receiver == "mcnc_su":
            su = make_realisation(cfg, h_list[u])
            pa_su = su.pa
            if precoding == "zf" and cfg.pa_kind in (PA_SOFT_LIMITER, PA_RAPP):
                # script :563-571 with zf_precoding: the single-user copy of the array (n_users = 1) takes the phase-only
                # branch of set_precoding_matrix (antenna_array.py:175-178), update_distortion fixes the PA saturation power
                # from THAT precoder (mean |v|^2 = 1), and McncReceiver.__init__ then overwrites the precoder with MRT
                # (corrector.py:135) without touching the PA: MRT loop, saturation power ibo * avg_sample_power * 1
                pa_su = (su.pa[0], np.power(10, cfg.ibo_db / 10) * cfg.avg_sample_power, su.pa[2])
            d = mcnc_receive(cfg, y, su.h_loop, su.v, pa_su, su.ak_hk_vk_agc, n_iter_max, trace=tr)
        else:
            d = mcnc_receive(cfg, y, h_list[u], real["v"], real["pa"], real["agc"][u], n_iter_max,
                             other_usr_idx=tx_idx[1 - u], usr_idx=u, trace=tr)
        rx_eq.append(y); det.append(d); corr.append(np.stack(tr["corr"]))
        err.append(count_bit_errors(tx_idx[u], d))
    return dict(rx_eq=np.stack(rx_eq), det=np.stack(det), err=np.array(err, dtype=np.int64), corr=np.stack(corr))


def frame_clean(cfg: LinkConfig, real: Realisation, tx_idx, noise_fd):
    """Clean ("no distortion") loop body, mp_model.py:158-173.  FFT(IFFT(x)) == x, so the whole chain is
    precode -> channel -> noise -> /hk_vk_agc -> detect (the final IFFT/+CP/FFT of :166-169 cancels)."""
    x = array_transmit(cfg, tx_idx, real.v, real.pa, skip_dist=True)
    rx = propagate(x, real.h_true) + noise_fd
    rx = np.divide(rx, place_on_grid(real.hk_vk_agc, cfg.n_fft, fill=1.0))
    rx = fft_ortho(ifft_ortho(rx))  # to_time_domain + demodulate (utilities.py:339, modulation.py:290)
    det = detect_indices(rx[cfg.bins], cfg.constel_size)
    return det, count_bit_errors(tx_idx, det[None, :])[0]


_POP6 = np.array([bin(i).count("1") for i in range(1 << 16)], dtype=np.int64)


def count_bit_errors(tx_idx, det_idx):
    """utilities.py:94-104 on the bit arrays == popcount(tx ^ rx) on the symbol values."""
    x = np.bitwise_xor(np.asarray(tx_idx, dtype=np.int64)[None, :], np.asarray(det_idx, dtype=np.int64))
    return _POP6[x].sum(axis=1)


def draw_rx_position(cfg: LinkConfig, loc_rng: np.random.Generator):
    """mp_model.py:143-148 / :193-198.  NB the reference adds BOTH offsets to rx_loc_x (cord_y quirk)."""
    x0 = cfg.rx_pos[0]
    px = x0 + loc_rng.uniform(low=-cfg.rx_loc_var / 2.0, high=cfg.rx_loc_var / 2.0)
    py = x0 + loc_rng.uniform(low=-cfg.rx_loc_var / 2.0, high=cfg.rx_loc_var / 2.0)
    return (px, py, cfg.rx_pos[2])


def draw_channel(cfg: LinkConfig, rx_pos, rayleigh_state=None):
    if cfg.channel == "los":
        return channel_los(cfg.tx_pos, rx_pos, cfg.freqs)
    if cfg.channel == "two_path":
        return channel_two_path(cfg.tx_pos, rx_pos, cfg.freqs)
    if cfg.channel == "rayleigh":
        rng, att = rayleigh_state
        return np.multiply(rayleigh_coeffs(rng, cfg.n_ant, cfg.n_fft), att)
    raise ValueError(cfg.channel)


def link_simulate(cfg: LinkConfig, snr_db: float, n_iter_max: int, seeds, n_frames: int, is_mcnc=True,
                  incl_clean_run=False, reroll_chan=True, csi_eps=None, rayleigh_seed=1234,
                  return_frames=False, h_fixed=None):
    """Fixed-frame-count version of Link.simulate (mp_model.py:89-222) for ONE worker, consuming the
    same numpy PCG64 streams in the same order: seeds = (bits, noise, location[, csi]).
    All iterations 0..n_iter_max are recorded for every frame (the reference stops recording an
    iteration once its error/bit budget is met, mp_model.py:181-187).

    Returns (errors [I+1 (+1 clean first)], bits_sent same shape).
    """
    bit_rng = np.random.default_rng(seeds[0])
    noise_rng = np.random.default_rng(seeds[1])
    loc_rng = np.random.default_rng(seeds[2])
    ray_state = None
    if cfg.channel == "rayleigh":
        r = np.random.default_rng(rayleigh_seed)
        att = fspl_matrix(_distances(cfg.tx_pos, cfg.rx_pos), cfg.freqs)
        rayleigh_coeffs(r, cfg.n_ant, cfg.n_fft)  # the constructor's own draw, channel.py:229
        ray_state = (r, att)
    rx_pos = cfg.rx_pos
    if h_fixed is not None:  # any channel object with reroll_chan=False: the matrix it holds (mp_model.py:190-200)
        assert not reroll_chan
        h = np.asarray(h_fixed)
    else:
        h = draw_channel(cfg, rx_pos, ray_state) if (cfg.channel != "rayleigh" or not reroll_chan) else None

    def next_realisation():
        nonlocal h, rx_pos
        if reroll_chan or h is None:
            if cfg.channel != "rayleigh":
                rx_pos = draw_rx_position(cfg, loc_rng)
            h = draw_channel(cfg, rx_pos, ray_state)
        h_loop = None
        if csi_eps is not None:
            h_loop = csi_error_channel(h, cfg.n_sub_carr, csi_eps, noise_rng)  # mp_model.py:272: my_noise.rng_gen
        return make_realisation(cfg, h, h_loop)

    n_res = n_iter_max + 1 + (1 if incl_clean_run else 0)
    errs = np.zeros(n_res, dtype=np.int64)
    bits = np.zeros(n_res, dtype=np.int64)
    frames = []
    real = None
    if incl_clean_run:
        for _ in range(n_frames):
            real = next_realisation()
            tx_bits = bit_rng.choice((0, 1), cfg.n_bits_per_ofdm_sym)
            tx_idx = bits_to_indices(tx_bits, cfg.n_bits_per_symbol)
            nz = awgn_noise(noise_rng, cfg.n_fft, cfg.avg_symbol_power * real.hk_vk_noise_scaler, snr_db)
            _, e = frame_clean(cfg, real, tx_idx, nz)
            errs[0] += e
            bits[0] += cfg.n_bits_per_ofdm_sym
    off = 1 if incl_clean_run else 0
    for _ in range(n_frames):
        real = next_realisation()
        tx_bits = bit_rng.choice((0, 1), cfg.n_bits_per_ofdm_sym)
        tx_idx = bits_to_indices(tx_bits, cfg.n_bits_per_symbol)
        nz = awgn_noise(noise_rng, cfg.n_fft, cfg.avg_symbol_power * real.ak_hk_vk_noise_scaler, snr_db)
        tr = {} if return_frames else None
        det, e = frame_distorted(cfg, real, tx_idx, nz, n_iter_max, is_mcnc=is_mcnc, trace=tr)
        errs[off:] += e
        bits[off:] += cfg.n_bits_per_ofdm_sym
        if return_frames:
            frames.append(dict(rx_pos=rx_pos, tx_idx=tx_idx, noise=nz, det=det, err=e, real=real,
                               rx_eq_sc=tr["rx_eq_sc"], corr=tr["corr"]))
    if return_frames:
        return errs, bits, frames
    return errs, bits
