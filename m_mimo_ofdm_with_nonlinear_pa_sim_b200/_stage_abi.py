"""ctypes binding of the stage-level entry points of include/mmimo_b200.h and a numpy-facing wrapper
(`StageOps`) used by the drop-in classes.  numpy complex128 in / complex128 out (the reference's dtype);
the device computes in complex64.  No CPU fallback: every method ends in a CUDA kernel."""
import ctypes as C
import os

import numpy as np

c_f_p = C.POINTER(C.c_float)
c_d_p = C.POINTER(C.c_double)
c_u8_p = C.POINTER(C.c_uint8)
c_u64_p = C.POINTER(C.c_uint64)

TX_SKIP_DIST, TX_TIME_DOMAIN, TX_PER_USER = 1, 2, 4
PREC_PHASE, PREC_MRT, PREC_ZF = 0, 1, 2
PA_NONE, PA_SOFT_LIMITER, PA_RAPP, PA_TOI = 0, 1, 2, 3
CHAN_LOS, CHAN_TWO_PATH = 0, 1


class ModemDesc(C.Structure):
    _fields_ = [("constel_size", C.c_int32), ("n_fft", C.c_int32), ("n_sub_carr", C.c_int32), ("cp_len", C.c_int32),
                ("constel_scale", C.c_double)]


class PaDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("reserved", C.c_int32), ("p0", C.c_double), ("p1", C.c_double)]


STAGE_EXPORTS = ["mmimo_fft", "mmimo_welch_psd", "mmimo_pa_process", "mmimo_qam_map", "mmimo_qam_detect", "mmimo_qam_soft_llr", "mmimo_array_transmit",
                 "mmimo_ofdm_demodulate", "mmimo_propagate", "mmimo_awgn_add", "mmimo_count_mismatched_bits",
                 "mmimo_channel_geometric", "mmimo_channel_fspl", "mmimo_channel_random_paths", "mmimo_precoding",
                 "mmimo_agc", "mmimo_cnc_receive", "mmimo_mcnc_receive"]


def bind(lib):
    vp, i, sz, d = C.c_void_p, C.c_int, C.c_size_t, C.c_double
    md, pd = C.POINTER(ModemDesc), C.POINTER(PaDesc)
    lib.mmimo_fft.argtypes = [vp, i, i, i, i, i, c_f_p, c_f_p]
    lib.mmimo_welch_psd.argtypes = [vp, sz, c_f_p, i, i, d, c_d_p, c_d_p]
    lib.mmimo_pa_process.argtypes = [vp, pd, sz, c_f_p, c_f_p]
    lib.mmimo_qam_map.argtypes = [vp, md, sz, c_u8_p, c_f_p]
    lib.mmimo_qam_detect.argtypes = [vp, md, sz, c_f_p, c_f_p, c_u8_p]
    lib.mmimo_qam_soft_llr.argtypes = [vp, md, sz, c_f_p, c_d_p, sz, c_d_p]
    lib.mmimo_array_transmit.argtypes = [vp, md, pd, i, i, c_u8_p, c_f_p, i, c_f_p, c_f_p]
    lib.mmimo_ofdm_demodulate.argtypes = [vp, md, c_f_p, c_f_p, c_u8_p]
    lib.mmimo_propagate.argtypes = [vp, i, i, c_f_p, c_f_p, i, c_f_p]
    lib.mmimo_awgn_add.argtypes = [vp, sz, c_f_p, c_f_p, d, c_f_p]
    lib.mmimo_count_mismatched_bits.argtypes = [vp, sz, c_u8_p, c_u8_p, c_u64_p]
    lib.mmimo_channel_geometric.argtypes = [vp, i, i, i, c_d_p, c_d_p, d, c_d_p, c_d_p, i, c_f_p]
    lib.mmimo_channel_fspl.argtypes = [vp, i, i, c_d_p, c_d_p, d, c_d_p, c_d_p, c_f_p, c_f_p]
    lib.mmimo_channel_random_paths.argtypes = [vp, i, i, c_d_p, c_d_p, i, c_d_p, c_d_p, c_f_p]
    lib.mmimo_precoding.argtypes = [vp, i, i, i, i, i, i, c_f_p, c_f_p]
    lib.mmimo_agc.argtypes = [vp, i, i, i, i, i, c_f_p, c_f_p, d, c_d_p, c_f_p, c_f_p, c_d_p, c_d_p, c_f_p]
    lib.mmimo_cnc_receive.argtypes = [vp, md, pd, d, i, c_f_p, c_f_p, c_u8_p, c_f_p]
    lib.mmimo_mcnc_receive.argtypes = [vp, md, pd, i, i, i, c_f_p, c_f_p, c_f_p, i, c_f_p, c_u8_p, c_u8_p, c_f_p]
    for name in STAGE_EXPORTS:
        getattr(lib, name).restype = C.c_int


def _c64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.complex64))


def _f64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


def _u8(a):
    return np.ascontiguousarray(np.asarray(a).astype(np.uint8, copy=False))


def _p(a, ct):
    return a.ctypes.data_as(C.POINTER(ct)) if a is not None else None


def modem_desc(constel_size, n_fft, n_sub_carr, cp_len, scale=1.0):
    return ModemDesc(int(constel_size), int(n_fft), int(n_sub_carr), int(cp_len), float(scale))


def pa_desc(kind, p0=0.0, p1=0.0):
    return PaDesc(int(kind), 0, float(p0), float(p1))


class StageOps:
    """numpy-level access to the stage kernels of one mmimo context (one process, one GPU)."""

    def __init__(self, lib, handle, check):
        self._lib, self._h, self._check = lib, handle, check

    def _call(self, fn, *args):
        self._check(self._lib, fn(self._h, *args))

    # -- transforms -----------------------------------------------------------------------------
    def fft(self, x, inverse=False, skip=0, cp=0):
        x = _c64(x)
        n = x.shape[-1] - skip
        rows = int(np.prod(x.shape[:-1])) if x.ndim > 1 else 1
        out = np.empty(x.shape[:-1] + (n + cp,), dtype=np.complex64)
        self._call(self._lib.mmimo_fft, int(bool(inverse)), n, rows, int(skip), int(cp), _p(x, C.c_float), _p(out, C.c_float))
        return out.astype(np.complex128)

    def pa_process(self, pa, x):
        x = np.asarray(x)
        xc = _c64(x)
        out = np.empty_like(xc)
        self._call(self._lib.mmimo_pa_process, C.byref(pa), xc.size, _p(xc, C.c_float), _p(out, C.c_float))
        return out.astype(np.complex128)

    # -- mapping / detection --------------------------------------------------------------------
    def qam_map(self, md, bits):
        bits = _u8(bits).reshape(-1)
        bps = int(np.log2(md.constel_size))
        n_sym = bits.size // bps
        out = np.empty(n_sym, dtype=np.complex64)
        self._call(self._lib.mmimo_qam_map, C.byref(md), n_sym, _p(bits, C.c_uint8), _p(out, C.c_float))
        return out.astype(np.complex128)

    def qam_detect(self, md, symbols, want_symbols=True, want_bits=True):
        s = _c64(symbols).reshape(-1)
        bps = int(np.log2(md.constel_size))
        sym = np.empty(s.size, dtype=np.complex64) if want_symbols else None
        bits = np.empty(s.size * bps, dtype=np.uint8) if want_bits else None
        self._call(self._lib.mmimo_qam_detect, C.byref(md), s.size, _p(s, C.c_float), _p(sym, C.c_float), _p(bits, C.c_uint8))
        return (sym.astype(np.complex128) if want_symbols else None, bits.view(np.int8) if want_bits else None)

    def welch_psd(self, x, fs, nfft, nperseg):
        """scipy.signal.welch(x, fs=fs, nfft=nfft, nperseg=nperseg, return_onesided=False) -> (freqs, psd), float64"""
        x = _c64(x).reshape(-1)
        freqs = np.empty(int(nfft), dtype=np.float64)
        psd = np.empty(int(nfft), dtype=np.float64)
        self._call(self._lib.mmimo_welch_psd, x.size, _p(x, C.c_float), int(nperseg), int(nfft), float(fs), _p(freqs, C.c_double),
                   _p(psd, C.c_double))
        return freqs, psd

    def qam_soft_llr(self, md, symbols, noise_var):
        """modulation.soft_decoding (modulation.py:29-59): float64 LLRs, MSB of each symbol first"""
        s = _c64(symbols).reshape(-1)
        nv = np.ascontiguousarray(np.atleast_1d(np.asarray(noise_var, dtype=np.float64)).reshape(-1))
        bps = int(np.log2(md.constel_size))
        out = np.empty(s.size * bps, dtype=np.float64)
        self._call(self._lib.mmimo_qam_soft_llr, C.byref(md), s.size, _p(s, C.c_float), _p(nv, C.c_double), nv.size, _p(out, C.c_double))
        return out

    def array_transmit(self, md, pa, n_ant, n_users, bits, precoder, skip_dist=False, time_domain=False, per_user=False,
                       return_both=False):
        bits = _u8(bits).reshape(n_users, -1)
        v = _c64(precoder).reshape(n_ant, n_users, md.n_sub_carr) if precoder is not None else None
        flags = (TX_SKIP_DIST if skip_dist else 0) | (TX_TIME_DOMAIN if time_domain else 0) | (TX_PER_USER if per_user else 0)
        g = n_users if (per_user and n_users > 1) else 1
        n_out = md.n_fft + (md.cp_len if time_domain else 0)
        out = np.empty((g, n_ant, n_out), dtype=np.complex64)
        clean = np.empty_like(out) if return_both else None
        self._call(self._lib.mmimo_array_transmit, C.byref(md), C.byref(pa) if pa is not None else None, int(n_ant), int(n_users),
                   _p(bits, C.c_uint8), _p(v, C.c_float), flags, _p(out, C.c_float), _p(clean, C.c_float))
        out = out.astype(np.complex128)
        if return_both:
            return out, clean.astype(np.complex128)
        return out

    def ofdm_demodulate(self, md, td, want_bits=True):
        td = _c64(td).reshape(-1)
        if td.size != md.n_fft + md.cp_len:
            raise ValueError("expected %d time-domain samples, got %d" % (md.n_fft + md.cp_len, td.size))
        bps = int(np.log2(md.constel_size))
        sym = np.empty(md.n_sub_carr, dtype=np.complex64)
        bits = np.empty(md.n_sub_carr * bps, dtype=np.uint8) if want_bits else None
        self._call(self._lib.mmimo_ofdm_demodulate, C.byref(md), _p(td, C.c_float), _p(sym, C.c_float), _p(bits, C.c_uint8))
        return sym.astype(np.complex128), (bits.view(np.int8) if want_bits else None)

    # -- channel side ---------------------------------------------------------------------------
    def propagate(self, x, h, sum_signals=True):
        x, h = _c64(x), _c64(h)
        if x.shape != h.shape:
            x, h = (np.ascontiguousarray(a) for a in np.broadcast_arrays(x, h))
        x2 = x.reshape(-1, x.shape[-1]) if x.ndim > 1 else x.reshape(1, -1)
        h2 = h.reshape(x2.shape)
        k, n = x2.shape
        out = np.empty(n if sum_signals else (k, n), dtype=np.complex64)
        self._call(self._lib.mmimo_propagate, k, n, _p(x2, C.c_float), _p(h2, C.c_float), int(bool(sum_signals)), _p(out, C.c_float))
        out = out.astype(np.complex128)
        return out if sum_signals else out.reshape(x.shape)

    def awgn_add(self, in_sig, unit_noise, scale):
        a, nz = _c64(in_sig), _c64(unit_noise)
        out = np.empty_like(a)
        self._call(self._lib.mmimo_awgn_add, a.size, _p(a, C.c_float), _p(nz, C.c_float), float(scale), _p(out, C.c_float))
        return out.astype(np.complex128)

    def count_mismatched_bits(self, a, b):
        a, b = _u8(a).reshape(-1), _u8(b).reshape(-1)
        if a.size != b.size:
            raise ValueError("bit arrays differ in length: %d vs %d" % (a.size, b.size))
        out = C.c_uint64(0)
        self._call(self._lib.mmimo_count_mismatched_bits, a.size, _p(a, C.c_uint8), _p(b, C.c_uint8), C.byref(out))
        return int(out.value)

    def channel_geometric(self, kind, tx_pos, tx_gain_db, rx_gain_db, rx_pos, freqs, skip_attenuation=False):
        tx_pos, tx_gain_db, rx_pos, freqs = _f64(tx_pos), _f64(tx_gain_db), _f64(rx_pos), _f64(freqs)
        k, n = tx_pos.shape[0], freqs.size
        out = np.empty((k, n), dtype=np.complex64)
        self._call(self._lib.mmimo_channel_geometric, int(kind), k, n, _p(tx_pos, C.c_double), _p(tx_gain_db, C.c_double),
                   float(rx_gain_db), _p(rx_pos, C.c_double), _p(freqs, C.c_double), int(bool(skip_attenuation)), _p(out, C.c_float))
        return out.astype(np.complex128)

    def channel_fspl(self, tx_pos, tx_gain_db, rx_gain_db, rx_pos, freqs, coeffs=None):
        tx_pos, tx_gain_db, rx_pos, freqs = _f64(tx_pos), _f64(tx_gain_db), _f64(rx_pos), _f64(freqs)
        k, n = tx_pos.shape[0], freqs.size
        co = _c64(coeffs).reshape(k, n) if coeffs is not None else None
        out = np.empty((k, n), dtype=np.complex64)
        self._call(self._lib.mmimo_channel_fspl, k, n, _p(tx_pos, C.c_double), _p(tx_gain_db, C.c_double), float(rx_gain_db),
                   _p(rx_pos, C.c_double), _p(freqs, C.c_double), _p(co, C.c_float), _p(out, C.c_float))
        return out.astype(np.complex128)

    def channel_random_paths(self, tx_pos, freqs, angles, taus):
        tx_pos, freqs, angles, taus = _f64(tx_pos), _f64(freqs), _f64(angles), _f64(taus)
        k, n = tx_pos.shape[0], freqs.size
        out = np.empty((k, n), dtype=np.complex64)
        self._call(self._lib.mmimo_channel_random_paths, k, n, _p(tx_pos, C.c_double), _p(freqs, C.c_double), angles.size,
                   _p(angles, C.c_double), _p(taus, C.c_double), _p(out, C.c_float))
        return out.astype(np.complex128)

    # -- precoding / AGC ------------------------------------------------------------------------
    def precoding(self, mode, h_list, n_sub_carr, sep_carriers=False):
        """h_list: [U][K][n_fft] -> [K][U][n_sc] (or [K][n_sc] with sep_carriers)."""
        h = _c64(h_list)
        if h.ndim == 2:
            h = h[None]
        u, k, n = h.shape
        v = np.empty((k, n_sub_carr) if sep_carriers else (k, u, n_sub_carr), dtype=np.complex64)
        self._call(self._lib.mmimo_precoding, int(mode), int(bool(sep_carriers)), u, k, n, int(n_sub_carr), _p(h, C.c_float),
                   _p(v, C.c_float))
        return v.astype(np.complex128)

    def agc(self, h, v, usr_idx, ibo_db, n_sub_carr, alpha_fixed=None):
        """Returns dict(agc_nfft, hkvk_nfft, scalers(2), vk_pow[K], alpha[K])."""
        h = _c64(h)
        k, n = h.shape
        v = _c64(v).reshape(k, -1, n_sub_carr)
        u = v.shape[1]
        agc = np.empty(n, dtype=np.complex64)
        hv = np.empty(n, dtype=np.complex64)
        sc = np.zeros(2, dtype=np.float64)
        pw = np.zeros(k, dtype=np.float64)
        al = np.zeros(k, dtype=np.float32)
        af = C.c_double(float(alpha_fixed)) if alpha_fixed is not None else None
        self._call(self._lib.mmimo_agc, k, u, int(usr_idx), n, int(n_sub_carr), _p(h, C.c_float), _p(v, C.c_float), float(ibo_db),
                   C.byref(af) if af is not None else None, _p(agc, C.c_float), _p(hv, C.c_float), _p(sc, C.c_double),
                   _p(pw, C.c_double), _p(al, C.c_float))
        return dict(agc_nfft=agc.astype(np.complex128), hkvk_nfft=hv.astype(np.complex128), scalers=sc, vk_pow=pw,
                    alpha=al.astype(np.float64))

    # -- receivers ------------------------------------------------------------------------------
    def cnc_receive(self, md, pa, alpha, n_iter, in_sig_fd, other_usr_symbols=None, want_bits=True, want_corr=False):
        x = _c64(in_sig_fd).reshape(-1)
        if x.size != md.n_fft:
            raise ValueError("in_sig_fd must hold n_fft = %d samples, got %d" % (md.n_fft, x.size))
        bps = int(np.log2(md.constel_size))
        other = _c64(other_usr_symbols).reshape(md.n_sub_carr) if other_usr_symbols is not None else None
        bits = np.empty((n_iter + 1, md.n_sub_carr * bps), dtype=np.uint8) if want_bits else None
        corr = np.empty((n_iter + 1, md.n_sub_carr), dtype=np.complex64) if want_corr else None
        self._call(self._lib.mmimo_cnc_receive, C.byref(md), C.byref(pa), float(alpha), int(n_iter), _p(x, C.c_float),
                   _p(other, C.c_float), _p(bits, C.c_uint8), _p(corr, C.c_float))
        return (bits.view(np.int8) if want_bits else None, corr.astype(np.complex128) if want_corr else None)

    def mcnc_receive(self, md, pa, h, v, agc_nfft, n_iter, in_sig_fd, usr_idx=0, other_usr_bits=None, want_bits=True,
                     want_corr=False):
        h = _c64(h)
        k, n = h.shape
        v = _c64(v).reshape(k, -1, md.n_sub_carr)
        u = v.shape[1]
        x = _c64(in_sig_fd).reshape(-1)
        if x.size != md.n_fft or n != md.n_fft:
            raise ValueError("in_sig_fd / channel must hold n_fft = %d samples" % md.n_fft)
        agc = _c64(agc_nfft).reshape(md.n_fft)
        bps = int(np.log2(md.constel_size))
        ob = _u8(other_usr_bits).reshape(u - 1, md.n_sub_carr * bps) if u > 1 else None
        bits = np.empty((n_iter + 1, md.n_sub_carr * bps), dtype=np.uint8) if want_bits else None
        corr = np.empty((n_iter + 1, md.n_sub_carr), dtype=np.complex64) if want_corr else None
        self._call(self._lib.mmimo_mcnc_receive, C.byref(md), C.byref(pa) if pa is not None else None, k, u, int(usr_idx),
                   _p(h, C.c_float), _p(v, C.c_float), _p(agc, C.c_float), int(n_iter), _p(x, C.c_float), _p(ob, C.c_uint8),
                   _p(bits, C.c_uint8), _p(corr, C.c_float))
        return (bits.view(np.int8) if want_bits else None, corr.astype(np.complex128) if want_corr else None)


_OPS = {}


def get_ops():
    """Process-wide StageOps on the GPU named by MMIMO_DEVICE / LOCAL_RANK (default 0).  Created lazily in the
    process that uses it (a CUDA context does not survive fork, so a forked worker gets its own)."""
    from . import _lib
    pid = os.getpid()
    ops = _OPS.get(pid)
    if ops is None:
        lib = _lib.load()
        h = C.c_void_p()
        dev = os.environ.get("MMIMO_DEVICE", os.environ.get("LOCAL_RANK"))
        if dev is None:
            n = lib.mmimo_device_count()
            if n <= 0:
                raise _lib.MmimoError("no CUDA device visible: %s" % lib.mmimo_last_error().decode("utf-8", "replace"))
            dev = os.getpid() % n
        dev = int(dev)
        _lib.check(lib, lib.mmimo_create(C.byref(h), dev))
        ops = StageOps(lib, h, _lib.check)
        _OPS.clear()
        _OPS[pid] = ops
    return ops
