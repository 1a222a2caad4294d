"""Host-side handle on one GPU context of libmmimo_b200.so: the fused Monte-Carlo link
(`Link.simulate`, mp_model.py:89-222) for batches of frames.  numpy in / numpy out; all compute is CUDA."""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import _lib

C_LIGHT = 299792458.0  # scipy.constants.c

PA_KINDS = {"none": 0, "softlim": 1, "rapp": 2, "toi": 3}
CHAN_KINDS = {"los": 0, "two_path": 1, "rayleigh": 2, "injected": 3}
RX_KINDS = {"cnc": 0, "mcnc": 1}
N_FFT, N_SC = 4096, 2048


def ula_positions(n_elements: int, center_freq: float, wav_len_spacing: float, cord_x=0.0, cord_y=0.0, cord_z=15.0):
    """Element coordinates of LinearArray (antenna_array.py:436-445): ULA along X centred on cord_x."""
    lam = C_LIGHT / center_freq
    half = (n_elements - 1) * wav_len_spacing * lam / 2
    x = np.linspace(-half, half, n_elements) + cord_x
    return np.stack((x, np.full(n_elements, float(cord_y)), np.full(n_elements, float(cord_z))), axis=1)


def subcarrier_freqs(n_fft: int, carrier_spacing: float, center_freq: float) -> np.ndarray:
    """Sub-carrier frequencies exactly as channel.py:51-52 produces them: torch.fft.fftfreq is float32 and
    stays float32 through `* spacing + fc`, i.e. the values sit on the float32 grid (256 Hz at 3.5 GHz)."""
    k = np.fft.fftfreq(n_fft, d=1.0 / n_fft).astype(np.float32)
    f32 = k * np.float32(carrier_spacing) + np.float32(center_freq)
    return f32.astype(np.float64)


def _ptr(arr, ctype):
    return arr.ctypes.data_as(C.POINTER(ctype)) if arr is not None else None


def _c64(a, shape):
    a = np.ascontiguousarray(np.asarray(a, dtype=np.complex64))
    if a.shape != tuple(shape):
        raise ValueError("expected shape %s, got %s" % (tuple(shape), a.shape))
    return a


class LinkEngine:
    """One mmimo_ctx (one GPU).  Not thread-safe; create it in the process that uses it (CUDA contexts do
    not survive fork)."""

    def __init__(self, device: int = 0):
        self._lib = _lib.load()
        if device < 0:  # spread worker processes over the visible GPUs
            import os
            n = self._lib.mmimo_device_count()
            if n <= 0:
                raise _lib.MmimoError("no CUDA device visible: %s" % self._lib.mmimo_last_error().decode("utf-8", "replace"))
            device = os.getpid() % n
        h = C.c_void_p()
        _lib.check(self._lib, self._lib.mmimo_create(C.byref(h), int(device)))
        self._h = h
        self.device = int(device)
        self.cfg = None
        self._keep = []

    def close(self):
        if getattr(self, "_h", None):
            self._lib.mmimo_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------------------------------
    def configure(self, *, constel_size=64, n_fft=N_FFT, n_sub_carr=N_SC, n_ant=64, pa="softlim", ibo_db=0.0,
                  p_hardness=4.0, channel="los", receiver="mcnc", skip_attenuation=False, center_freq=3.5e9,
                  carrier_spacing=15e3, wav_len_spacing=0.5, array_pos=(0.0, 0.0, 15.0), rx_pos=(212.0, 212.0, 1.5),
                  rx_loc_var=10.0, ant_gain_db=0.0, tx_pos=None, freqs=None):
        if tx_pos is None:
            tx_pos = ula_positions(n_ant, center_freq, wav_len_spacing, *array_pos)
        tx_pos = np.ascontiguousarray(np.asarray(tx_pos, dtype=np.float64).reshape(n_ant, 3))
        if freqs is None:
            freqs = subcarrier_freqs(n_fft, carrier_spacing, center_freq)
        freqs = np.ascontiguousarray(np.asarray(freqs, dtype=np.float64).reshape(n_fft))
        cfg = _lib.LinkConfig()
        cfg.constel_size, cfg.n_fft, cfg.n_sub_carr, cfg.n_ant = int(constel_size), int(n_fft), int(n_sub_carr), int(n_ant)
        cfg.pa_kind = PA_KINDS[pa] if isinstance(pa, str) else int(pa)
        cfg.channel_kind = CHAN_KINDS[channel] if isinstance(channel, str) else int(channel)
        cfg.receiver_kind = RX_KINDS[receiver] if isinstance(receiver, str) else int(receiver)
        cfg.skip_attenuation = int(bool(skip_attenuation))
        cfg.ibo_db, cfg.p_hardness = float(ibo_db), float(p_hardness)
        cfg.center_freq, cfg.carrier_spacing = float(center_freq), float(carrier_spacing)
        cfg.rx_pos = (C.c_double * 3)(*[float(v) for v in rx_pos])
        cfg.rx_loc_var, cfg.ant_gain_db = float(rx_loc_var), float(ant_gain_db)
        cfg.tx_pos = _ptr(tx_pos, C.c_double)
        cfg.freqs = _ptr(freqs, C.c_double)
        _lib.check(self._lib, self._lib.mmimo_configure(self._h, C.byref(cfg)))
        self.cfg = dict(constel_size=int(constel_size), n_fft=int(n_fft), n_sub_carr=int(n_sub_carr), n_ant=int(n_ant),
                        pa=pa, ibo_db=float(ibo_db), channel=channel, receiver=receiver, tx_pos=tx_pos, freqs=freqs)
        self.bits_per_frame = int(math.log2(constel_size)) * n_sub_carr
        return self

    def set_ibo(self, ibo_db: float):
        _lib.check(self._lib, self._lib.mmimo_set_ibo(self._h, float(ibo_db)))
        self.cfg["ibo_db"] = float(ibo_db)

    # ------------------------------------------------------------------------------------------
    def _request(self, n_frames, n_iter, snr_db, seed, frame_id0, incl_clean_run, csi_eps, inject, noise_unit=False):
        K = self.cfg["n_ant"]
        rq = _lib.SimRequest()
        rq.n_frames, rq.n_iter = int(n_frames), int(n_iter)
        rq.incl_clean_run = 2 if incl_clean_run == "only" else int(bool(incl_clean_run))
        rq.flags = 1 if noise_unit else 0
        rq.snr_db, rq.csi_eps = float(snr_db), float(csi_eps or 0.0)
        rq.seed, rq.frame_id0 = int(seed) & (2 ** 64 - 1), int(frame_id0)
        keep = []
        inject = inject or {}
        unknown = set(inject) - {"rx_pos", "h_true", "h_loop", "tx_idx", "noise", "clean_rx_pos", "clean_h_true",
                                 "clean_h_loop", "clean_tx_idx", "clean_noise", "csi_noise", "clean_csi_noise"}
        if unknown:
            raise ValueError("unknown injected inputs: %s" % sorted(unknown))
        for pre in ("", "clean_"):
            v = inject.get(pre + "rx_pos")
            if v is not None:
                a = np.ascontiguousarray(np.asarray(v, dtype=np.float64).reshape(n_frames, 3))
                keep.append(a)
                setattr(rq, pre + "rx_pos", _ptr(a, C.c_double))
            for name in ("h_true", "h_loop", "csi_noise"):
                v = inject.get(pre + name)
                if v is not None:
                    a = _c64(v, (n_frames, K, N_SC))
                    keep.append(a)
                    setattr(rq, pre + name, _ptr(a, C.c_float))
            v = inject.get(pre + "tx_idx")
            if v is not None:
                a = np.ascontiguousarray(np.asarray(v, dtype=np.uint8).reshape(n_frames, N_SC))
                keep.append(a)
                setattr(rq, pre + "tx_idx", _ptr(a, C.c_uint8))
            v = inject.get(pre + "noise")
            if v is not None:
                a = _c64(v, (n_frames, N_SC))
                keep.append(a)
                setattr(rq, pre + "noise", _ptr(a, C.c_float))
        return rq, keep

    def simulate(self, n_frames, n_iter, snr_db, *, seed=0, frame_id0=0, incl_clean_run=False, csi_eps=0.0,
                 inject=None, dumps=(), noise_unit=False):
        """Run `n_frames` frames; returns dict(err=[n_iter+1], bits=..., clean_err, + requested dumps).

        dumps: any of det, err_per_frame, rx_eq, corr, h_true, v, agc, pa_p0, alpha, rx_pos, clean_det,
        clean_err_per_frame.  incl_clean_run: False | True | "only" (just the no-distortion frames).
        noise_unit: injected noise arrays hold raw standard normals, scaled on the device (noise.py:64-66)."""
        if self.cfg is None:
            raise _lib.MmimoError("LinkEngine.configure() has not been called")
        K, I = self.cfg["n_ant"], int(n_iter)
        rq, keep = self._request(n_frames, n_iter, snr_db, seed, frame_id0, incl_clean_run, csi_eps, inject, noise_unit)
        rs = _lib.SimResult()
        out = {"err": np.zeros(I + 1, dtype=np.uint64), "clean_err": np.zeros(1, dtype=np.uint64)}
        rs.err, rs.clean_err = _ptr(out["err"], C.c_uint64), _ptr(out["clean_err"], C.c_uint64)
        shapes = {"det": ((n_frames, I + 1, N_SC), np.uint8, C.c_uint8),
                  "err_per_frame": ((n_frames, I + 1), np.uint32, C.c_uint32),
                  "rx_eq": ((n_frames, N_SC), np.complex64, C.c_float),
                  "corr": ((n_frames, I + 1, N_SC), np.complex64, C.c_float),
                  "h_true": ((n_frames, K, N_SC), np.complex64, C.c_float),
                  "v": ((n_frames, K, N_SC), np.complex64, C.c_float),
                  "agc": ((n_frames, N_SC), np.complex64, C.c_float),
                  "pa_p0": ((n_frames,), np.float32, C.c_float),
                  "alpha": ((n_frames, K), np.float32, C.c_float),
                  "rx_pos": ((n_frames, 3), np.float64, C.c_double),
                  "clean_det": ((n_frames, N_SC), np.uint8, C.c_uint8),
                  "clean_err_per_frame": ((n_frames,), np.uint32, C.c_uint32)}
        for name in dumps:
            shape, dt, ct = shapes[name]
            out[name] = np.zeros(shape, dtype=dt)
            setattr(rs, name, _ptr(out[name], ct))
        _lib.check(self._lib, self._lib.mmimo_link_simulate(self._h, C.byref(rq), C.byref(rs)))
        del keep
        out["bits_per_frame"] = int(rs.bits_per_frame)
        out["bits"] = int(rs.bits_per_frame) * int(n_frames)
        return out

    def sdr(self, n_frames, *, seed=0, frame_id0=0, inject=None):
        """Signal-to-distortion snapshots of the configured link (main_beampatterns_plotting/sdr_mp.py:124-151):
        returns (desired[n_frames], distortion[n_frames]) powers; SDR per snapshot = desired / distortion.
        inject: optional rx_pos [n,3], h_true [n,K,2048], tx_idx [n,2048]."""
        if self.cfg is None:
            raise _lib.MmimoError("LinkEngine.configure() has not been called")
        K = self.cfg["n_ant"]
        inject = inject or {}
        rq = _lib.SdrRequest()
        rq.n_frames, rq.seed, rq.frame_id0 = int(n_frames), int(seed) & (2 ** 64 - 1), int(frame_id0)
        keep = []
        if inject.get("rx_pos") is not None:
            a = np.ascontiguousarray(np.asarray(inject["rx_pos"], dtype=np.float64).reshape(n_frames, 3)); keep.append(a)
            rq.rx_pos = _ptr(a, C.c_double)
        if inject.get("h_true") is not None:
            a = _c64(inject["h_true"], (n_frames, K, N_SC)); keep.append(a)
            rq.h_true = _ptr(a, C.c_float)
        if inject.get("tx_idx") is not None:
            a = np.ascontiguousarray(np.asarray(inject["tx_idx"], dtype=np.uint8).reshape(n_frames, N_SC)); keep.append(a)
            rq.tx_idx = _ptr(a, C.c_uint8)
        des = np.zeros(n_frames, dtype=np.float64)
        dis = np.zeros(n_frames, dtype=np.float64)
        _lib.check(self._lib, self._lib.mmimo_sdr_simulate(self._h, C.byref(rq), _ptr(des, C.c_double), _ptr(dis, C.c_double)))
        return des, dis

    def sweep(self, snr_db, ibo_db, frames_per_point, n_iter, *, seed=0, frame_id0=0, csi_eps=0.0):
        """Multi-point batch (mmimo_link_sweep): point p runs frames [p*fpp, (p+1)*fpp) of one batch at snr_db[p] /
        ibo_db[p] (either may be None = the configured value / a scalar broadcast).  Returns err [n_points, n_iter+1].
        The inner (IBO, Eb/N0) loops of main_mp_miso_mcnc_constant_ber_req_ebn0_vs_ibo.py:116-150 in one launch chain."""
        if self.cfg is None:
            raise _lib.MmimoError("LinkEngine.configure() has not been called")
        snr = None if snr_db is None else np.atleast_1d(np.asarray(snr_db, dtype=np.float64))
        ibo = None if ibo_db is None else np.atleast_1d(np.asarray(ibo_db, dtype=np.float64))
        n_pts = max(len(snr) if snr is not None else 1, len(ibo) if ibo is not None else 1)
        if snr is not None:
            snr = np.ascontiguousarray(np.broadcast_to(snr, (n_pts,)))
        if ibo is not None:
            ibo = np.ascontiguousarray(np.broadcast_to(ibo, (n_pts,)))
        rq, _keep = self._request(0, n_iter, 0.0 if snr is None else float(snr[0]), seed, frame_id0, False, csi_eps, None)
        err = np.zeros((n_pts, int(n_iter) + 1), dtype=np.uint64)
        _lib.check(self._lib, self._lib.mmimo_link_sweep(self._h, C.byref(rq), n_pts, int(frames_per_point),
                                                         _ptr(snr, C.c_double), _ptr(ibo, C.c_double), _ptr(err, C.c_uint64)))
        return err

    def mu_simulate(self, n_frames, n_iter, snr_db, *, precoding="mrt", usr_pos=((212.0, 212.0, 1.5), (150.0, 250.0, 1.5)),
                    loop_single_user=False, receiver=None, seed=0, frame_id0=0, inject=None, dumps=()):
        """Fused two-user frames (main_multiuser/main_multiuser_mcnc_ber_vs_ebn0.py:514-642, corrector.py:348-489):
        n_frames channel realisations, one frame per user each.  precoding "mrt" | "zf" (antenna_array.py:202-257).
        receiver: "mcnc_mu" (McncMuReceiver, default) | "mcnc_su" (single-user McncReceiver, = loop_single_user=True) |
        "cnc" (CncReceiver) | "cnc_mu" (CncMuReceiver, corrector.py:248-345).
        inject: optional rx_pos [n,2,3], h [n,2,K,2048], tx_idx [n,2,2048], noise [n,2,2048], rx_eq_in [n,2,2048].
        dumps: any of "det", "rx_eq", "v", "agc", "pa_p0".  Returns dict with err [2, n_iter+1], bits (per user)."""
        if self.cfg is None:
            raise _lib.MmimoError("LinkEngine.configure() has not been called")
        K = self.cfg["n_ant"]
        inject = inject or {}
        n = int(n_frames)
        rq = _lib.MuRequest()
        rq.n_frames, rq.n_iter, rq.snr_db = n, int(n_iter), float(snr_db)
        rq.precoding = {"mrt": 0, "zf": 1}[precoding]
        receiver = receiver or ("mcnc_su" if loop_single_user else "mcnc_mu")
        rq.receiver = {"mcnc_mu": 0, "mcnc_su": 1, "cnc": 2, "cnc_mu": 3}[receiver]
        rq.seed, rq.frame_id0 = int(seed) & (2 ** 64 - 1), int(frame_id0)
        for u in range(2):
            for i in range(3):
                rq.usr_pos[u][i] = float(usr_pos[u][i])
        keep = []
        if inject.get("rx_pos") is not None:
            a = np.ascontiguousarray(np.asarray(inject["rx_pos"], dtype=np.float64).reshape(n, 2, 3)); keep.append(a)
            rq.rx_pos = _ptr(a, C.c_double)
        if inject.get("h") is not None:
            a = _c64(inject["h"], (n, 2, K, N_SC)); keep.append(a)
            rq.h = _ptr(a, C.c_float)
        if inject.get("tx_idx") is not None:
            a = np.ascontiguousarray(np.asarray(inject["tx_idx"], dtype=np.uint8).reshape(n, 2, N_SC)); keep.append(a)
            rq.tx_idx = _ptr(a, C.c_uint8)
        for key in ("noise", "rx_eq_in"):
            if inject.get(key) is not None:
                a = _c64(inject[key], (n, 2, N_SC)); keep.append(a)
                setattr(rq, key, _ptr(a, C.c_float))
        rs = _lib.MuResult()
        err = np.zeros((2, int(n_iter) + 1), dtype=np.uint64)
        rs.err = _ptr(err, C.c_uint64)
        out = {}
        if "det" in dumps:
            out["det"] = np.zeros((n, 2, int(n_iter) + 1, N_SC), dtype=np.uint8); rs.det = _ptr(out["det"], C.c_uint8)
        if "rx_eq" in dumps:
            out["rx_eq"] = np.zeros((n, 2, N_SC), dtype=np.complex64); rs.rx_eq = _ptr(out["rx_eq"], C.c_float)
        if "v" in dumps:
            out["v"] = np.zeros((n, 2, K, N_SC), dtype=np.complex64); rs.v = _ptr(out["v"], C.c_float)
        if "agc" in dumps:
            out["agc"] = np.zeros((n, 2, N_SC), dtype=np.complex64); rs.agc = _ptr(out["agc"], C.c_float)
        if "pa_p0" in dumps:
            out["pa_p0"] = np.zeros(n, dtype=np.float32); rs.pa_p0 = _ptr(out["pa_p0"], C.c_float)
        _lib.check(self._lib, self._lib.mmimo_mu_link_simulate(self._h, C.byref(rq), C.byref(rs)))
        out["err"] = err
        out["bits"] = np.full(2, n * int(rs.bits_per_frame), dtype=np.uint64)
        return out

    def enqueue(self, n_frames, n_iter, snr_db, *, seed=0, frame_id0=0, incl_clean_run=False, csi_eps=0.0,
                inject=None, noise_unit=False):
        """Asynchronous: H2D of injected inputs + kernels, counters accumulate on the device until fetch().
        Injected host arrays are kept alive until the next fetch()/synchronize()."""
        rq, keep = self._request(n_frames, n_iter, snr_db, seed, frame_id0, incl_clean_run, csi_eps, inject, noise_unit)
        if len(self._keep) >= 32:  # a caller that never fetches: drain once in a while instead of growing without bound
            self.synchronize()
        self._keep.append(keep)
        _lib.check(self._lib, self._lib.mmimo_link_enqueue(self._h, C.byref(rq)))

    def fetch(self, n_iter):
        err = np.zeros(int(n_iter) + 1, dtype=np.uint64)
        clean = np.zeros(1, dtype=np.uint64)
        _lib.check(self._lib, self._lib.mmimo_link_fetch(self._h, int(n_iter), _ptr(err, C.c_uint64), _ptr(clean, C.c_uint64)))
        self._keep.clear()
        return err, clean

    def set_stream(self, cuda_stream_ptr: int = 0):
        """Enqueue all further work on the given cudaStream_t (e.g. torch.cuda.current_stream().cuda_stream)."""
        _lib.check(self._lib, self._lib.mmimo_set_stream(self._h, C.c_void_p(int(cuda_stream_ptr) or None)))

    def counters_to_device(self, dst_dev_ptr: int, clear: bool = True):
        """Stream-ordered D2D copy of the 65 uint64 error counters into caller-owned device memory."""
        _lib.check(self._lib, self._lib.mmimo_counters_to_device(self._h, C.c_void_p(int(dst_dev_ptr)), int(bool(clear))))

    def synchronize(self):
        _lib.check(self._lib, self._lib.mmimo_synchronize(self._h))
        self._keep.clear()

    def profile(self, enable: bool):
        _lib.check(self._lib, self._lib.mmimo_profile(self._h, int(bool(enable))))

    def profile_read(self):
        """(ms per kind, launches per kind) for kinds (channel, prep, frame, other); clears the records."""
        ms = np.zeros(4, dtype=np.float64)
        n = np.zeros(4, dtype=np.uint64)
        _lib.check(self._lib, self._lib.mmimo_profile_read(self._h, _ptr(ms, C.c_double), _ptr(n, C.c_uint64)))
        return ms, n

    def measure_fp32_peak(self) -> float:
        v = C.c_double()
        _lib.check(self._lib, self._lib.mmimo_measure_fp32_peak(self._h, C.byref(v)))
        return float(v.value)

    STAGE_KERNELS = ("map_precode", "ifft", "pa_soft_limiter", "fft", "propagate")
    # algorithmic HBM bytes per row of 4096 complex64 samples: map reads the 16 KiB precoder row and writes the grid row;
    # the transforms and the PA read and write one row; propagate reads two rows and writes one
    STAGE_BYTES_PER_ROW = (16384 + 32768, 65536, 65536, 65536, 98304)

    def measure_stage_kernels(self, rows: int = 16384, reps: int = 5):
        """ms per launch of the five unfused building-block kernels on `rows` device-resident rows (mmimo_measure_stage_kernels)."""
        ms = np.zeros(5, np.float64)
        _lib.check(self._lib, self._lib.mmimo_measure_stage_kernels(self._h, int(rows), int(reps), _ptr(ms, C.c_double)))
        return dict(zip(self.STAGE_KERNELS, ms.tolist()))

    @property
    def stream_ptr(self) -> int:
        return int(self._lib.mmimo_stream(self._h) or 0)

    @property
    def launch_count(self) -> int:
        return int(self._lib.mmimo_launch_count(self._h))
