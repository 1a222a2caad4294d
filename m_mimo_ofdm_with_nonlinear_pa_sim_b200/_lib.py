"""ctypes binding of include/mmimo_b200.h.  Loading fails loudly: there is no CPU fallback."""
import ctypes as C
import os

from . import _build

c_f_p = C.POINTER(C.c_float)
c_d_p = C.POINTER(C.c_double)
c_u8_p = C.POINTER(C.c_uint8)
c_u32_p = C.POINTER(C.c_uint32)
c_u64_p = C.POINTER(C.c_uint64)


class LinkConfig(C.Structure):
    _fields_ = [("constel_size", C.c_int32), ("n_fft", C.c_int32), ("n_sub_carr", C.c_int32), ("n_ant", C.c_int32),
                ("pa_kind", C.c_int32), ("channel_kind", C.c_int32), ("receiver_kind", C.c_int32),
                ("skip_attenuation", C.c_int32), ("ibo_db", C.c_double), ("p_hardness", C.c_double),
                ("center_freq", C.c_double), ("carrier_spacing", C.c_double), ("rx_pos", C.c_double * 3),
                ("rx_loc_var", C.c_double), ("ant_gain_db", C.c_double), ("tx_pos", c_d_p), ("freqs", c_d_p)]


class SimRequest(C.Structure):
    _fields_ = [("n_frames", C.c_int32), ("n_iter", C.c_int32), ("incl_clean_run", C.c_int32), ("flags", C.c_int32),
                ("snr_db", C.c_double), ("csi_eps", C.c_double), ("seed", C.c_uint64), ("frame_id0", C.c_uint64),
                ("rx_pos", c_d_p), ("h_true", c_f_p), ("h_loop", c_f_p), ("tx_idx", c_u8_p), ("noise", c_f_p),
                ("clean_rx_pos", c_d_p), ("clean_h_true", c_f_p), ("clean_h_loop", c_f_p), ("clean_tx_idx", c_u8_p),
                ("clean_noise", c_f_p), ("csi_noise", c_f_p), ("clean_csi_noise", c_f_p)]


class SimResult(C.Structure):
    _fields_ = [("err", c_u64_p), ("clean_err", c_u64_p), ("bits_per_frame", C.c_uint64), ("det", c_u8_p),
                ("err_per_frame", c_u32_p), ("rx_eq", c_f_p), ("corr", c_f_p), ("h_true", c_f_p), ("v", c_f_p),
                ("agc", c_f_p), ("pa_p0", c_f_p), ("alpha", c_f_p), ("rx_pos", c_d_p), ("clean_det", c_u8_p), ("clean_err_per_frame", c_u32_p)]


class SdrRequest(C.Structure):
    _fields_ = [("n_frames", C.c_int32), ("reserved", C.c_int32), ("seed", C.c_uint64), ("frame_id0", C.c_uint64),
                ("rx_pos", c_d_p), ("h_true", c_f_p), ("tx_idx", c_u8_p)]


class MuRequest(C.Structure):
    _fields_ = [("n_frames", C.c_int32), ("n_iter", C.c_int32), ("precoding", C.c_int32), ("receiver", C.c_int32),
                ("snr_db", C.c_double), ("seed", C.c_uint64), ("frame_id0", C.c_uint64), ("usr_pos", (C.c_double * 3) * 2),
                ("rx_pos", c_d_p), ("h", c_f_p), ("tx_idx", c_u8_p), ("noise", c_f_p), ("rx_eq_in", c_f_p)]


class MuResult(C.Structure):
    _fields_ = [("err", c_u64_p), ("bits_per_frame", C.c_uint64), ("det", c_u8_p), ("rx_eq", c_f_p), ("v", c_f_p),
                ("agc", c_f_p), ("pa_p0", c_f_p)]


class MmimoError(RuntimeError):
    pass


_LIB = None

# every symbol include/mmimo_b200.h declares (tests/test_abi.py checks the header against this list)
EXPORTS = ["mmimo_device_count", "mmimo_create", "mmimo_destroy", "mmimo_last_error", "mmimo_version", "mmimo_stream", "mmimo_synchronize",
           "mmimo_set_stream", "mmimo_counters_to_device",
           "mmimo_configure", "mmimo_set_ibo", "mmimo_sdr_simulate", "mmimo_mu_link_simulate", "mmimo_link_sweep", "mmimo_link_simulate", "mmimo_link_enqueue", "mmimo_link_fetch",
           "mmimo_launch_count", "mmimo_profile", "mmimo_profile_read", "mmimo_measure_fp32_peak", "mmimo_measure_stage_kernels",
           # stage-level operations (see _stage_abi.STAGE_EXPORTS)
           "mmimo_fft", "mmimo_welch_psd", "mmimo_welch_psd_noalloc", "mmimo_pa_process", "mmimo_qam_map", "mmimo_qam_detect", "mmimo_qam_soft_llr",
           "mmimo_array_transmit", "mmimo_ofdm_demodulate", "mmimo_propagate", "mmimo_awgn_add", "mmimo_count_mismatched_bits",
           "mmimo_channel_geometric", "mmimo_channel_fspl", "mmimo_channel_random_paths", "mmimo_precoding",
           "mmimo_agc", "mmimo_cnc_receive", "mmimo_mcnc_receive"]


def lib_path() -> str:
    """The in-tree library; MMIMO_LIB points at a measurement variant built by tools/build_variants.py instead."""
    return os.environ.get("MMIMO_LIB") or _build.LIB_PATH


def load(build_if_missing: bool = True):
    """dlopen libmmimo_b200.so (building it in-tree with nvcc when absent or stale and nvcc exists)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if build_if_missing and not os.environ.get("MMIMO_LIB") and _build.is_stale():
        try:
            _build.build()
        except Exception as exc:  # no nvcc on this box: a prebuilt .so is still fine
            if not os.path.isfile(path):
                raise MmimoError("libmmimo_b200.so is missing and could not be built: %s" % exc) from exc
            import warnings
            warnings.warn("libmmimo_b200.so is older than its sources and could not be rebuilt (%s): using the stale "
                          "library" % str(exc).splitlines()[0])
    if not os.path.isfile(path):
        raise MmimoError("libmmimo_b200.so not found at %s - run `python __graft_entry__.py build`" % path)
    lib = C.CDLL(path)
    lib.mmimo_device_count.restype = C.c_int
    lib.mmimo_create.argtypes = [C.POINTER(C.c_void_p), C.c_int]
    lib.mmimo_create.restype = C.c_int
    lib.mmimo_destroy.argtypes = [C.c_void_p]
    lib.mmimo_destroy.restype = None
    lib.mmimo_last_error.restype = C.c_char_p
    lib.mmimo_version.restype = C.c_char_p
    lib.mmimo_stream.argtypes = [C.c_void_p]
    lib.mmimo_stream.restype = C.c_void_p
    lib.mmimo_synchronize.argtypes = [C.c_void_p]
    lib.mmimo_set_stream.argtypes = [C.c_void_p, C.c_void_p]
    lib.mmimo_counters_to_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    lib.mmimo_configure.argtypes = [C.c_void_p, C.POINTER(LinkConfig)]
    lib.mmimo_set_ibo.argtypes = [C.c_void_p, C.c_double]
    lib.mmimo_link_simulate.argtypes = [C.c_void_p, C.POINTER(SimRequest), C.POINTER(SimResult)]
    lib.mmimo_link_enqueue.argtypes = [C.c_void_p, C.POINTER(SimRequest)]
    lib.mmimo_link_fetch.argtypes = [C.c_void_p, C.c_int32, c_u64_p, c_u64_p]
    lib.mmimo_sdr_simulate.argtypes = [C.c_void_p, C.POINTER(SdrRequest), c_d_p, c_d_p]
    lib.mmimo_mu_link_simulate.argtypes = [C.c_void_p, C.POINTER(MuRequest), C.POINTER(MuResult)]
    lib.mmimo_mu_link_simulate.restype = C.c_int
    # device pointers as integers (size_t n, x_dev, nperseg, nfft, fs, psd_dev, stream)
    lib.mmimo_welch_psd_noalloc.argtypes = [C.c_size_t, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_void_p]
    lib.mmimo_welch_psd_noalloc.restype = C.c_int
    lib.mmimo_link_sweep.argtypes = [C.c_void_p, C.POINTER(SimRequest), C.c_int32, C.c_int32, c_d_p, c_d_p, c_u64_p]
    lib.mmimo_link_sweep.restype = C.c_int
    lib.mmimo_launch_count.argtypes = [C.c_void_p]
    lib.mmimo_launch_count.restype = C.c_uint64
    lib.mmimo_profile.argtypes = [C.c_void_p, C.c_int]
    lib.mmimo_profile_read.argtypes = [C.c_void_p, c_d_p, c_u64_p]
    lib.mmimo_measure_fp32_peak.argtypes = [C.c_void_p, c_d_p]
    lib.mmimo_measure_stage_kernels.argtypes = [C.c_void_p, C.c_int, C.c_int, c_d_p]
    _bind_stage_ops(lib)
    _LIB = lib
    return lib


def _bind_stage_ops(lib):
    """argtypes of the stage-level entry points."""
    from . import _stage_abi
    _stage_abi.bind(lib)


def check(lib, rc):
    if rc != 0:
        raise MmimoError(lib.mmimo_last_error().decode("utf-8", "replace"))
