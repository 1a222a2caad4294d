"""Drop-in for the reference's utilities.py: same function names and argument meaning.

The per-sample numeric helpers on the link path (FFT wrappers, bit-error count) run on the GPU through
libmmimo_b200.so; bit <-> integer helpers are vectorised numpy (pure index bookkeeping); geometry, CSV and
progress-bar helpers are plain Python like the reference's.  Plotting helpers import matplotlib lazily."""
import csv
import sys

import numpy as np
from numpy import ndarray

from m_mimo_ofdm_with_nonlinear_pa_sim_b200._stage_abi import get_ops


# ---- bit helpers (utilities.py:18-67) ------------------------------------------------------------
def dec2bitarray(in_number, bit_width: int) -> ndarray:
    """Integers -> int8 bit array, MSB first (utilities.py:18-51)."""
    if isinstance(in_number, (np.integer, int)):
        return decimal2bitarray(int(in_number), bit_width)
    num = np.asarray(in_number).astype(np.int64).reshape(-1)
    shifts = np.arange(bit_width - 1, -1, -1, dtype=np.int64)
    return ((num[:, None] >> shifts[None, :]) & 1).astype(np.int8).reshape(-1)


def decimal2bitarray(number: int, bit_width: int) -> ndarray:
    shifts = np.arange(bit_width - 1, -1, -1, dtype=np.int64)
    return ((int(number) >> shifts) & 1).astype(np.int8)


def bitarray2dec(in_bitarray) -> int:
    """Bit array (MSB first) -> integer (utilities.py:54-67)."""
    number = 0
    for b in np.asarray(in_bitarray).reshape(-1):
        number = (number << 1) + int(b)
    return number


# ---- powers / conversions (utilities.py:70-143) --------------------------------------------------
def td_signal_power(signal: ndarray) -> float:
    return np.mean(np.abs(signal) ** 2)


def fd_signal_power(signal: ndarray) -> float:
    return np.sum(np.abs(signal) ** 2)


def count_mismatched_bits(tx_bits_arr: ndarray, rx_bits_arr: ndarray) -> int:
    """Number of differing bits (utilities.py:94-104), counted on the GPU."""
    return get_ops().count_mismatched_bits(tx_bits_arr, rx_bits_arr)


def ebn0_to_snr(eb_per_n0, n_fft: int, n_sub_carr: int, constel_size: int):
    return 10 * np.log10(10 ** (eb_per_n0 / 10) * n_sub_carr * np.log2(constel_size) / n_fft)


def snr_to_ebn0(snr, n_fft: int, n_sub_carr: int, constel_size: int):
    return 10 * np.log10(10 ** (snr / 10) * (n_fft / (n_sub_carr * np.log2(constel_size))))


def to_db(samples):
    return 10 * np.log10(samples)


# ---- geometry (utilities.py:146-192) -------------------------------------------------------------
def pts_on_circum(radius: float, n_points: int = 100) -> list:
    ang = 2 * np.pi / n_points * np.arange(0, n_points + 1)
    return [(np.cos(a) * radius, np.sin(a) * radius) for a in ang]


def pts_on_semicircum(radius: float, n_points: int = 100) -> list:
    ang = np.pi / n_points * np.arange(0, n_points + 1)
    return [(np.cos(a) * radius, np.sin(a) * radius) for a in ang]


def pts_on_semisphere(radius: float, n_points: int = 100, center_x: float = 0, center_y: float = 0,
                      center_z: float = 0):
    side = int(np.sqrt(n_points))
    az_vec = np.deg2rad(np.linspace(0, 180, side, endpoint=True))
    el_vec = np.deg2rad(np.linspace(0, 180, side, endpoint=True))
    pts = []
    for az in az_vec:
        for el in el_vec:
            pts.append((-radius * np.sin(el) * np.cos(az) + center_x, -radius * np.sin(el) * np.sin(az) + center_y,
                        -radius * np.cos(el) + center_z))
    return pts


# ---- transforms (utilities.py:311-339) -----------------------------------------------------------
def to_freq_domain(in_sig_td: ndarray, remove_cp: bool = True, cp_len: int = None) -> ndarray:
    """Row-wise orthonormal FFT, optionally dropping the cyclic prefix first."""
    skip = 0
    if remove_cp:
        if cp_len is None:
            raise TypeError("cp_len is required with remove_cp=True")
        skip = int(cp_len)
    return get_ops().fft(np.asarray(in_sig_td), inverse=False, skip=skip)


def to_time_domain(in_sig_mat_fd: ndarray) -> ndarray:
    return get_ops().fft(np.asarray(in_sig_mat_fd), inverse=True)


# ---- result files / console (utilities.py:342-392) -----------------------------------------------
def welch(x, fs: float = 1.0, nfft: int = None, nperseg: int = None, return_onesided: bool = False):
    """Power spectral density on the device with the call signature the reference's PSD blocks use for scipy.signal.welch
    (main_multiuser/main_multiuser_mrt_precoding_radiation_pattern.py:371-381): complex input, two-sided, periodic Hann window,
    50 % overlap, constant detrend, density scaling.  `from utilities import welch` in place of `from scipy.signal import
    welch` moves the PSD blocks of the beampattern scripts to the GPU.  Returns (freqs, psd) like scipy."""
    if return_onesided:
        raise ValueError("the device PSD is two-sided (the reference always passes return_onesided=False for its complex signals)")
    x = np.asarray(x).ravel()
    nperseg = int(nperseg) if nperseg is not None else min(256, x.size)
    nfft = int(nfft) if nfft is not None else nperseg
    return get_ops().welch_psd(x, float(fs), nfft, nperseg)


def save_to_csv(data_lst: list, filename: str) -> None:
    with open("figs/csv_results/%s.csv" % filename, 'w', newline='') as csv_file:
        csv.writer(csv_file).writerows(data_lst)


def read_from_csv(filename: str) -> list:
    with open("../figs/csv_results/%s.csv" % filename, 'r', newline='') as csv_file:
        return list(csv.reader(csv_file, quoting=csv.QUOTE_NONNUMERIC))


def print_progress_bar(iteration: int, total: int, prefix: str = '', suffix: str = '', decimals: int = 1,
                       bar_length: int = 100) -> None:
    frac = iteration / float(total)
    filled = int(round(bar_length * frac))
    sys.stdout.write('\r%s |%s| %s%% %s' % (prefix, '|' * filled + '-' * (bar_length - filled),
                                            ("{0:." + str(decimals) + "f}").format(100 * frac), suffix))
    if iteration == total:
        sys.stdout.write('\n')
    sys.stdout.flush()


# ---- plotting (utilities.py:195-308): out of the hot path, matplotlib imported on use ------------
def plot_spatial_config(ant_array, rx_transceiver=None, rx_points_lst: list = None, plot_3d: bool = True) -> None:
    import matplotlib.pyplot as plt
    fig = plt.figure()
    ax = fig.add_subplot(projection='3d') if plot_3d else fig.add_subplot()
    xs = [t.cord_x for t in ant_array.array_elements]
    ys = [t.cord_y for t in ant_array.array_elements]
    zs = [t.cord_z for t in ant_array.array_elements]
    if plot_3d:
        ax.scatter(xs, ys, zs, marker='^', label="TX")
    else:
        ax.scatter(xs, ys, marker='^', label="TX")
    pts = []
    if rx_transceiver is not None:
        pts = [(rx_transceiver.cord_x, rx_transceiver.cord_y, rx_transceiver.cord_z)]
    elif rx_points_lst is not None:
        pts = list(rx_points_lst)
    if pts:
        p = np.asarray(pts, dtype=float)
        if plot_3d:
            ax.scatter(p[:, 0], p[:, 1], p[:, 2], marker='o', label="RX")
        else:
            ax.scatter(p[:, 0], p[:, 1], marker='o', label="RX")
    ax.set_xlabel("X plane [m]")
    ax.set_ylabel("Y plane [m]")
    ax.legend()
    ax.grid()
    plt.show()


def plot_array_config(ant_array, plot_3d: bool = False) -> None:
    plot_spatial_config(ant_array, plot_3d=plot_3d)
