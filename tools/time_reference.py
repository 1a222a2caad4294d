#!/usr/bin/env python
"""Time the UNMODIFIED reference on the bench workload (build container only: needs /root/reference).

Two lines, as SURVEY.md 8(d) asks:
  1. `mp_model.Link.simulate` in ONE process (mp_model.py:89-222), fixed frame budget, after one warm-up frame;
  2. the launcher pattern of main_mp_miso_mcnc_ber_vs_ebn0.py:121-134: `mp.Process(target=link.simulate)` x
     `mp.cpu_count()` workers sharing two `mp.Array` counters, fixed frame budget.

Workload = bench.py's: K=64 LOS ULA, MRT, soft limiter IBO 0 dB, 64-QAM, n_fft 4096 / 2048 sub-carriers, Eb/N0 15 dB,
MCNC iterations [0..5], receiver position re-drawn per frame (plus the CNC receiver as a second single-process row).

    python tools/time_reference.py [--frames 8] [--mp-frames 48] > profiles/r02_reference_timing.json

The GPU box has no reference checkout; bench.py attaches this file's content as `cpu_baseline.reference_measured`.
"""
import argparse
import copy
import ctypes
import json
import multiprocessing as mp
import os
import platform
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.refharness import import_reference  # noqa: E402

R = import_reference()


def build_link(n_ant, n_frames, is_mcnc):
    modulation, distortion, transceiver, antenna_array, channel, noise, mp_model, utilities = (
        R[k] for k in ("modulation", "distortion", "transceiver", "antenna_array", "channel", "noise", "mp_model",
                       "utilities"))
    my_mod = modulation.OfdmQamModem(constel_size=64, n_fft=4096, n_sub_carr=2048, cp_len=128)
    my_dist = distortion.SoftLimiter(0, my_mod.avg_sample_power)
    my_tx = transceiver.Transceiver(modem=copy.deepcopy(my_mod), impairment=copy.deepcopy(my_dist),
                                    center_freq=int(3.5e9), carrier_spacing=int(15e3))
    my_rx = transceiver.Transceiver(modem=copy.deepcopy(my_mod), impairment=copy.deepcopy(my_dist), cord_x=212,
                                    cord_y=212, cord_z=1.5, center_freq=int(3.5e9), carrier_spacing=int(15e3))
    my_array = antenna_array.LinearArray(n_elements=n_ant, base_transceiver=my_tx, center_freq=int(3.5e9),
                                         wav_len_spacing=0.5, cord_x=0, cord_y=0, cord_z=15)
    ch = channel.MisoLosFd()
    ch.calc_channel_mat(tx_transceivers=my_array.array_elements, rx_transceiver=my_rx, skip_attenuation=False)
    nz = noise.Awgn(snr_db=10, seed=1234)
    link = mp_model.Link(mod_obj=my_mod, array_obj=my_array, std_rx_obj=my_rx, chan_obj=ch, noise_obj=nz,
                         rx_loc_var=10.0, n_err_min=int(1e12), bits_sent_max=my_mod.n_bits_per_ofdm_sym * n_frames,
                         is_mcnc=is_mcnc)
    link.update_distortion(ibo_val_db=0.0)
    link.set_snr(snr_db_val=float(utilities.ebn0_to_snr(15.0, 2048, 2048, 64)))
    return link


def run_single(n_ant, n_frames, is_mcnc, iters):
    def once(n):
        link = build_link(n_ant, n, is_mcnc)
        e = mp.Array(ctypes.c_double, len(iters), lock=True)
        b = mp.Array(ctypes.c_double, len(iters), lock=True)
        t0 = time.perf_counter()
        link.simulate(False, True, np.asarray(iters), [1, 2, 3], e, b)
        return time.perf_counter() - t0, list(e), list(b)
    once(1)  # numba warm-up
    dt, e, b = once(n_frames)
    return {"frames": n_frames, "seconds": dt, "frames_per_s": n_frames / dt,
            "ber": [x / y for x, y in zip(e, b)]}


def run_mp(n_ant, total_frames, iters, n_proc):
    """main_mp_miso_mcnc_ber_vs_ebn0.py:121-134: every worker runs link.simulate against the SHARED counters and stops
    when the shared bit count reaches bits_sent_max."""
    link = build_link(n_ant, total_frames, True)
    e = mp.Array(ctypes.c_double, len(iters), lock=True)
    b = mp.Array(ctypes.c_double, len(iters), lock=True)
    seeds = np.random.default_rng(2137).integers(0, high=sys.maxsize, size=(n_proc, 3))
    procs = [mp.Process(target=link.simulate, args=(False, True, np.asarray(iters), seeds[i], e, b))
             for i in range(n_proc)]
    t0 = time.perf_counter()
    for p in procs:
        p.start()
    for p in procs:
        p.join()
    dt = time.perf_counter() - t0
    frames = b[0] / 12288.0
    return {"processes": n_proc, "frames": frames, "seconds": dt, "frames_per_s": frames / dt,
            "note": "wall clock includes process start-up and the per-process numba compilation, as in the script",
            "ber": [x / y for x, y in zip(list(e), list(b))]}


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return platform.processor()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=8)
    ap.add_argument("--mp-frames", type=int, default=48)
    a = ap.parse_args()
    iters = [0, 1, 2, 3, 4, 5]
    import torch
    out = {
        "where": "build container (the GPU box has no reference checkout)",
        "cpu": cpu_model(), "vcpus": mp.cpu_count(),
        "versions": {"python": platform.python_version(), "numpy": np.__version__, "torch": torch.__version__},
        "workload": "K=64 LOS ULA, MRT, soft limiter IBO 0 dB, 64-QAM, n_fft 4096 / 2048 sub-carriers, Eb/N0 15 dB, "
                    "iterations [0..5], receiver position re-drawn per frame",
        "single_process_mcnc": run_single(64, a.frames, True, iters),
        "single_process_cnc": run_single(64, 4 * a.frames, False, iters),
        "multiprocess_mcnc": run_mp(64, a.mp_frames, iters, mp.cpu_count()),
    }
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
