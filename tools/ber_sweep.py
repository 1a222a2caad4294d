#!/usr/bin/env python
"""BER-vs-Eb/N0 (or vs IBO) sweep on the fused B200 link kernels - the job of
main_mp_clipping_noise_cancellation/main_mp_miso_{cnc,mcnc}_ber_vs_{ebn0,ibo}.py: same stopping rule
(n_err_min / bits_sent_max per result row), same CSV row layout (row 0 swept axis, row 1 no distortion, row 2
standard receiver, rows 3.. iterations 1..I; utilities.save_to_csv).  One process per GPU under torchrun; frames are
sharded by global frame index and the counters all-reduced (parallel.monte_carlo_point).

  python tools/ber_sweep.py --ebn0 5:20:0.5 --n-iter 8 --out ber.csv
  torchrun --nproc-per-node 8 tools/ber_sweep.py --sweep ibo --ibo -9:9:0.5 --ebn0 15 --out ber_vs_ibo.csv
  python tools/ber_sweep.py --sweep fixed-ber --ebn0 10:22:0.5 --ibo 0:7.5:0.5 --out fixed_ber.csv      # BER grid, reference layout
  python tools/ber_sweep.py --sweep nant-chan --ebn0 15 --out ber_vs_nant.csv
"""
import argparse
import csv
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def write_csv(path, data):
    """utilities.save_to_csv (utilities.py:342-352): one csv.writer row per data vector"""
    with open(path, "w", newline="") as fh:
        csv.writer(fh).writerows(data)


def read_csv(path):
    """utilities.read_from_csv (utilities.py:355-365), as the final_plots/*.py readers call it"""
    with open(path, "r", newline="") as fh:
        return list(csv.reader(fh, quoting=csv.QUOTE_NONNUMERIC))


def csv_rows(mode, ax, rows, ber_grid=None):
    """The reference's row layouts.  ebn0 / ibo / nant-chan: row 0 the swept axis, then one row per result slot
    (main_mp_miso_mcnc_ber_vs_ebn0.py:175-179; nant-chan: per channel, ..._ber_vs_nant_vs_chan.py:205-210).  fixed-ber: row 0
    the IBO axis, then the BER of every (IBO, Eb/N0) pair, IBO-major, one value per iteration
    (main_mp_miso_mcnc_constant_ber_req_ebn0_vs_ibo.py:210-215; final_plots/fixed_ber_ebno_vs_ibo.py:46-57 re-derives the
    required Eb/N0 from that grid)."""
    if mode == "fixed-ber":
        return [list(ax)] + [list(ber_grid[i, j]) for i in range(ber_grid.shape[0]) for j in range(ber_grid.shape[1])]
    return [list(ax)] + [list(r) for r in rows]


def axis(spec):
    p = [float(v) for v in spec.split(":")]
    if len(p) == 1:
        return np.array(p)
    return np.arange(p[0], p[1] + p[2] / 2, p[2])


def run_sweep(eng, a, rank=0, world=1, exchange=None):
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200.parallel import link_batch_runner, monte_carlo_point
    ebn0s, ibos = axis(a.ebn0), axis(a.ibo)
    pts = [(e, ibos[0]) for e in ebn0s] if a.sweep == "ebn0" else [(ebn0s[0], i) for i in ibos]
    bits_frame = int(math.log2(a.constel)) * 2048
    rows = np.zeros((a.n_iter + 2, len(pts)))
    frames = 0
    for c, (ebn0, ibo) in enumerate(pts):
        eng.set_ibo(float(ibo))
        snr = 10 * math.log10(10 ** (ebn0 / 10) * math.log2(a.constel))  # ebn0_to_snr with n_fft := n_sub_carr
        runner = link_batch_runner(eng, a.n_iter, snr, seed=a.seed + 7919 * c, csi_eps=a.csi_eps, incl_clean_run=True)
        err, bits, n = monte_carlo_point(runner, a.n_iter + 2, bits_frame, int(a.err_min), int(a.bits_max), a.frames_per_step,
                                         rank, world, exchange)
        rows[:, c] = err / bits
        frames += n
    ax = ebn0s if a.sweep == "ebn0" else ibos
    return ax, rows, frames


def run_fixed_ber(eng, a, rank=0, world=1, exchange=None):
    """Required Eb/N0 for a target BER versus IBO (main_mp_miso_mcnc_constant_ber_req_ebn0_vs_ibo.py:99-180): BER on the
    whole IBO x Eb/N0 grid, every IBO row as ONE multi-point batch per round (LinkEngine.sweep), the reference's stopping
    rule per (point, iteration), then interp1d(BER -> Eb/N0) per IBO and iteration (inf where the target is not crossed).
    Returns (ibo axis, req_ebn0 [n_iter+1, n_ibo], ber [n_ibo, n_ebn0, n_iter+1], frames)."""
    from scipy import interpolate

    from m_mimo_ofdm_with_nonlinear_pa_sim_b200.parallel import shard_range
    ebn0s, ibos = axis(a.ebn0), axis(a.ibo)
    bits_frame = int(math.log2(a.constel)) * 2048
    snr = 10 * np.log10(10 ** (ebn0s / 10) * math.log2(a.constel))
    n_it = a.n_iter + 1
    ber = np.full((len(ibos), len(ebn0s), n_it), np.nan)
    frames = 0
    fpp = a.frames_per_step   # frames per point and round
    for i, ibo in enumerate(ibos):
        err = np.zeros((len(ebn0s), n_it)); bits = np.zeros((len(ebn0s), n_it))
        rnd = 0
        while True:
            active = (err < a.err_min) & (bits < a.bits_max)          # mp_model.py:181-187 per result slot
            if not active.any():
                break
            pts = np.nonzero(active.any(axis=1))[0]
            # the active points of this round form ONE batch with frame ids base + j * fpp + frame (j = position in `pts`);
            # every rank runs a contiguous block of points of that batch, so the counters do not depend on the rank count
            j0, cnt = shard_range(len(pts), rank, world)
            base = ((i * 4096 + rnd) * len(ebn0s)) * fpp
            e = np.zeros((len(pts), n_it))
            if cnt > 0:
                e[j0:j0 + cnt] = eng.sweep(snr[pts[j0:j0 + cnt]], np.full(cnt, ibo), fpp, a.n_iter, seed=a.seed,
                                           frame_id0=base + j0 * fpp, csi_eps=a.csi_eps)
            if exchange is not None:
                e = exchange.allreduce(e.ravel()).reshape(len(pts), n_it).astype(np.float64)
            for j, p in enumerate(pts):
                err[p] += np.where(active[p], e[j], 0)
                bits[p] += np.where(active[p], fpp * bits_frame, 0)
            frames += fpp * len(pts)
            rnd += 1
        ber[i] = err / np.maximum(bits, 1)
    req = np.full((n_it, len(ibos)), np.inf)
    for it in range(n_it):
        for i in range(len(ibos)):
            try:
                req[it, i] = float(interpolate.interp1d(ber[i, :, it], ebn0s)(a.target_ber))
            except ValueError:
                pass  # target not crossed on this Eb/N0 range: the reference stores inf (script :171-175)
    return ibos, req, ber, frames


def run_nant_chan(eng, a, rank=0, world=1, exchange=None):
    """BER versus number of antennas versus channel type at one Eb/N0
    (main_mp_clipping_noise_cancellation/main_mp_miso_mcnc_ber_vs_nant_vs_chan.py:80-150).  CSV layout of that script:
    row 0 the antenna counts, then per channel (los, two_path, rayleigh) one row per result slot (no distortion,
    standard receiver, iterations 1..I)."""
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200.parallel import link_batch_runner, monte_carlo_point
    n_ants = [int(v) for v in a.n_ants.split(",")]
    chans = a.channels.split(",")
    ebn0 = axis(a.ebn0)[0]
    bits_frame = int(math.log2(a.constel)) * 2048
    snr = 10 * math.log10(10 ** (ebn0 / 10) * math.log2(a.constel))
    rows = np.zeros((len(chans), a.n_iter + 2, len(n_ants)))
    frames = 0
    for ci, ch in enumerate(chans):
        for ni, n_ant in enumerate(n_ants):
            eng.configure(constel_size=a.constel, n_ant=n_ant, pa=a.pa, ibo_db=axis(a.ibo)[0], channel=ch, receiver=a.receiver)
            runner = link_batch_runner(eng, a.n_iter, snr, seed=a.seed + 7919 * (ci * 64 + ni), csi_eps=a.csi_eps, incl_clean_run=True)
            err, bits, n = monte_carlo_point(runner, a.n_iter + 2, bits_frame, int(a.err_min), int(a.bits_max), a.frames_per_step,
                                             rank, world, exchange)
            rows[ci, :, ni] = err / bits
            frames += n
    return np.array(n_ants, dtype=float), rows.reshape(-1, len(n_ants)), frames


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sweep", choices=["ebn0", "ibo", "fixed-ber", "nant-chan"], default="ebn0")
    ap.add_argument("--target-ber", type=float, default=1e-2)
    ap.add_argument("--n-ants", default="1,2,4,8,16,32,64,128", help="nant-chan sweep: antenna counts")
    ap.add_argument("--channels", default="los,two_path,rayleigh", help="nant-chan sweep: channel types")
    ap.add_argument("--ebn0", default="5:20:0.5")
    ap.add_argument("--ibo", default="0")
    ap.add_argument("--n-ant", type=int, default=64)
    ap.add_argument("--constel", type=int, default=64)
    ap.add_argument("--channel", default="los", choices=["los", "two_path", "rayleigh"])
    ap.add_argument("--receiver", default="mcnc", choices=["cnc", "mcnc"])
    ap.add_argument("--pa", default="softlim", choices=["softlim", "rapp"])
    ap.add_argument("--n-iter", type=int, default=8)
    ap.add_argument("--csi-eps", type=float, default=0.0)
    ap.add_argument("--bits-max", type=float, default=1e7)
    ap.add_argument("--err-min", type=float, default=1e6)
    ap.add_argument("--frames-per-step", type=int, default=1024)
    ap.add_argument("--seed", type=int, default=2137)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()

    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import LinkEngine
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200.parallel import CounterExchange
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    exchange = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        exchange = CounterExchange(world, "cuda")
    eng = LinkEngine(local)
    eng.configure(constel_size=a.constel, n_ant=a.n_ant, pa=a.pa, ibo_db=axis(a.ibo)[0], channel=a.channel, receiver=a.receiver)
    t0 = time.time()
    _ber = None
    if a.sweep == "nant-chan":
        ax, rows, frames = run_nant_chan(eng, a, rank, world, exchange)
    elif a.sweep == "fixed-ber":
        ax, rows, _ber, frames = run_fixed_ber(eng, a, rank, world, exchange)
    else:
        ax, rows, frames = run_sweep(eng, a, rank, world, exchange)
    dt = time.time() - t0
    if rank == 0:
        if a.out:
            write_csv(a.out, csv_rows(a.sweep, ax, rows, _ber if a.sweep == "fixed-ber" else None))
        print("--- Computation time: %f ---  (%d frames, %.0f frames/s, %d GPU)" % (dt, frames, frames / dt, world))
        for r in [list(ax)] + [list(r) for r in rows]:   # fixed-ber: the required Eb/N0 per iteration (rows) and IBO (columns)
            print(" ".join("%.3e" % v for v in r))
    eng.close()
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
