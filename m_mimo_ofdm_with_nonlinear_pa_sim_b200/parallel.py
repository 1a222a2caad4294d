"""Frame sharding across GPUs and the one exchange of the path: a sum of error / bit counters.

The reference fans `Link.simulate` out over OS processes that add into two shared `multiprocessing.Array`s
(main_mp_miso_mcnc_ber_vs_ebn0.py:121-134, mp_model.py:172-173,221-222).  Here one process drives one GPU;
Monte-Carlo frames are split by GLOBAL frame index (the Philox counter), so the set of simulated frames - and
with it every counter - does not depend on the number of GPUs, and the shared-array accumulation becomes a
single all-reduce of `n_slots + 1` integers per step (NCCL on GPUs, gloo in the CPU tests).  No tensor data
crosses GPUs."""
from __future__ import annotations

import numpy as np


def shard_range(n_total: int, rank: int, world: int):
    """Contiguous split of `n_total` frames: (first frame offset, frame count) of `rank`.  The first
    `n_total % world` ranks get one frame more."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world: %d/%d" % (rank, world))
    base, rem = divmod(int(n_total), world)
    start = rank * base + min(rank, rem)
    return start, base + (1 if rank < rem else 0)


class CounterExchange:
    """Sum of int64 counter vectors over all ranks.  `device` is where torch.distributed expects the tensor
    ("cuda" for NCCL, "cpu" for gloo); world size 1 needs no process group."""

    def __init__(self, world: int = 1, device: str = "cpu"):
        self.world = int(world)
        self.device = device
        self._buf = None

    def allreduce(self, values) -> np.ndarray:
        v = np.asarray(values, dtype=np.int64)
        if self.world == 1:
            return v.copy()
        import torch
        import torch.distributed as dist
        if self._buf is None or self._buf.numel() != v.size:
            self._buf = torch.zeros(v.size, dtype=torch.int64, device=self.device)
        self._buf.copy_(torch.from_numpy(v))
        dist.all_reduce(self._buf)
        return self._buf.cpu().numpy().copy()


def monte_carlo_point(run_batch, n_slots: int, bits_per_frame: int, n_err_min: int, bits_sent_max: int,
                      frames_per_step: int, rank: int = 0, world: int = 1, exchange: CounterExchange = None,
                      max_steps: int = 1 << 30):
    """One sweep point with the reference's stopping rule (mp_model.py:181-187) evaluated on the REDUCED counters
    between steps: every result slot counts frames until it holds `n_err_min` errors or `bits_sent_max` bits.

    run_batch(first_global_frame, n_frames, n_slots_active_mask) -> per-slot bit errors (length n_slots) of that
    block of global frames on this rank's GPU.  A step covers `frames_per_step` GLOBAL frames (split over the ranks by
    shard_range), clipped to what the bit budget of the least advanced active slot still allows - the stop rule is
    applied between steps, so the step size is global on purpose: the set of simulated frames, and with it every counter,
    is the same for any number of ranks.
    Returns (errors[n_slots], bits[n_slots], frames_simulated_globally)."""
    exchange = exchange or CounterExchange(world)
    err = np.zeros(n_slots, dtype=np.int64)
    bits = np.zeros(n_slots, dtype=np.int64)
    next_frame = 0
    for _ in range(max_steps):
        active = (err < n_err_min) & (bits < bits_sent_max)
        if not active.any():
            break
        left = int(np.ceil((bits_sent_max - int(bits[active].min())) / bits_per_frame))
        n_glob = max(1, min(frames_per_step, left))
        start, count = shard_range(n_glob, rank, world)
        e = np.zeros(n_slots, dtype=np.int64)
        if count > 0:
            e[:] = np.asarray(run_batch(next_frame + start, count, active), dtype=np.int64)
        tot = exchange.allreduce(np.concatenate((e, [count])))
        if int(tot[-1]) != n_glob:
            raise RuntimeError("ranks disagree on the step size: %d vs %d" % (int(tot[-1]), n_glob))
        err[active] += tot[:-1][active]
        bits[active] += n_glob * bits_per_frame
        next_frame += n_glob
    return err, bits, next_frame


def link_batch_runner(engine, n_iter: int, snr_db: float, seed: int, csi_eps: float = 0.0, incl_clean_run: bool = False):
    """run_batch for `monte_carlo_point` over a configured LinkEngine: slots = [clean] + iterations 0..n_iter."""
    def run(first_frame, n_frames, active):
        out = engine.simulate(n_frames, n_iter, snr_db, seed=seed, frame_id0=first_frame, csi_eps=csi_eps,
                              incl_clean_run=incl_clean_run)
        e = out["err"].astype(np.int64)
        return np.concatenate((out["clean_err"].astype(np.int64), e)) if incl_clean_run else e
    return run
