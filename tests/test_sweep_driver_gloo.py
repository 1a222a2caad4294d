"""tools/ber_sweep.run_fixed_ber on CPU: the host logic of the fixed-BER driver (multi-point batches, stopping rule,
interp1d) with a deterministic stand-in for LinkEngine.sweep keyed on the global frame id - the property the Philox-keyed
kernels have - single rank and world_size 2 over gloo: the counters must not depend on the number of ranks."""
import argparse
import importlib.util
import multiprocessing as mp
import os
import socket

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_driver():
    spec = importlib.util.spec_from_file_location("ber_sweep", os.path.join(ROOT, "tools", "ber_sweep.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class FakeEngine:
    """errors of a frame = f(frame id, SNR, IBO, iteration): fewer at high SNR / IBO / iteration"""

    def sweep(self, snr_db, ibo_db, fpp, n_iter, *, seed=0, frame_id0=0, csi_eps=0.0):
        snr, ibo = np.asarray(snr_db, float), np.asarray(ibo_db, float)
        out = np.zeros((len(snr), n_iter + 1), dtype=np.uint64)
        for p in range(len(snr)):
            ids = frame_id0 + p * fpp + np.arange(fpp, dtype=np.int64)
            noise = (ids * 2654435761 + seed) % 97
            for it in range(n_iter + 1):
                mean = 12288 * 0.05 * 10 ** (-(snr[p] - 24 + 1.5 * it + 0.5 * ibo[p]) / 8.0)
                out[p, it] = int(np.sum(np.floor(mean * (0.8 + 0.4 * noise / 97.0))))
        return out


def args():
    return argparse.Namespace(ebn0="16:26:1", ibo="0:4:2", constel=64, n_iter=3, err_min=3e3, bits_max=12288 * 40,
                              frames_per_step=8, seed=3, csi_eps=0.0, target_ber=1e-2)


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200.parallel import CounterExchange
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        drv = load_driver()
        ibos, req, ber, frames = drv.run_fixed_ber(FakeEngine(), args(), rank, world, CounterExchange(world, "cpu"))
        q.put((rank, req.tolist(), ber.tolist(), frames))
    finally:
        dist.destroy_process_group()


def test_fixed_ber_driver_single_rank():
    drv = load_driver()
    ibos, req, ber, frames = drv.run_fixed_ber(FakeEngine(), args())
    assert ibos.tolist() == [0.0, 2.0, 4.0] and req.shape == (4, 3) and ber.shape == (3, 11, 4)
    assert np.all(np.diff(ber, axis=1) < 0)                      # BER falls with Eb/N0
    fin = np.isfinite(req)
    assert fin.any() and np.all(req[fin] >= 16.0) and np.all(req[fin] <= 26.0)
    assert np.all(np.diff(req[1:, 0]) < 0)                       # later iterations need less Eb/N0 for the target BER
    # the stopping rule: low Eb/N0 points stop on n_err_min early, high Eb/N0 points run to bits_sent_max
    assert frames < 3 * 11 * 40 and frames % 8 == 0


def test_fixed_ber_driver_two_ranks_match_one_rank():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    drv = load_driver()
    ibos, req, ber, frames = drv.run_fixed_ber(FakeEngine(), args())
    for rank, r_req, r_ber, r_frames in res:
        assert r_frames == frames
        assert np.array_equal(np.array(r_ber), ber)
        assert np.array_equal(np.array(r_req), req)
