"""Host logic of the multi-GPU path on CPU: frame sharding by global frame index and the counter all-reduce,
world_size 2 over the gloo backend (the GPU build uses the same code with NCCL).  The per-frame "errors" come from a
deterministic stand-in keyed on the global frame index - the same property the Philox-keyed kernels have - so the
reduced counters must not depend on the number of ranks."""
import multiprocessing as mp
import os
import socket

import numpy as np
import pytest

from m_mimo_ofdm_with_nonlinear_pa_sim_b200.parallel import CounterExchange, monte_carlo_point, shard_range

N_SLOTS, BITS = 4, 12288


def fake_errors(first, n, active):
    f = np.arange(first, first + n, dtype=np.int64)[:, None]
    j = np.arange(N_SLOTS, dtype=np.int64)[None, :]
    e = ((f * 2654435761 + j * 40503) % 997) // (1 + 3 * j)  # fewer errors in later slots, like CNC iterations
    return e.sum(axis=0)


def test_shard_range_covers_everything():
    for total in (0, 1, 7, 64, 4097):
        for world in (1, 2, 3, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == total
            for (s0, c0), (s1, _) in zip(spans, spans[1:]):
                assert s0 + c0 == s1
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_single_rank_stop_rule():
    err, bits, frames = monte_carlo_point(fake_errors, N_SLOTS, BITS, n_err_min=10 ** 12, bits_sent_max=BITS * 100,
                                          frames_per_step=32)
    assert frames == 100 and np.all(bits == BITS * 100)
    assert np.array_equal(err, fake_errors(0, 100, None))
    # error threshold: slot 0 collects errors fastest and stops first; later slots keep going
    err2, bits2, _ = monte_carlo_point(fake_errors, N_SLOTS, BITS, n_err_min=20000, bits_sent_max=BITS * 10000,
                                       frames_per_step=8)
    assert np.all(err2 >= 20000) and bits2[0] < bits2[-1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ex = CounterExchange(world, "cpu")
        a = monte_carlo_point(fake_errors, N_SLOTS, BITS, 10 ** 12, BITS * 101, 16, rank, world, ex)
        b = monte_carlo_point(fake_errors, N_SLOTS, BITS, 20000, BITS * 10000, 4, rank, world, ex)
        q.put((rank, [x.tolist() if hasattr(x, "tolist") else x for x in a], [x.tolist() if hasattr(x, "tolist") else x for x in b]))
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo_match_one_rank():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ref_a = monte_carlo_point(fake_errors, N_SLOTS, BITS, 10 ** 12, BITS * 101, 16)   # the SAME arguments on one rank
    ref_b = monte_carlo_point(fake_errors, N_SLOTS, BITS, 20000, BITS * 10000, 4)
    for rank, a, b in res:
        assert a[0] == ref_a[0].tolist() and a[1] == ref_a[1].tolist() and a[2] == ref_a[2]
        assert b[0] == ref_b[0].tolist() and b[1] == ref_b[1].tolist() and b[2] == ref_b[2]
