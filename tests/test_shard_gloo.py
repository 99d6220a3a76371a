"""N > 1 host logic on CPU: two gloo ranks shard independent sequences (id mod world), time
"their" work, and the whole-job rate uses the slowest rank; per-sequence results are reassembled in id order.
No data-path collective exists on this path (sequences never exchange pixels)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "direct-visual-odometry_b200"))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, n_seq, q):
    import torch.distributed as dist
    from dvo_b200 import shard
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ids = shard.shard_ids(n_seq, world, rank)
        # stand-in for per-sequence poses: a function of the sequence id only (independent sequences)
        vals = np.stack([np.full((4, 4), float(i), np.float32) for i in ids]) if len(ids) else np.zeros((0, 4, 4), np.float32)
        ms = 10.0 * (rank + 1)                     # rank 1 is the slow one
        slow = shard.max_over_ranks(ms, dist)
        rate = shard.whole_job_rate(len(ids) * 5, ms, dist)   # 5 frames per sequence
        allv = shard.gather_by_sequence(ids, vals, n_seq, dist)
        dist.barrier()
        q.put((rank, ids.tolist(), slow, rate, allv[:, 0, 0].tolist()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_seq", [256, 7])
def test_two_ranks_shard_and_aggregate(n_seq):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_seq, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    ids0, ids1 = out[0][1], out[1][1]
    assert sorted(ids0 + ids1) == list(range(n_seq))          # every sequence exactly once
    assert all(i % 2 == 0 for i in ids0) and all(i % 2 == 1 for i in ids1)
    for rank, ids, slow, rate, allv in out:
        assert slow == 20.0                                    # max over ranks
        assert abs(rate - n_seq * 5 / 20e-3) < 1e-6 * rate     # all ranks' units / slowest time
        assert allv == [float(i) for i in range(n_seq)]        # reassembled in id order on every rank


def test_single_process_paths():
    from dvo_b200 import shard
    assert shard.shard_ids(10, 1, 0).tolist() == list(range(10))
    assert shard.owner_of(13, 4) == 1
    assert shard.max_over_ranks(3.5) == 3.5
    assert shard.whole_job_rate(100, 50.0) == 2000.0
    v = shard.gather_by_sequence([2, 0, 1], np.array([[2.0], [0.0], [1.0]]), 3)
    assert v[:, 0].tolist() == [0.0, 1.0, 2.0]
    with pytest.raises(ValueError):
        shard.shard_ids(4, 2, 2)
