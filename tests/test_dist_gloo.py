"""CPU, world_size 2 over gloo: the N>1 host logic of bench.py (sharding, max-over-ranks, whole-job rate).
The scoring path itself has no collective, so this is all there is to test without GPUs."""
import os
import socket
import sys

import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    from flink_jpmml_b200 import dist as D
    dist = D.init("gloo")
    assert D.env_rank() == (rank, world, rank)
    lo, hi = D.shard(1001, world, rank)
    ms = [10.0 + 5.0 * rank, 3.0 - rank]              # rank 1 is slower on the first timer, faster on the second
    mx = D.max_over_ranks(dist, ms)
    total = D.sum_over_ranks(dist, [hi - lo])[0]
    dist.barrier()
    q.put((rank, lo, hi, mx, total, D.whole_job_rate(1000, world, 4, mx[0])))
    dist.destroy_process_group()


def test_two_rank_gloo_plumbing():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    [p.start() for p in procs]
    out = sorted(q.get(timeout=120) for _ in range(world))
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    (r0, lo0, hi0, mx0, tot0, rate0), (r1, lo1, hi1, mx1, tot1, rate1) = out
    assert (lo0, hi0, lo1, hi1) == (0, 501, 501, 1001)           # contiguous, sizes differ by at most one
    assert mx0 == mx1 == [15.0, 3.0]                              # the slowest rank defines the time
    assert tot0 == tot1 == 1001.0
    assert rate0 == rate1 == pytest.approx(1000 * 2 * 4 / 0.015)  # whole-job units / max time


def test_shard_covers_everything():
    from flink_jpmml_b200 import dist as D
    for n, w in [(0, 3), (7, 8), (1 << 22, 8), (10, 3)]:
        parts = [D.shard(n, w, r) for r in range(w)]
        assert parts[0][0] == 0 and parts[-1][1] == n
        assert all(parts[i][1] == parts[i + 1][0] for i in range(w - 1))
        assert max(b - a for a, b in parts) - min(b - a for a, b in parts) <= 1
