"""CPU, world_size 2 over gloo: the N>1 plumbing of bench.py -- segment assignment covers every piece exactly once and
the whole-job throughput is (sum of units) / (slowest rank's time)."""
import os
import socket

import pytest
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, q):
    import torch.distributed as dist
    from govideocompressor_b200 import farm
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    segs = farm.segments_for_rank(8, rank, world)
    seconds = 2.0 if rank == 0 else 3.0                 # rank 1 is the straggler
    fps = farm.job_throughput(38 * len(segs), seconds)
    dist.barrier()
    q.put((rank, segs, fps, farm.max_over_ranks(seconds)))
    dist.destroy_process_group()


def test_two_ranks_share_eight_segments():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    (r0, s0, f0, m0), (r1, s1, f1, m1) = res
    assert sorted(s0 + s1) == list(range(8)) and not set(s0) & set(s1)
    assert s0[0] == 7 and s1[0] == 6                    # highest index first, like the dispatcher
    assert m0 == m1 == 3.0
    assert f0 == f1 == pytest.approx(38 * 8 / 3.0)


def test_single_process_needs_no_process_group():
    from govideocompressor_b200 import farm
    assert farm.segments_for_rank(5, 0, 1) == [4, 3, 2, 1, 0]
    assert farm.job_throughput(100, 4.0) == 25.0
