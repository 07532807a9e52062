"""Multi-GPU plumbing: one process per GPU, whole segments per rank, no collective on the data path.

The reference scales by handing whole pieces to whichever worker connects next (server.go:176-190); pieces share
nothing.  On one 8xB200 box the same thing is one worker process per GPU.  The only cross-rank operations are for
measurement: a barrier and the MAX of the elapsed time (bench.py contract)."""


def segments_for_rank(n_segments, rank, world):
    """Static equivalent of the pull-based dispatcher for a benchmark: pieces are handed out highest index first
    (server.go:170-184), round-robin over the ranks."""
    order = list(range(n_segments - 1, -1, -1))
    return order[rank::world]


def max_over_ranks(value, device=None):
    """MAX-reduce a float over all ranks (torch.distributed must be initialised; any backend)."""
    import torch
    import torch.distributed as dist
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value, device=None):
    import torch
    import torch.distributed as dist
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def job_throughput(units_local, seconds_local, device=None):
    """Whole-job throughput: units processed by all ranks / the slowest rank's time."""
    return sum_over_ranks(units_local, device) / max_over_ranks(seconds_local, device)
