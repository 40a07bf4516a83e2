"""How the path spreads over several GPUs: by document.

Every cluster manager is one document (`<root>/<name>/main.tf.json`, backend/local/backend.go:14-19)
and no operation reads across documents, so N workers each take a share of the managers and run
the whole path on their own GPU.  There is no exchange step and therefore no collective in the
data path; the only communication is bookkeeping (who takes what, and the slowest rank's time).
"""
from __future__ import annotations


def assign(names, world: int):
    """Managers per rank: sorted names dealt round-robin (deterministic on every rank)."""
    names = sorted(names)
    return [names[r::world] for r in range(world)]


def job_time_ms(local_ms: float, dist=None, device=None) -> float:
    """The job's time is the slowest rank's (max over ranks)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local_ms
    import torch
    t = torch.tensor([local_ms], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def job_value(ops_per_rank: int, world: int, ms: float) -> float:
    """Whole-job throughput: every rank's operations over the slowest rank's time."""
    return world * ops_per_rank / (ms * 1e-3)
