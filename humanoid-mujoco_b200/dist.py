"""Multi-GPU plumbing of the two shardable paths: one process per GPU, torch.distributed.

  * env stepping shards by contiguous blocks with no data-path collective (`shard_range`);
  * the planner has ONE exchange step per iteration: agree on the arg-min of the per-sample returns and hand
    every rank the winning spline (`exchange_winner`) — the parallel form of the `partial_sort` +
    `CopyCandidateToPolicy` pair in mjpc/planners/sampling/planner.cc:151-187, 525-534.

The collectives carry 3 + P*nu doubles; they run on whatever backend the process group was created with
(NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

import numpy as np


def world():
    """(rank, world_size) of the default process group, (0, 1) when torch.distributed is not initialised."""
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except Exception:
        pass
    return 0, 1


def shard_range(n_total: int, rank: int, world_size: int):
    """Contiguous block [lo, hi) of `n_total` units owned by `rank`; earlier ranks take the remainder."""
    base, rem = divmod(int(n_total), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def exchange_winner(cost: float, local_index: int, nominal_return: float, win_values: np.ndarray, first_global: int,
                    device=None):
    """All ranks call this with their local best.  Returns (cost, global_index, nominal_return_of_rank0, values) of
    the global winner: the lowest cost; ties go to the lowest global index (std::partial_sort keeps the first minimum
    of a stable scan, planner.cc:176-181).  NaN costs never win unless every rank holds NaN."""
    import torch
    import torch.distributed as dist

    rank, ws = world()
    if ws == 1:
        return float(cost), int(first_global + local_index), float(nominal_return), np.array(win_values, dtype=np.float64)
    if device is None:
        device = torch.device("cpu")
    key = float(cost) if cost == cost else float("inf")
    mine = torch.tensor([key, float(first_global + local_index), float(nominal_return)], dtype=torch.float64, device=device)
    parts = [torch.zeros_like(mine) for _ in range(ws)]
    dist.all_gather(parts, mine)
    g = torch.stack(parts).cpu().numpy()
    # lexicographic (cost, global index)
    owner = int(np.lexsort((g[:, 1], g[:, 0]))[0])
    spline = torch.tensor(np.ascontiguousarray(win_values, dtype=np.float64), device=device)
    dist.broadcast(spline, src=owner)
    return float(g[owner, 0]), int(g[owner, 1]), float(g[0, 2]), spline.cpu().numpy()
