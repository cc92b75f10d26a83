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


def select_winner(records: np.ndarray) -> np.ndarray:
    """Host statement of the device reduction `k_plan_select` (csrc/hmj_aux.cu): `records` [W][3 + P*12] holds one
    winner record per rank {cost, global index, return of the rank's sample 0, spline values}.  The global winner is
    the lexicographic minimum of (cost, global index); NaN costs never win unless every rank holds NaN; the nominal
    return is rank 0's (it owns global sample 0)."""
    records = np.asarray(records, dtype=np.float64)
    key = np.where(np.isnan(records[:, 0]), np.inf, records[:, 0])
    owner = int(np.lexsort((records[:, 1], key))[0])
    out = records[owner].copy()
    out[2] = records[0, 2]
    return out


def exchange_records(record: np.ndarray, batch=None, rec_local=None, rec_all=None) -> np.ndarray:
    """The planner's one exchange step.  NCCL: `rec_local` / `rec_all` are torch CUDA tensors ([R] written by
    hmj_plan_iterate, [W][R]); the all-gather is issued on the batch's own stream and the selection runs on the device
    right behind it (hmj_plan_select) — one synchronisation, no host copy between collective and selection.  Other
    backends (gloo in the CPU tests): the host record is all-gathered and `select_winner` picks."""
    import torch
    import torch.distributed as dist

    rank, ws = world()
    if ws == 1:
        return record
    if rec_all is not None and dist.get_backend() == "nccl":
        stream = torch.cuda.ExternalStream(batch.stream, device=rec_all.device)
        with torch.cuda.stream(stream):
            dist.all_gather_into_tensor(rec_all, rec_local)
        nnode = (rec_local.numel() - 3) // 12
        return batch.plan_select(nnode, rec_all.data_ptr(), ws)
    mine = torch.from_numpy(np.ascontiguousarray(record, dtype=np.float64))
    parts = [torch.zeros_like(mine) for _ in range(ws)]
    dist.all_gather(parts, mine)
    return select_winner(torch.stack(parts).numpy())


def gather_rows(rows: np.ndarray, device_index: int = 0) -> np.ndarray:
    """All ranks contribute a [k][w] float64 array of the same shape; everyone gets the [world * k][w] stack in rank order
    (the cross-entropy planner's elite exchange).  NCCL: tensors on the rank's GPU; gloo: CPU tensors."""
    import torch
    import torch.distributed as dist

    rank, ws = world()
    rows = np.ascontiguousarray(rows, dtype=np.float64)
    if ws == 1:
        return rows
    dev = torch.device("cuda", device_index) if dist.get_backend() == "nccl" else torch.device("cpu")
    mine = torch.from_numpy(rows).to(dev)
    parts = [torch.zeros_like(mine) for _ in range(ws)]
    dist.all_gather(parts, mine)
    return torch.cat(parts).cpu().numpy()


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
