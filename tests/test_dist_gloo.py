"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: block sharding of envs / samples and the
planner's one exchange step (arg-min of the returns + broadcast of the winning spline).  The device work itself is
covered by the -m gpu tests; here every rank fabricates its local results."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import humanoid_mujoco_b200  # noqa: F401
from humanoid_mujoco_b200.dist import exchange_winner, shard_range


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, case, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n_local, P = 8, 3
        rng = np.random.default_rng(100 + case)       # same stream on every rank: the global population
        returns = rng.uniform(1.0, 2.0, world * n_local)
        values = rng.standard_normal((world * n_local, P, 12))
        if case == 1:                                  # tie across ranks: lowest global index must win
            returns[3] = returns[n_local + 5] = 0.5
        if case == 2:                                  # winner on the last rank
            returns[world * n_local - 1] = 0.25
        if case == 3:                                  # NaN on one rank never wins
            returns[:n_local] = np.nan
        lo, hi = shard_range(world * n_local, rank, world)
        mine = returns[lo:hi]
        li = int(np.nanargmin(mine)) if not np.all(np.isnan(mine)) else 0
        cost, gidx, nominal, win = exchange_winner(float(mine[li]), li, float(returns[lo]), values[lo + li], lo)
        q.put((rank, cost, gidx, nominal, win.tolist(), returns.tolist(), values.tolist()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("case", [0, 1, 2, 3])
def test_exchange_winner_world2(case):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    returns, values = np.array(out[0][5]), np.array(out[0][6])
    want = int(np.nanargmin(returns))                  # first minimum == lowest global index on ties
    for rank, cost, gidx, nominal, win, _, _ in out:
        assert gidx == want, (rank, gidx, want)
        assert cost == returns[want]
        assert np.array_equal(np.array(win), values[want])
        assert nominal == returns[0] or (np.isnan(nominal) and np.isnan(returns[0]))


def test_shard_range_covers_everything_once():
    for n in (0, 1, 7, 8, 4096, 16384, 8192 + 3):
        for w in (1, 2, 4, 8):
            got = [shard_range(n, r, w) for r in range(w)]
            assert got[0][0] == 0 and got[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(got, got[1:]))
            sizes = [b - a for a, b in got]
            assert max(sizes) - min(sizes) <= 1


def test_single_process_exchange_is_identity():
    v = np.arange(36.0).reshape(3, 12)
    cost, gidx, nominal, win = exchange_winner(1.5, 4, 2.5, v, 16)
    assert (cost, gidx, nominal) == (1.5, 20, 2.5) and np.array_equal(win, v)


def _worker_rows(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from humanoid_mujoco_b200.dist import exchange_records, gather_rows, select_winner

        rows = np.arange(6.0).reshape(2, 3) + 100 * rank
        got = gather_rows(rows)
        rec = np.concatenate([[2.0 - rank, 1024.0 * rank + 7, 5.0 + rank], np.full(36, float(rank))])
        win = exchange_records(rec)        # gloo path: host all-gather + select_winner
        q.put((rank, got.tolist(), win.tolist(), select_winner(np.stack([rec, rec])).tolist()))
    finally:
        dist.destroy_process_group()


def test_elite_rows_and_record_exchange_world2():
    """The cross-entropy planner's elite exchange (gather_rows) and the record form of the winner exchange on gloo."""
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_rows, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    want = np.concatenate([np.arange(6.0).reshape(2, 3), np.arange(6.0).reshape(2, 3) + 100])
    for rank, got, win, _ in out:
        assert np.array_equal(np.array(got), want)
        win = np.array(win)
        assert win[0] == 1.0 and win[1] == 1031.0 and win[2] == 5.0      # rank 1 wins; the nominal return is rank 0's
        assert np.all(win[3:] == 1.0)
