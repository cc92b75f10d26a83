"""Run under torchrun (>= 2 GPUs): the planner's winner exchange through the library's own NCCL communicator
(hmj_comm_* + hmj_plan_exchange, what a C++ host uses) against the torch.distributed path of SamplingPlanner."""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [REPO, os.path.join(REPO, "tools")]
import humanoid_mujoco_b200 as hmj  # noqa: E402
from humanoid_mujoco_b200._lib import check  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = hmj.load_library()
    lib.hmj_comm_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
    lib.hmj_comm_destroy.argtypes = [C.c_void_p]
    lib.hmj_comm_destroy.restype = None
    lib.hmj_plan_exchange.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    idbuf = (C.c_char * 128)()
    if rank == 0:
        check(lib.hmj_comm_unique_id(idbuf))
    t = torch.tensor(list(bytes(idbuf)), dtype=torch.uint8, device="cuda")
    dist.broadcast(t, 0)
    idbuf = (C.c_char * 128).from_buffer_copy(bytes(t.cpu().numpy().tolist()))
    comm = C.c_void_p()
    check(lib.hmj_comm_create(idbuf, rank, world, local, C.byref(comm)))

    n, H = 64, 18
    pl = hmj.SamplingPlanner(num_trajectory=n, device=local, seed=13)
    assert (pl.rank, pl.world) == (rank, world)
    b = pl.batch
    q = np.array(b.model["qpos0"][0]); v = np.zeros(18)
    pl.SetState(q, v, 0.0)
    ok = True
    for it in range(4):
        pl.UpdateNominalPolicy(H)
        times, nominal = pl.policy.plan.arrays()
        # torch.distributed path (all_gather_into_tensor on the batch stream + device selection)
        pl.iteration = it
        pl.OptimizePolicyCandidates(1, H)
        want = (pl.winner, pl.winner_return, pl.nominal_return, pl._winner_values.copy())
        # library path: same iteration redone, exchange through hmj_plan_exchange
        b.plan_iterate(pl.state, pl.time, times, nominal, H, pl.interpolation_, noise=None, noise_std=pl.noise_exploration[0],
                       seed=pl.seed, iteration=it, first_global=rank * n)
        rec = np.zeros(3 + 3 * 12)
        check(lib.hmj_plan_exchange(b._h, comm, 3, rec.ctypes.data))
        same = (int(rec[1]) == want[0] and rec[0] == want[1] and rec[2] == want[2] and np.array_equal(rec[3:].reshape(3, 12), want[3]))
        ok = ok and same
        # every rank must hold the same winner
        g = torch.tensor(rec, device="cuda")
        parts = [torch.zeros_like(g) for _ in range(world)]
        dist.all_gather(parts, g)
        ok = ok and all(torch.equal(p, parts[0]) for p in parts)
        pl.CopyCandidateToPolicy(0)
        if rank == 0:
            print(f"iteration {it}: winner {want[0]} (owner rank {want[0] // n}) return {want[1]:.6f} nominal {want[2]:.6f} same={same}", flush=True)
    lib.hmj_comm_destroy(comm)
    pl.batch.close()
    dist.destroy_process_group()
    if not ok:
        raise SystemExit(1)
    if rank == 0:
        print("EXCHANGE OK", flush=True)


if __name__ == "__main__":
    main()
