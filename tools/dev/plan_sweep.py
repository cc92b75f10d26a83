"""Developer tool: planner iterations/s vs sample count (GPU) — where the GPU overtakes the host threads."""
import os, sys, time
import numpy as np
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))]
import bench
import humanoid_mujoco_b200 as hmj

q, v, t = bench.plan_start_state(0)
for n, H, nit in ((10, 18, 100), (64, 18, 50), (256, 18, 50), (1024, 18, 30), (1024, 100, 10), (4096, 100, 5)):
    pl = hmj.SamplingPlanner(num_trajectory=n, device=0, seed=1)
    pl.SetState(q, v, t)
    for _ in range(3):
        pl.OptimizePolicy(H)
    t0 = time.perf_counter()
    for _ in range(nit):
        pl.OptimizePolicy(H)
    dt = (time.perf_counter() - t0) / nit
    print(f"n={n:5d} H={H:3d}: {1/dt:8.1f} it/s  {1e3*dt:7.3f} ms/it  {1e6*dt/(H-1):7.1f} us/step  {n*(H-1)/dt:.3e} sample-steps/s", flush=True)
    pl.batch.close()
