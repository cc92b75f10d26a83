"""Developer tool: per-stage CUDA-event times of a physics step at planner conditions (dt = 0.02, standing start) for small batches."""
import os, sys
import numpy as np
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))]
import bench
import humanoid_mujoco_b200 as hmj

q, v, t = bench.plan_start_state(0)
for n in (1, 10, 128, 1024):
    b = hmj.HumanoidBatch(n)
    p = b.default_params(); p["timestep"] = 0.02; b.set_params(p)
    rng = np.random.default_rng(1)
    b.set_state(np.tile(q, (n, 1)), np.tile(v, (n, 1)), np.zeros(n), np.zeros((n, 18)))
    ctrl = rng.normal(0, 0.05, (n, 12))
    b.step(ctrl, 5)
    b.profile(True)
    for _ in range(17):
        b.step(ctrl, 1)
    pr = b.profile_read()
    b.profile(False)
    k = pr["steps"]
    print(f"n={n:5d}: smooth {1e3*pr['smooth_ms']/k:6.1f} us  collide+factor {1e3*pr['collide_ms']/k:6.1f} us  solve {1e3*pr['solve_ms']/k:6.1f} us"
          f"  ncon {b.get('ncon').mean():.2f} nefc {b.get('nefc').mean():.1f} niter {b.get('niter').mean():.2f} max {b.get('niter').max()}", flush=True)
    b.close()
