import os, sys
import numpy as np
sys.path[:0] = ['/root/repo']
import bench
import humanoid_mujoco_b200 as hmj
q, v, t = bench.plan_start_state(0)
n = int(sys.argv[1])
b = hmj.HumanoidBatch(n)
p = b.default_params(); p["timestep"] = 0.02; b.set_params(p)
rng = np.random.default_rng(1)
b.set_state(np.tile(q, (n, 1)), np.tile(v, (n, 1)), np.zeros(n), np.zeros((n, 18)))
ctrl = rng.normal(0, 0.05, (n, 12))
b.step(ctrl, 12)
print("ncon", b.get("ncon"), "niter", b.get("niter"))
