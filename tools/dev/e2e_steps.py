import os, sys, time
import numpy as np
sys.path[:0] = ['/root/repo']
import bench
import humanoid_mujoco_b200 as hmj
n = 4096
env = hmj.BatchedHumanoidEnv(n, randomization_factor=0.0, seed=1)
b = env.batch
z18 = np.zeros((n, 18))
def probe(params, qpos):
    b.set_params(params); b.set_state(qpos, z18, np.zeros(n), z18); b.step(np.zeros((n, 12)), 1); return b.get("ncon") > 0
params, qpos, attempts = bench.make_workload(b.default_params(), b.model["qpos0"][0], n, 1, probe)
b.set("status", np.zeros(n, dtype=np.int32))
env.reset_to(params, qpos)
b.step(np.zeros((n, 12)), bench.SETTLE_STEPS)
ring = bench.control_ring(0, 64, n)
for s in range(150): b.step(ring[s % 64], 1)
acts = hmj.pinned_empty((64, n, 12))
acts[...] = np.random.default_rng(5).uniform(-1, 1, acts.shape)
ts = []
for s in range(45):
    t0 = time.perf_counter(); env.step(acts[s % 64]); ts.append((time.perf_counter() - t0) * 1e6)
print("first 45 e2e steps (us):", " ".join(f"{t:.0f}" for t in ts))
