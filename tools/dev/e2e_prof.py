"""Where the end-to-end control step spends its time (4096 envs): python staging, the fused C call, result handling."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import bench
import humanoid_mujoco_b200 as hmj

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
env = hmj.BatchedHumanoidEnv(n, randomization_factor=1.0, device=0, seed=1)
b = env.batch
params, qpos = bench.make_workload(b.default_params(), b.model["qpos0"][0], n, 1)
env.params[:] = params
b.set_params(params)
b.set_state(qpos, np.zeros((n, 18)), np.zeros(n), np.zeros((n, 18)))
b.step(np.zeros((n, 12)), bench.SETTLE_STEPS)
acts = hmj.pinned_empty((64, n, 12))
acts[...] = np.random.default_rng(5).uniform(-1, 1, acts.shape)
for s in range(20):
    env.step(acts[s])
K = 300
T0 = time.perf_counter()
for s in range(K):
    env.step(acts[s % 64])
T1 = time.perf_counter()
print("env.step per step us:", round(1e6 * (T1 - T0) / K, 1))
T0 = time.perf_counter()
for s in range(K):
    b.env_step_fused(acts[s % 64], False, env.randomization_factor, env.max_simulation_time, env.seed, 1)
T1 = time.perf_counter()
print("fused C call (pinned action) per step us:", round(1e6 * (T1 - T0) / K, 1))
import torch
st = torch.cuda.ExternalStream(b.stream)
ring = torch.as_tensor(acts[:64]).cuda()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for s in range(K):
    b.step_dev(ring[s % 64].data_ptr(), 1)
e1.record(st)
b.sync()
print("device-only physics step (random ctrl ring) us:", round(1e3 * e0.elapsed_time(e1) / K, 1))
t0 = time.perf_counter()
for s in range(K):
    b.step_dev(ring[s % 64].data_ptr(), 1)
    b.sync()
print("physics step + sync each step us:", round(1e6 * (time.perf_counter() - t0) / K, 1))
import cProfile, pstats
pr = cProfile.Profile()
pr.enable()
for s in range(K):
    env.step(acts[s % 64])
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
