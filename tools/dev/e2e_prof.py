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
acts = np.random.default_rng(5).uniform(-1, 1, (64, n, 12))
for s in range(20):
    env.step(acts[s])
K = 300
t = {"stage": 0.0, "call": 0.0, "finish": 0.0, "infos": 0.0}
io = env.io
T0 = time.perf_counter()
for s in range(K):
    a = time.perf_counter()
    io["action"][...] = acts[s % 64]
    c = time.perf_counter()
    b.env_step_fused(True, False, env.randomization_factor, env.max_simulation_time, env.seed, 1)
    d = time.perf_counter()
    r = env._finish(io["obs"], io["reward"], io["terminated"].astype(bool), io["truncated"].astype(bool), io["self_collision"].astype(bool))
    e = time.perf_counter()
    infos = [env._empty_info] * n
    done = r[2] | r[3]
    if done.any():
        for i in np.nonzero(done)[0]:
            infos[i] = {"is_success": bool(r[3][i])}
    f = time.perf_counter()
    t["stage"] += c - a; t["call"] += d - c; t["finish"] += e - d; t["infos"] += f - e
T1 = time.perf_counter()
print("per step us:", {k: round(1e6 * v / K, 1) for k, v in t.items()}, "total", round(1e6 * (T1 - T0) / K, 1))
# the same C call with no host copies: physics only via step_dev on the resident ctrl
import torch
st = torch.cuda.ExternalStream(b.stream)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ctrl = torch.zeros((n, 12), dtype=torch.float64, device="cuda")
e0.record(st)
for s in range(K):
    b.step_dev(ctrl.data_ptr(), 1)
e1.record(st)
b.sync()
print("device-only physics step us:", round(1e3 * e0.elapsed_time(e1) / K, 1))
t0 = time.perf_counter()
for s in range(K):
    b.step_dev(ctrl.data_ptr(), 1)
    b.sync()
print("physics step + sync each step us:", round(1e6 * (time.perf_counter() - t0) / K, 1))
t0 = time.perf_counter()
for s in range(K):
    b.env_step_fused(False, False, 1.0, env.max_simulation_time, env.seed, 1)
print("fused call, no action copy us:", round(1e6 * (time.perf_counter() - t0) / K, 1))
