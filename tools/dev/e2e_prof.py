"""Developer tool: where the end-to-end control step spends its time beyond the device step (bench workload C3).
usage: e2e_prof.py [scale] [steps]"""
import os, sys, time
import numpy as np
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))]
import bench
import humanoid_mujoco_b200 as hmj

scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
K = int(sys.argv[2]) if len(sys.argv) > 2 else 300
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
acts = hmj.pinned_empty((64, n, 12))
acts[...] = np.random.default_rng(5).uniform(-scale, scale, acts.shape)
for s in range(150):
    env.step(acts[s % 64])
lib, h = b.lib, b._h
calls = []
orig = lib.hmj_env_step_fused_from
def timed(*a):
    t = time.perf_counter(); r = orig(*a); calls.append(time.perf_counter() - t); return r
b.lib.hmj_env_step_fused_from = timed
t0 = time.perf_counter()
for s in range(K):
    env.step(acts[s % 64])
tot = (time.perf_counter() - t0) / K
b.lib.hmj_env_step_fused_from = orig
c = float(np.mean(calls))
# device-only steps (same controls already on the device)
b.sync()
t0 = time.perf_counter()
for s in range(K):
    b.step_dev(None, 1)
b.sync()
dev = (time.perf_counter() - t0) / K
# replayed controls (no H2D)
t0 = time.perf_counter()
for s in range(K):
    env.step(None)
rep = (time.perf_counter() - t0) / K
print(f"env.step {tot*1e6:.1f} us | C call {c*1e6:.1f} us | python around it {(tot-c)*1e6:.1f} us | device-only step loop {dev*1e6:.1f} us | env.step(None) {rep*1e6:.1f} us")
