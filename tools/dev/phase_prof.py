"""Developer tool: per-phase latency of k_solve (build with -DHMJ_PHASE_PROF, load via HMJ_B200_LIB)."""
import ctypes as C, os, sys
import numpy as np
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))]
import bench
import humanoid_mujoco_b200 as hmj

n = 4096
env = hmj.BatchedHumanoidEnv(n, randomization_factor=1.0, seed=1)
b = env.batch
params, qpos = bench.make_workload(b.default_params(), b.model["qpos0"][0], n, 1)
b.set_params(params)
b.set_state(qpos, np.zeros((n, 18)), np.zeros(n), np.zeros((n, 18)))
b.step(np.zeros((n, 12)), 300)
rng = np.random.default_rng(2)
for _ in range(20):
    b.step(rng.uniform(-1, 1, (n, 12)), 1)
out = (C.c_ulonglong * 16)()
b.lib.hmj_debug_phase_cycles(out, 1)
K = 50
for _ in range(K):
    b.step(rng.uniform(-1, 1, (n, 12)), 1)
b.lib.hmj_debug_phase_cycles(out, 0)
names = {1: "load+chol0+solve", 2: "contacts+limits", 3: "Jb+rows", 4: "warmstart", 5: "eval0: solve tail", 6: "newton-pre (Mv,basis)",
         7: "line search", 8: "evalN: solve tail", 9: "post-newton", 10: "results+checkacc", 11: "euler chol+solve",
         12: "eval: cost+H assembly", 13: "eval: chol+solve", 14: "eval: qfc+grad"}
tot = sum(out)
for i in range(1, 15):
    print(f"{names[i]:28s} {out[i] / (n * K):10.0f} cycles/env-step  {100 * out[i] / tot:5.1f}%")
print("total per env-step", tot / (n * K), "niter mean", b.get("niter").mean(), "ncon mean", b.get("ncon").mean())
