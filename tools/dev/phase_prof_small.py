"""Developer tool: per-phase latency of k_solve for a SMALL settled batch (little SM contention, contact-rich)."""
import ctypes as C, os, sys
import numpy as np
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))]
import bench
import humanoid_mujoco_b200 as hmj

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
b = hmj.HumanoidBatch(n)
params, qpos = bench.make_workload(b.default_params(), b.model["qpos0"][0], n, 1)
b.set_params(params)
b.set_state(qpos, np.zeros((n, 18)), np.zeros(n), np.zeros((n, 18)))
b.step(np.zeros((n, 12)), 300)
rng = np.random.default_rng(2)
out = (C.c_ulonglong * 16)()
b.lib.hmj_debug_phase_cycles(out, 1)
K = 50
it = 0
nc = 0
for _ in range(K):
    b.step(rng.uniform(-0.2, 0.2, (n, 12)), 1)
    it += b.get("niter").sum(); nc += b.get("ncon").sum()
b.lib.hmj_debug_phase_cycles(out, 0)
names = {1: "load+chol0+solve", 2: "contacts+limits", 3: "Jb+rows", 4: "warmstart", 5: "eval0: solve tail", 6: "newton-pre (Mv,basis)",
         7: "line search", 8: "evalN: solve tail", 9: "post-newton", 10: "results+checkacc", 11: "euler chol+solve",
         12: "eval: cost+H assembly", 13: "eval: chol+solve", 14: "eval: qfc+grad"}
tot = sum(out)
evals = it + n * K
print(f"n={n} mean niter {it / (n * K):.2f} mean ncon {nc / (n * K):.2f}")
for i in range(1, 15):
    per = out[i] / (n * K)
    extra = ""
    if i in (12, 13, 14): extra = f"  -> {out[i] / evals:8.0f} per evaluate"
    if i in (6, 7): extra = f"  -> {out[i] / max(it, 1):8.0f} per iteration"
    print(f"{names[i]:28s} {per:10.0f} cycles/env-step  {100 * out[i] / tot:5.1f}%{extra}")
print("total per env-step", tot / (n * K))
