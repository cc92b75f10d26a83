"""Developer tool: per-phase latency of k_solve for the SLOWEST envs of one 4096-env step (the single-wave tail).
Build: make -C humanoid-mujoco_b200/csrc OUT=../lib_prof/libhmj_b200.so OBJ=../lib_prof/obj XFLAGS=-DHMJ_PHASE_PROF
Run:   HMJ_B200_LIB=humanoid-mujoco_b200/lib_prof/libhmj_b200.so HMJ_PIPES=1 python tools/dev/phase_tail.py"""
import ctypes as C, os, sys
import numpy as np
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))]
import bench
import humanoid_mujoco_b200 as hmj

n = 4096
scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
b = hmj.HumanoidBatch(n)
z18 = np.zeros((n, 18))
def probe(params, qpos):
    b.set_params(params); b.set_state(qpos, z18, np.zeros(n), z18); b.step(np.zeros((n, 12)), 1); return b.get("ncon") > 0
params, qpos, _att = bench.make_workload(b.default_params(), b.model["qpos0"][0], n, 1, probe)
b.set("status", np.zeros(n, dtype=np.int32))
b.set_state(qpos, z18, np.zeros(n), z18)
b.step(np.zeros((n, 12)), 300)
rng = np.random.default_rng(2)
for _ in range(150):
    b.step(rng.uniform(-scale, scale, (n, 12)), 1)
names = {1: "load", 2: "contacts+limits", 3: "Jb+rows", 4: "warmstart", 6: "newton-pre", 7: "line search", 5: "test+W", 8: "H init", 9: "post-newton",
         10: "results", 11: "euler", 12: "H assembly", 13: "chol+solve", 14: "cost/qfc/grad"}
out = np.zeros((n, 16), dtype=np.uint32)
b.lib.hmj_debug_phase_env.argtypes = [C.c_void_p, C.c_int, C.c_int]
for rep in range(3):
    b.lib.hmj_debug_phase_env(out.ctypes.data, n, 1)
    b.step(rng.uniform(-scale, scale, (n, 12)), 1)
    b.lib.hmj_debug_phase_env(out.ctypes.data, n, 0)
    tot = out[:, :15].sum(1)
    niter, ncon, nefc = b.get("niter"), b.get("ncon"), b.get("nefc")
    order = np.argsort(-tot.astype(np.int64))
    print(f"step {rep}: cycles per env  max {tot.max()}  p99 {np.percentile(tot, 99):.0f}  p90 {np.percentile(tot, 90):.0f}  median {np.median(tot):.0f}")
    cls = np.where((ncon == 0) & (nefc == 0), 0, np.where(ncon <= 4, 1, np.where(ncon <= 8, 2, 3)))
    print("   class sizes", np.bincount(cls, minlength=4).tolist(), " niter hist", np.bincount(niter, minlength=12).tolist())
    for c in (1, 2):
        sel = cls == c
        if sel.any():
            print(f"   class {c}: max {tot[sel].max()} median {np.median(tot[sel]):.0f}; mean per-phase: "
                  + ", ".join(f"{names[i]} {out[sel][:, i].mean():.0f}" for i in names))
    for e in order[:6]:
        ph = ", ".join(f"{names[i]} {out[e, i]}" for i in names if out[e, i] > 0.03 * tot[e])
        print(f"  env {e}: total {tot[e]}  niter {niter[e]} ncon {ncon[e]} nefc {nefc[e]} ls evals {out[e, 15]} | {ph}")
