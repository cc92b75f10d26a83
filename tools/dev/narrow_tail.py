"""Developer tool: per-group item statistics of k_narrow over a few 4096-env steps at the steady state of a bench regime
(prof build: tools/dev/build_variant.sh prof -DHMJ_PHASE_PROF, run with HMJ_B200_LIB=.../lib_prof/libhmj_b200.so HMJ_PIPES=1).
usage: narrow_tail.py [control scale] [lead-in steps]"""
import ctypes as C, os, sys
import numpy as np
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))]
import bench
import humanoid_mujoco_b200 as hmj

scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
lead = int(sys.argv[2]) if len(sys.argv) > 2 else 150
n = 4096
b = hmj.HumanoidBatch(n)
z18 = np.zeros((n, 18))
def probe(params, qpos):
    b.set_params(params); b.set_state(qpos, z18, np.zeros(n), z18); b.step(np.zeros((n, 12)), 1); return b.get("ncon") > 0
params, qpos, attempts = bench.make_workload(b.default_params(), b.model["qpos0"][0], n, 1, probe)
b.set("status", np.zeros(n, dtype=np.int32))
b.set_state(qpos, z18, np.zeros(n), z18)
b.step(np.zeros((n, 12)), bench.SETTLE_STEPS)
rng = np.random.default_rng(2)
for _ in range(lead):
    b.step(scale * rng.uniform(-1, 1, (n, 12)), 1)
out = np.zeros((64, 5), dtype=np.uint64)
warp = np.zeros(12, dtype=np.uint64)
b.lib.hmj_debug_narrow_prof.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
b.lib.hmj_debug_narrow_prof(out.ctypes.data, warp.ctypes.data, 1)
K = 10
for _ in range(K):
    b.step(scale * rng.uniform(-1, 1, (n, 12)), 1)
b.lib.hmj_debug_narrow_prof(out.ctypes.data, warp.ctypes.data, 0)
print(f"per launch: busy warps {warp[2] / K:.0f}, max warp busy {warp[0]} cycles (max over {K} launches), mean {warp[1] / max(warp[2], 1):.0f}, staging max {warp[3]}")
print("launches by item size 4/8/16/32:", warp[6:10].tolist(), " items/launch", warp[10] / K, " queries/launch", warp[11] / K)
print("group  items/launch  mean cycles  max cycles  mean trips  max trips")
for g in range(64):
    if out[g, 0]:
        print(f"{g:5d} {out[g, 0] / K:12.1f} {out[g, 1] / out[g, 0]:12.0f} {out[g, 2]:11d} {out[g, 3] / out[g, 0]:11.1f} {out[g, 4]:10d}")
q = np.zeros(12, dtype=np.uint64)
if hasattr(b.lib, "hmj_debug_narrow_q"):
    b.lib.hmj_debug_narrow_q.argtypes = [C.c_void_p, C.c_int]
    b.lib.hmj_debug_narrow_q(q.ctypes.data, 0)
    q = q.astype(np.float64)
    # counters run since process start: report fractions
    print(f"queries with contact: {q[0]:.0f} (mean trips {q[1]/max(q[0],1):.1f}, max {q[2]:.0f}); without: {q[3]:.0f} (mean trips {q[4]/max(q[3],1):.1f}, max {q[5]:.0f})")
    print("no-contact exits proving a separation gap > 1e-9 / 1e-6 / 1e-4 / 1e-3:", (q[6:10] / max(q[3], 1)).round(3).tolist())
if hasattr(b.lib, "hmj_debug_sup_stat"):
    st = np.zeros(10, dtype=np.uint64)
    b.lib.hmj_debug_sup_stat.argtypes = [C.c_void_p]
    b.lib.hmj_debug_sup_stat(st.ctypes.data)
    st = st.astype(np.float64)
    print(f"support calls {st[0]:.0f}, mean candidates {st[1]/max(st[0],1):.1f}; list length histogram by 8: {(st[2:]/max(st[0],1)).round(3).tolist()}")
