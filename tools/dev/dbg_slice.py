"""Developer tool: find where a bench-workload env diverges between GPU and oracle over 10 steps."""
import os, sys
import numpy as np
R = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [R, os.path.join(R, "tools"), os.path.join(R, "oracle")]
import bench
import humanoid_mujoco_b200 as hmj
from oracle_py import Oracle

oracle = Oracle()
n = 96
b = hmj.HumanoidBatch(n)
z18 = np.zeros((n, 18))
def probe(p, q):
    b.set_params(p); b.set_state(q, z18, np.zeros(n), z18); b.step(np.zeros((n, 12)), 1); return b.get("ncon") > 0
params, qpos, _ = bench.make_workload(b.default_params(), b.model["qpos0"][0], n, 1, probe)
b.set_params(params); b.set("status", np.zeros(n, dtype=np.int32)); b.set_state(qpos, z18, np.zeros(n), z18)
b.step(np.zeros((n, 12)), 300)
ring = bench.control_ring(0, 64, n)
lead = 150
for s in range(lead):
    b.step(ring[s % 64], 1)
q0, v0, t0, w0 = b.get_state()
states = []
for s in range(10):
    b.step(ring[(lead + s) % 64], 1)
    q, v, _, w = b.get_state()
    states.append((q.copy(), v.copy(), b.get("ncon").copy(), b.get("nefc").copy(), b.get("niter").copy(), b.get("status").copy(), w.copy()))
for i in range(0, n, 2):
    oracle.reset(); oracle.params[0] = params[i]
    oracle.set_state(q0[i], v0[i], ring[lead % 64][i], warmstart=w0[i], time=t0[i])
    bad = False
    for s in range(10):
        oracle.data["ctrl"][0] = ring[(lead + s) % 64][i]
        oracle.step(1)
        q, v, nc, ne, ni, st, w = states[s]
        e = max(np.max(np.abs(q[i] - oracle.d["qpos"])), np.max(np.abs(v[i] - oracle.d["qvel"])))
        if e > 1e-9 or bad:
            bad = True
            print(f"env {i} step {s}: err {e:.3e} ncon gpu {nc[i]} orc {int(oracle.d['ncon'])} nefc {ne[i]}/{int(oracle.d['nefc'])} niter {ni[i]}/{int(oracle.d['solver_niter'])} status {st[i]}/{int(oracle.d['status'])} "
                  f"werr {np.max(np.abs(w[i]-oracle.d['qacc_warmstart'])):.3e} max|qacc| {np.max(np.abs(w[i])):.3e}")

# sensitivity of the ORACLE ITSELF at the diverging step: same state, inputs nudged in the last bits
i, s = 28, 6
q, v, nc, ne, ni, st, w = states[s - 1]
def orc_from(qq, vv, ww):
    oracle.reset(); oracle.params[0] = params[i]
    oracle.set_state(qq, vv, ring[(lead + s) % 64][i], warmstart=ww, time=0.0)
    oracle.step(1)
    return oracle.d["qacc_warmstart"].copy(), int(oracle.d["solver_niter"]), int(oracle.d["ncon"])
base, nit0, nc0 = orc_from(q[i], v[i], w[i])
rng = np.random.default_rng(0)
for k in range(8):
    qq = q[i] * (1 + rng.uniform(-1, 1, 19) * 2e-16); vv = v[i] * (1 + rng.uniform(-1, 1, 18) * 2e-16)
    a, nit, ncc = orc_from(qq, vv, w[i])
    print(f"oracle, inputs nudged by 1 ulp: max|dqacc| {np.max(np.abs(a - base)):.3e} of {np.max(np.abs(base)):.3e}  niter {nit0}->{nit} ncon {nc0}->{ncc}")
