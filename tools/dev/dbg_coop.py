import os, sys
import numpy as np
sys.path[:0] = ['/root/repo', '/root/repo/tools']
import bench
import humanoid_mujoco_b200 as hmj
n = 96
b = hmj.HumanoidBatch(n)
z18 = np.zeros((n, 18))
def probe(p, q):
    b.set_params(p); b.set_state(q, z18, np.zeros(n), z18); b.step(np.zeros((n, 12)), 1); return b.get("ncon") > 0
os.environ["HMJ_BROAD_COOP"] = "0"
params, qpos, att = bench.make_workload(b.default_params(), b.model["qpos0"][0], n, 1, probe)
b.set_params(params); b.set("status", np.zeros(n, dtype=np.int32)); b.set_state(qpos, z18, np.zeros(n), z18)
b.step(np.zeros((n, 12)), 300)
ring = bench.control_ring(0, 64, n)
ndiff = 0
for s in range(170):
    st = b.get_state()
    res = {}
    for mode in ("0", "1"):
        os.environ["HMJ_BROAD_COOP"] = mode
        b.set_state(*st)
        b.set_ctrl(ring[s % 64])
        b.run_stage(1); b.run_stage(2)
        res[mode] = (b.get_raw("slot_mask", np.uint32, (n, 4)).copy(), b.get_raw("wq_count", np.int32, (64 * 9,))[:64].copy())
    if not np.array_equal(res["0"][0], res["1"][0]) or not np.array_equal(res["0"][1], res["1"][1]):
        ndiff += 1
        bad = np.nonzero((res["0"][0] != res["1"][0]).any(axis=1))[0]
        print("step", s, "slot masks differ for envs", bad[:6], "counts old", res["0"][1][:8], "coop", res["1"][1][:8])
        for e in bad[:2]:
            print("   env", e, "old", [hex(x) for x in res["0"][0][e]], "coop", [hex(x) for x in res["1"][0][e]], "status", b.get("status")[e])
        if ndiff > 3: break
    os.environ["HMJ_BROAD_COOP"] = "0"
    b.set_state(*st)
    b.step(ring[s % 64], 1)
print("done, differing steps:", ndiff)
