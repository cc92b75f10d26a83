"""Developer tool: bring the bench workload to its steady state and run a few more steps (target of ncu captures).
usage: prof_steps.py <scale> <lead_in> <steps>   (HMJ_PIPES=1 HMJ_GRAPH=0 for a readable launch list)"""
import os, sys
import numpy as np
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))]
import bench
import humanoid_mujoco_b200 as hmj

scale, lead, steps = float(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
n = 4096
b = hmj.HumanoidBatch(n)
z18 = np.zeros((n, 18))
def probe(params, qpos):
    b.set_params(params); b.set_state(qpos, z18, np.zeros(n), z18); b.step(np.zeros((n, 12)), 1); return b.get("ncon") > 0
params, qpos, attempts = bench.make_workload(b.default_params(), b.model["qpos0"][0], n, 1, probe)
b.set("status", np.zeros(n, dtype=np.int32))
b.set_state(qpos, z18, np.zeros(n), z18)
b.step(np.zeros((n, 12)), bench.SETTLE_STEPS)
ring = bench.control_ring(0, 64, n) * scale
for s in range(lead + steps):
    b.step(ring[s % 64], 1)
print("attempts", attempts, "launches before the last", steps, "steps:", attempts + bench.SETTLE_STEPS + lead,
      "ncon", b.get("ncon").mean(), "nefc", b.get("nefc").mean())
