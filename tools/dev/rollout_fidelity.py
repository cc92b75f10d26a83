"""Developer tool: how many of N predictive-sampling rollouts (H = 18, C2 start) differ from the oracle by more than 1e-6
relative (make/break-contact events amplify last-bit differences).  Oracle returns are cached in /tmp between runs."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [REPO, os.path.join(REPO, "oracle"), os.path.join(REPO, "tests")]
import humanoid_mujoco_b200 as hmj
from oracle_py import Oracle
from test_gpu_planner import settled_task_state, oracle_returns

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
o = Oracle()
state, time = settled_task_state(o)
pl = hmj.SamplingPlanner(num_trajectory=n, device=0, seed=0)
pl.SetState(state[:19], state[19:], time)
H = pl.steps()
pl.UpdateNominalPolicy(H)
rng = np.random.default_rng(11)
scale = 0.5 * (pl.ctrlrange[:, 1] - pl.ctrlrange[:, 0])
noise = rng.standard_normal((n, pl.num_spline_points, 12)) * scale * 0.3
pl.Rollouts(H, noise=noise)
cache = f"/tmp/fid_{n}.npy"
if os.path.exists(cache):
    ref = np.load(cache)
else:
    ref, _ = oracle_returns(o, state, time, pl.candidate_times, pl.candidate_values, H, pl.interpolation_)
    np.save(cache, ref)
rel = np.abs(pl.total_return - ref) / np.maximum(np.abs(ref), 1.0)
print(os.environ.get("HMJ_B200_LIB", "default lib"), f"n={n}: rel>1e-6: {(rel > 1e-6).sum()}  rel>1e-9: {(rel > 1e-9).sum()}  max {rel.max():.3e}  median {np.median(rel):.3e}")
