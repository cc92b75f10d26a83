"""humanoid-mujoco_b200 — B200-native batched humanoid physics-and-rollout engine.

Host-side mirror of the two reference seams:
  * `BatchedHumanoidEnv` — the CPUEnv reset/step/observation contract (simulation/cpu_env.py) over N
    lock-step environments;
  * `B200VecEnv` — the `DummyVecEnv` seam (rl/train.py:134-151): auto-reset, `terminal_observation`, `set_attr`;
  * `SamplingPlanner` — mjpc's predictive-sampling planner (mjpc/planners/sampling/planner.cc) with
    `Rollouts` running as a CUDA grid over samples.
Both call the same C ABI (include/hmj_b200.h) through ctypes.  There is no CPU fallback: importing the
package is cheap, creating a batch without the CUDA library or a GPU raises.

The directory name has a hyphen (it is the name the build contract asks for); import it as
`humanoid_mujoco_b200` (see humanoid_mujoco_b200.py at the repo root).
"""
from ._lib import HmjError, lib_path, load_library  # noqa: F401
from .batch import HumanoidBatch, default_params, load_model, pinned_empty  # noqa: F401
from .env import BatchedHumanoidEnv  # noqa: F401
from .planner import SamplingPlanner, SamplingPolicy, TimeSpline  # noqa: F401
from .planner_ce import CrossEntropyPlanner  # noqa: F401
from .planner_robust import RobustPlanner  # noqa: F401
from .vec_env import B200VecEnv  # noqa: F401

__all__ = [
    "HmjError", "load_library", "lib_path", "HumanoidBatch", "default_params", "load_model",
    "BatchedHumanoidEnv", "B200VecEnv", "SamplingPlanner", "SamplingPolicy", "TimeSpline", "CrossEntropyPlanner", "RobustPlanner",
]
