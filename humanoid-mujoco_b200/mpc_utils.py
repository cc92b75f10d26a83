"""The MPC state-estimation workaround of the reference (controllers/mpc_utils.py:36-55) on the batched engine.

`find_ideal_distance_to_ground` sweeps 200 candidate heights of the robot and keeps the one where the vertical component of
`mj_inverse`'s force at zero acceleration is smallest — 200 `mj_resetDataKeyframe` + `mj_forward` + `mj_inverse` calls per control
step in the reference.  Here the 200 candidates are the environments of one batch and the sweep is ONE `hmj_batch_inverse` call.
"""
from __future__ import annotations

import numpy as np

from .batch import HumanoidBatch

N_OFFSETS = 200


def find_ideal_distance_to_ground(batch: HumanoidBatch, base_qpos, joint_positions, torso_orientation, joint_qpos_idx):
    """`batch` must hold N_OFFSETS envs.  `base_qpos` [19] plays the keyframe's role (mj_resetDataKeyframe, mpc_utils.py:42);
    returns the height offset in [-0.1, 0.1] with the smallest |qfrc_inverse[2]| (mpc_utils.py:50-55)."""
    if batch.n != N_OFFSETS:
        raise ValueError(f"the height sweep uses {N_OFFSETS} environments")
    offsets = np.linspace(-0.1, 0.1, N_OFFSETS)
    q = np.tile(np.asarray(base_qpos, dtype=np.float64), (N_OFFSETS, 1))
    q[:, 3:7] = torso_orientation
    q[:, joint_qpos_idx] = joint_positions
    q[:, 2] += offsets
    n = N_OFFSETS
    batch.set_state(q, np.zeros((n, 18)), np.zeros(n), np.zeros((n, 18)))
    batch.set_ctrl(np.zeros((n, 12)))
    vertical = batch.inverse(None)[:, 2]            # mj_data.qacc = 0 (mpc_utils.py:44)
    return float(offsets[int(np.argmin(np.abs(vertical)))]), vertical
