#!/usr/bin/env python
"""Generate golden vectors for the CPUEnv reward / observation epilogue from the REFERENCE'S OWN Python.

Runs in the build container only (needs /root/reference).  What is executed, unmodified and where it lies:

  * simulation/reward_functions.py (standupReward :247-374 and its helpers :17-113) and
    simulation/simulation_parameters.py — imported as `simulation.reward_functions` through a bare
    `simulation` package object (the real simulation/__init__.py imports gymnasium, absent here);
  * the `inverseRotateVectors` lambda of simulation/cpu_env.py:26-28 — cpu_env.py cannot be imported
    (mujoco, gymnasium, cv2, perlin_noise are absent), so the one assignment statement is located with `ast`
    and evaluated from the file's own source text.

The only stand-in is `jax.scipy.spatial.transform.Rotation` -> `scipy.spatial.transform.Rotation` (JAX is not
installed; scipy's class is the fp64 model of the same API: from_quat([x, y, z, w]) normalises, inv(), apply()).
JAX would run it in float32 unless jax_enable_x64 is set; the golden is the fp64 answer.

The observation's arithmetic (`_get_obs`, cpu_env.py:465-571: noisy joint angles / velocities / gyro in
JOINT_NAMES order + inverse-rotated gravity of the noisy quaternion) is restated here term by term with the
reference's own JOINT_NAMES and the joint order of simulation/assets/humanoid.xml, and uses the reference's
inverseRotateVectors for the gravity part.

Writes tests/golden/env_golden.npz (inputs + outputs), committed so the pin travels to machines without the
reference tree.
"""
import ast
import os
import sys
import types
import xml.etree.ElementTree as ET

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("HMJ_REFERENCE", "/root/reference")
N = 2048


def import_reference():
    from scipy.spatial.transform import Rotation

    for name in ("jax", "jax.scipy", "jax.scipy.spatial", "jax.scipy.spatial.transform"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["jax.scipy.spatial.transform"].Rotation = Rotation
    pkg = types.ModuleType("simulation")
    pkg.__path__ = [os.path.join(REF, "simulation")]   # bare package: submodules load from the reference tree
    sys.modules["simulation"] = pkg
    import simulation.reward_functions as rf           # noqa: E402  (the reference's file, unmodified)
    import simulation.simulation_parameters as sp      # noqa: E402

    src = open(os.path.join(REF, "simulation", "cpu_env.py")).read()
    node = next(n for n in ast.parse(src).body
                if isinstance(n, ast.Assign) and getattr(n.targets[0], "id", "") == "inverseRotateVectors")
    ns = {"Rotation": Rotation}
    exec(compile(ast.Module(body=[node], type_ignores=[]), "cpu_env.py", "exec"), ns)   # noqa: S102 (reference source)
    return rf, sp, ns["inverseRotateVectors"]


def joint_indices(sp):
    """joint_qpos_idx / joint_dof_idx of cpu_env.py:111-122 from the MJCF joint order (free joint first)."""
    root = ET.parse(os.path.join(REF, "simulation", "assets", "humanoid.xml")).getroot()
    names = [j.get("name") for j in root.iter("joint") if j.get("type") != "free"]
    qadr = {n: 7 + i for i, n in enumerate(names)}
    dadr = {n: 6 + i for i, n in enumerate(names)}
    return np.array([qadr[n] for n in sp.JOINT_NAMES]), np.array([dadr[n] for n in sp.JOINT_NAMES])


def main():
    rf, sp, inv_rot = import_reference()
    jq, jd = joint_indices(sp)
    rng = np.random.default_rng(20261105)
    qpos = np.zeros((N, 19)); qvel = np.zeros((N, 18))
    qpos[:, :3] = rng.uniform(-1, 1, (N, 3))
    qpos[:, 3:7] = rng.standard_normal((N, 4))
    qpos[:, 7:] = rng.uniform(-1.6, 1.6, (N, 12))
    qvel[:] = rng.standard_normal((N, 18)) * rng.choice([0.05, 0.5, 5.0], (N, 1))
    # data.xquat[torso] is unit (one step stale); a third of the cases carry non-unit "noisy" quaternions as well
    quat = rng.standard_normal((N, 4))
    quat /= np.linalg.norm(quat, axis=1, keepdims=True)
    quat[::3] *= rng.uniform(0.5, 1.5, (len(quat[::3]), 1))
    quat[1::7] = np.array([-0.5, -0.5, 0.5, 0.5]) + rng.uniform(-0.1, 0.1, (len(quat[1::7]), 4))   # standup start
    quat[2::7] = np.array([-0.70710678, 0, 0, 0.70710678]) + rng.uniform(-0.05, 0.05, (len(quat[2::7]), 4))  # upright
    z = rng.uniform(-0.75, -0.25, N)
    z[::11] = -0.375; z[1::11] = -0.6; z[2::11] = rng.uniform(-0.3749, -0.30, len(z[2::11]))
    # joint torques on both sides of MAX__SAFE_JOINT_TORQUE = 1 N m, split into the two arrays the env adds
    tau = rng.standard_normal((N, 18)) * rng.choice([0.3, 1.0, 3.0], (N, 1))
    qfrc_constraint = rng.standard_normal((N, 18)) * 0.5
    qfrc_smooth = tau - qfrc_constraint
    prev = rng.uniform(-1.2, 1.2, (N, 12))              # un-clipped actions (cpu_env.py:670-671)
    latest = prev + rng.standard_normal((N, 12)) * rng.choice([1e-4, 1e-3, 1e-2, 0.3], (N, 1))
    latest[::5] = rng.uniform(-1.2, 1.2, (len(latest[::5]), 12))
    selfc = rng.random(N) < 0.3
    time = rng.uniform(0, 12, N)
    time[::13] = 10.0; time[1::13] = 10.0 - 1e-12
    max_time = 10.0
    noise = np.zeros((N, 31))
    noise[N // 4:] = rng.standard_normal((N - N // 4, 31)) * np.deg2rad(5)

    reward = np.zeros(N); term = np.zeros(N, dtype=bool); trunc = np.zeros(N, dtype=bool)
    obs = np.zeros((N, 30)); grav = np.zeros((N, 3))
    for i in range(N):
        tq = (qfrc_constraint[i] + qfrc_smooth[i])[jd]                       # _get_joint_torques, cpu_env.py:424-428
        r, t_, u_ = rf.standupReward(qvel[i, :3], np.array([0, 0]), quat[i], np.array([0]), z[i], tq,
                                     prev[i] / (np.pi / 2), latest[i] / (np.pi / 2), bool(selfc[i]), time[i], max_time)
        reward[i], term[i], trunc[i] = float(r), bool(t_), bool(u_)
        grav[i] = inv_rot(quat[i], np.array([0, 0, -1]))
        nz = noise[i]
        obs[i] = np.concatenate((qpos[i][jq] + nz[0:12], qvel[i][jd] + nz[12:24], qvel[i, 3:6] + nz[24:27],
                                 inv_rot(quat[i] + nz[27:31], np.array([0, 0, -1]))))
    out = os.path.join(HERE, "env_golden.npz")
    np.savez_compressed(out, qpos=qpos, qvel=qvel, quat=quat, z=z, qfrc_constraint=qfrc_constraint, qfrc_smooth=qfrc_smooth,
                        prev_action=prev, latest_action=latest, self_collision=selfc, time=time, max_time=np.float64(max_time),
                        noise=noise, reward=reward, terminated=term, truncated=trunc, obs=obs, local_gravity=grav,
                        joint_qpos_idx=jq, joint_dof_idx=jd)
    print(f"wrote {out}: {N} cases, reward in [{reward.min():.3f}, {reward.max():.3f}], "
          f"terminated {int(term.sum())}, truncated {int(trunc.sum())}, self-colliding {int(selfc.sum())}")


if __name__ == "__main__":
    main()
