"""BatchedHumanoidEnv — the CPUEnv contract (simulation/cpu_env.py) over N lock-step environments.

Same constructor arguments, same `reset` / `step` semantics and the same observation (30,) /
action (12,) layout as `CPUEnv` (cpu_env.py:34-63,374-416,676-693), with a leading batch axis.
`mujoco.mj_step` (cpu_env.py:684) is replaced by ONE call into the C ABI for all envs; the
observation / standupReward epilogue (cpu_env.py:465-616, reward_functions.py:247-374) runs as a
CUDA kernel reading the same one-step-stale quantities as the reference.

Reference quirks kept on purpose: per-episode CONSTANT noise vectors (the JAX key
is never split, cpu_env.py:471-513,664), un-clipped actions (:670-671), DR without mj_setConst
(:253-265), `gainprm[0] = 2 +- 0.5 rf` even at rf = 0 (:215-221), the reset rejection loop on
first-step contact (:411-414), "terminated" = time-out with reward -100, "truncated" = success.
Differences, documented: Python `random` / `perlin_noise` / JAX threefry draws are replaced by a
seeded numpy Generator (those streams are unpinned in the reference: unseeded `random`, JAX absent
here); Rotation math is fp64 instead of JAX's default fp32.
"""
from __future__ import annotations

import numpy as np

from .batch import DEFINES, NQ, NU, NV, HumanoidBatch

# simulation/simulation_parameters.py
FLOOR_FRICTION_MIN_MULTIPLIER, FLOOR_FRICTION_MAX_MULTIPLIER = 0.5, 1.0
MIN_DELAY, MAX_DELAY = 0.01, 0.05
MAX_MASS_CHANGE_PER_LIMB, MAX_EXTERNAL_MASS_ADDED = 0.05, 0.2
MIN_EXTERNAL_FORCE_DURATION, MAX_EXTERNAL_FORCE_DURATION = 0.05, 0.15
MIN_EXTERNAL_FORCE_MAGNITUDE, MAX_EXTERNAL_FORCE_MAGNITUDE = 5, 15
MIN_EXTERNAL_FORCE_INTERVAL, MAX_EXTERNAL_FORCE_INTERVAL = 1, 3
JOINT_INITIAL_OFFSET_MAX = 0.2
Z_POSITION_INITIAL_OFFSET_MAX = 0.1
QUAT_INITIAL_OFFSET_MAX = 0.1
JOINT_ARMATURE_MAX_CHANGE = 0.0005
JOINT_RANGE_MAX_CHANGE = 0.1
JOINT_STIFFNESS_MAX_CHANGE = 0
JOINT_MARGIN_MAX_CHANGE = 0.05
JOINT_FORCE_LIMIT_MAX_CHANGE = 0.05
JOINT_P_GAIN, JOINT_P_GAIN_MAX_CHANGE = 2, 0.5
JOINT_ANGLE_NOISE_STDDEV, JOINT_VELOCITY_NOISE_STDDEV = 2, 5
IMU_NOISE_STDDEV, GYRO_NOISE_STDDEV = 5, 2
JOINT_ACTION_NOISE_STDDEV = 0.5
MAX_FLOOR_BUMP_HEIGHT = 0.1
TIMESTEP = 0.002
MAX_SIM_TIME_STANDUP = 10.0
Z_INITIAL_POS_STANDUP = -0.6
INITIAL_QUAT_STANDUP = [-0.5, -0.5, 0.5, 0.5]
MAX_DELAY_STEPS = 25


def _smooth_heightfield(rng: np.random.Generator, n: int) -> np.ndarray:
    """Documented substitute for PerlinNoise(octaves=15) on the 8x8 grid (cpu_env.py:267-280): a smooth
    seeded pseudo-noise normalised to [0, 1].  Returns [n][64] float32."""
    coarse = rng.uniform(0, 1, (n, 4, 4))
    up = np.repeat(np.repeat(coarse, 2, axis=1), 2, axis=2)
    k = np.array([0.25, 0.5, 0.25])
    pad = np.pad(up, ((0, 0), (1, 1), (1, 1)), mode="edge")
    sm = sum(k[i] * k[j] * pad[:, i:i + 8, j:j + 8] for i in range(3) for j in range(3))
    sm = sm + 0.05 * rng.uniform(0, 1, (n, 8, 8))
    sm -= sm.min(axis=(1, 2), keepdims=True)
    sm /= np.maximum(sm.max(axis=(1, 2), keepdims=True), 1e-12)
    return sm.reshape(n, 64).astype(np.float32)


def randomize_params(base: np.ndarray, rng: np.random.Generator, rf: float, timestep: float = TIMESTEP) -> np.ndarray:
    """CPUEnv._randomize_dynamics / _randomize_joint_properties / _randomize_floor_heightmap (cpu_env.py:189-280) for a
    batch: `base` is hmj_params_t[k] holding the un-randomised MjModel fields; returns the randomised copy.  Pure numpy."""
    k = base.shape[0]
    p = base.copy()
    mult = 1.0 * (1.0 - rf) + rng.uniform(FLOOR_FRICTION_MIN_MULTIPLIER, FLOOR_FRICTION_MAX_MULTIPLIER, (k, 1)) * rf
    p["floor_friction"] = base["floor_friction"] * mult
    mass = base["body_mass"].copy()
    mass[:, 1:] = np.maximum(1e-5, mass[:, 1:] + rng.uniform(-MAX_MASS_CHANGE_PER_LIMB * rf, MAX_MASS_CHANGE_PER_LIMB * rf, (k, 14)))
    extra_body = rng.integers(1, 15, k)
    mass[np.arange(k), extra_body] += rng.uniform(0, MAX_EXTERNAL_MASS_ADDED * rf, k)
    p["body_mass"] = mass
    hinge_dofs = np.arange(6, NV)
    arm = base["dof_armature"].copy()
    arm[:, hinge_dofs] += rng.uniform(0, JOINT_ARMATURE_MAX_CHANGE, (k, 12)) * rf
    p["dof_armature"] = arm
    stiff = base["jnt_stiffness"].copy()
    stiff[:, 1:] += rng.uniform(0, JOINT_STIFFNESS_MAX_CHANGE, (k, 12)) * rf
    p["jnt_stiffness"] = stiff
    margin = base["jnt_margin"].copy()
    margin[:, 1:] += rng.uniform(0, JOINT_MARGIN_MAX_CHANGE, (k, 12)) * rf
    p["jnt_margin"] = margin
    rngs = base["jnt_range"].copy()
    rngs[:, 1:, :] += rng.uniform(-JOINT_RANGE_MAX_CHANGE, JOINT_RANGE_MAX_CHANGE, (k, 12, 2)) * rf
    p["jnt_range"] = rngs
    p["act_gainprm0"] = JOINT_P_GAIN + rng.uniform(-JOINT_P_GAIN_MAX_CHANGE, JOINT_P_GAIN_MAX_CHANGE, (k, NU)) * rf
    p["act_forcerange"] = base["act_forcerange"] + rng.uniform(
        -JOINT_FORCE_LIMIT_MAX_CHANGE, JOINT_FORCE_LIMIT_MAX_CHANGE, (k, NU, 2)) * rf
    p["hfield_data"] = _smooth_heightfield(rng, k) * np.float32(MAX_FLOOR_BUMP_HEIGHT * rf)
    p["timestep"] = timestep
    return p


def standup_start_state(qpos0: np.ndarray, rng: np.random.Generator, rf: float, k: int) -> np.ndarray:
    """_randomize_joint_positions + _randomize_starting_position (cpu_env.py:282-328), standup reward selected."""
    qpos = np.tile(np.asarray(qpos0, dtype=np.float64), (k, 1))
    qpos[:, 7:] += rng.uniform(-JOINT_INITIAL_OFFSET_MAX * rf, JOINT_INITIAL_OFFSET_MAX * rf, (k, 12))
    qpos[:, 0] = 0.0
    qpos[:, 1] = 0.0
    qpos[:, 2] = Z_INITIAL_POS_STANDUP + rng.uniform(0, Z_POSITION_INITIAL_OFFSET_MAX * rf, k)
    qpos[:, 3:7] = np.array(INITIAL_QUAT_STANDUP) + rng.uniform(-QUAT_INITIAL_OFFSET_MAX * rf, QUAT_INITIAL_OFFSET_MAX * rf, (k, 4))
    return qpos


class BatchedHumanoidEnv:
    metadata = {"render_modes": [], "render_fps": 30}

    def __init__(self, n_envs: int, randomization_factor: float = 0.0, use_potential_rewards: bool = False,
                 max_simulation_time_override=None, reward_override=None, device: int = 0, seed: int = 0):
        self.num_envs = int(n_envs)
        self.randomization_factor = float(randomization_factor)
        self.use_potential_rewards = bool(use_potential_rewards)
        self.reward_override = reward_override
        self.timestep = TIMESTEP
        mst = MAX_SIM_TIME_STANDUP if max_simulation_time_override is None else max_simulation_time_override
        self.max_simulation_time = np.inf if mst < 0 else float(mst)
        self.rng = np.random.default_rng(seed)
        self.batch = HumanoidBatch(self.num_envs, device)
        self.model = self.batch.model
        self.base_params = self.batch.default_params()
        self.action_shape = (self.num_envs, NU)
        self.observation_shape = (self.num_envs, 30)
        n = self.num_envs
        self.params = self.base_params.copy()
        self.previous_action = np.zeros((n, NU))
        self.latest_action = np.zeros((n, NU))
        self.ctrl = np.zeros((n, NU))
        self.noise_action = np.zeros((n, NU))
        self.noise_obs = np.zeros((n, 31))
        self.action_delay = np.zeros(n, dtype=np.int64)
        self.obs_delay = np.zeros((n, 3), dtype=np.int64)  # joint, gyro, gravity
        self.action_hist = np.zeros((n, MAX_DELAY_STEPS + 1, NU))
        self.obs_hist = np.zeros((n, MAX_DELAY_STEPS + 1, 30))
        self.action_count = np.zeros(n, dtype=np.int64)
        self.obs_count = np.zeros(n, dtype=np.int64)
        self.previous_reward = np.zeros(n)
        self.force_start = np.zeros(n); self.force_duration = np.zeros(n); self.force_magnitude = np.zeros(n)
        self.force_direction = np.zeros((n, 2)); self.force_body = np.zeros(n, dtype=np.int32)
        self.xfrc = np.zeros((n, 6))
        self.torso_quat = np.zeros((n, 4))
        self.time = np.zeros(n)

    # ------------------------------------------------------------------ randomisers (cpu_env.py:135-362)
    def _randomize_model(self, idx):
        self.params[idx] = randomize_params(self.base_params[idx], self.rng, self.randomization_factor, self.timestep)

    def _randomize_delays(self, idx):
        rf, rng, k = self.randomization_factor, self.rng, len(idx)
        draw = lambda: np.rint(rng.uniform(MIN_DELAY * rf, MAX_DELAY * rf, k) / self.timestep).astype(np.int64)  # noqa: E731
        self.action_delay[idx] = draw()
        self.obs_delay[idx, 0] = draw()  # joint angles + velocities
        self.obs_delay[idx, 1] = draw()  # gyro
        draw()                           # torso_local_velocity_delay (drawn, unused by the observation)
        self.obs_delay[idx, 2] = draw()  # gravity vector
        draw()                           # contact_sensor_delay (unused)
        self.action_hist[idx] = 0.0
        self.obs_hist[idx] = 0.0
        self.obs_hist[idx, :, 27:30] = [0, 0, -1]
        self.action_count[idx] = 0
        self.obs_count[idx] = 0

    def _start_state(self, idx):
        return standup_start_state(self.model["qpos0"][0], self.rng, self.randomization_factor, len(idx))

    # ------------------------------------------------------------------ reset (cpu_env.py:374-416)
    def reset(self, seed=None, options=None, env_ids=None):
        if seed is not None:
            self.rng = np.random.default_rng(seed)
        idx = np.arange(self.num_envs) if env_ids is None else np.asarray(env_ids, dtype=np.int64)
        pending = idx
        obs_out = np.zeros((self.num_envs, 30))
        for _attempt in range(100):
            if len(pending) == 0:
                break
            rf = self.randomization_factor
            self._randomize_model(pending)
            self.batch.set_params(self.params)
            self._randomize_delays(pending)
            k = len(pending)
            # per-episode constant noise vectors (the JAX key is never split)
            self.noise_action[pending] = rf * np.deg2rad(JOINT_ACTION_NOISE_STDDEV) * self.rng.standard_normal((k, NU))
            nz = np.zeros((k, 31))
            same12 = self.rng.standard_normal((k, NU))
            nz[:, 0:12] = rf * np.deg2rad(JOINT_ANGLE_NOISE_STDDEV) * same12
            nz[:, 12:24] = rf * np.deg2rad(JOINT_VELOCITY_NOISE_STDDEV) * same12
            nz[:, 24:27] = rf * np.deg2rad(GYRO_NOISE_STDDEV) * self.rng.standard_normal((k, 3))
            nz[:, 27:31] = rf * np.deg2rad(IMU_NOISE_STDDEV) * self.rng.standard_normal((k, 4))
            self.noise_obs[pending] = nz
            qpos, qvel, time, ws = self.batch.get_state()
            qpos[pending] = self._start_state(pending)
            qvel[pending] = 0.0; time[pending] = 0.0; ws[pending] = 0.0
            self.batch.set_state(qpos, qvel, time, ws)
            self.ctrl[pending] = 0.0
            self.previous_action[pending] = 0.0; self.latest_action[pending] = 0.0
            self.force_start[pending] = 0; self.force_duration[pending] = 0; self.force_magnitude[pending] = 0
            self.force_direction[pending] = 0; self.force_body[pending] = 0; self.xfrc[pending] = 0
            self.previous_reward[pending] = 0.0
            # first step with action=None; reject starts that touch anything or end the episode
            mask = np.zeros(self.num_envs, dtype=bool); mask[pending] = True
            obs, _, term, trunc, _ = self._step_impl(None, only=mask)
            if not mask.all():  # the lock-step call advanced everyone: put the bystanders back
                q2, v2, t2, w2 = self.batch.get_state()
                keep = ~mask
                q2[keep], v2[keep], t2[keep], w2[keep] = qpos[keep], qvel[keep], time[keep], ws[keep]
                self.batch.set_state(q2, v2, t2, w2)
                self.time = t2
            ncon = self.batch.get("ncon")
            bad = mask & ((ncon > 0) | term | trunc)
            accepted = mask & ~bad
            if accepted.any():
                obs2 = self._observe(only=accepted)  # reset returns a SECOND _get_obs() (cpu_env.py:416)
                obs_out[accepted] = obs2[accepted]
            pending = np.nonzero(bad)[0]
        if len(pending):
            raise RuntimeError(f"reset: {len(pending)} environments still in contact after 100 attempts")
        return (obs_out if env_ids is None else obs_out[idx]), {}

    # ------------------------------------------------------------------ step (cpu_env.py:676-693)
    def step(self, action=None):
        obs, reward, term, trunc, _ = self._step_impl(action, only=None)
        infos = [({"is_success": bool(trunc[i])} if (term[i] or trunc[i]) else {}) for i in range(self.num_envs)]
        return obs, reward, term, trunc, infos

    def _apply_action(self, action, only):
        n = self.num_envs
        sel = np.ones(n, dtype=bool) if only is None else only
        if action is None:
            a = self.ctrl.copy()
        else:
            a = np.asarray(action, dtype=np.float64).reshape(n, NU) + self.noise_action
        ids = np.nonzero(sel)[0]
        slot = self.action_count[ids] % (MAX_DELAY_STEPS + 1)
        self.action_hist[ids, slot] = a[ids]
        d = self.action_delay[ids]
        src = (self.action_count[ids] - d) % (MAX_DELAY_STEPS + 1)
        take = np.where((self.action_count[ids] >= d)[:, None], self.action_hist[ids, src], 0.0)
        self.action_count[ids] += 1
        self.ctrl[ids] = take
        self.previous_action[ids] = self.latest_action[ids]
        self.latest_action[ids] = take

    def _apply_external_forces(self, only):
        rf, rng = self.randomization_factor, self.rng
        sel = np.ones(self.num_envs, dtype=bool) if only is None else only
        t = self.time
        renew = sel & (t >= self.force_start + self.force_duration)
        k = int(renew.sum())
        if k:
            ids = np.nonzero(renew)[0]
            self.force_start[ids] = t[ids] + rng.uniform(MIN_EXTERNAL_FORCE_INTERVAL, MAX_EXTERNAL_FORCE_INTERVAL, k)
            self.force_duration[ids] = rng.uniform(MIN_EXTERNAL_FORCE_DURATION, MAX_EXTERNAL_FORCE_DURATION, k)
            self.force_magnitude[ids] = rng.uniform(MIN_EXTERNAL_FORCE_MAGNITUDE * rf, MAX_EXTERNAL_FORCE_MAGNITUDE * rf, k)
            dirs = rng.uniform(-1, 1, (k, 2))
            nrm = np.linalg.norm(dirs, axis=1, keepdims=True)
            dirs = np.where(nrm == 0, np.array([[1.0, 0.0]]), dirs / np.maximum(nrm, 1e-300))
            self.force_direction[ids] = dirs
            self.xfrc[ids, 0:2] = 0.0
            self.force_body[ids] = rng.integers(1, 15, k)
        active = sel & (t > self.force_start) & (t < self.force_start + self.force_duration)
        ids = np.nonzero(active)[0]
        self.xfrc[ids, 0:2] = self.force_direction[ids] * self.force_magnitude[ids, None]

    def _observe(self, only=None, res=None):
        """_get_obs + _get_reward on the device; `res` = results of a fused hmj_env_step, else a separate epilogue call."""
        if res is None:
            obs_now, reward, term, trunc, selfc = self.batch.obs_reward(
                self.previous_action, self.latest_action, self.noise_obs, self.max_simulation_time)
            root_quat = self.batch.get("sens")[:, 3:7]
        else:
            obs_now, reward, root_quat = res["obs"], res["reward"], res["root_quat"]
            term, trunc, selfc = res["terminated"].astype(bool), res["truncated"].astype(bool), res["self_collision"].astype(bool)
        sel = np.ones(self.num_envs, dtype=bool) if only is None else only
        ids = np.nonzero(sel)[0]
        self.torso_quat[ids] = root_quat[ids] + self.noise_obs[ids, 27:31]
        # delay FIFOs (joint angles+velocities share a delay; gyro; gravity), cpu_env.py:135-187,528-547
        slot = self.obs_count[ids] % (MAX_DELAY_STEPS + 1)
        self.obs_hist[ids, slot] = obs_now[ids]
        out = obs_now.copy()
        init = np.zeros(30); init[27:30] = [0, 0, -1]
        for grp, cols in ((0, slice(0, 24)), (1, slice(24, 27)), (2, slice(27, 30))):
            d = self.obs_delay[ids, grp]
            src = (self.obs_count[ids] - d) % (MAX_DELAY_STEPS + 1)
            val = self.obs_hist[ids, src][:, cols]
            out[ids, cols] = np.where((self.obs_count[ids] >= d)[:, None], val, init[cols][None, :])
        self.obs_count[ids] += 1
        self._last = (reward, term, trunc, selfc)
        return out

    def _step_impl(self, action, only):
        self._apply_action(action, only)
        self._apply_external_forces(only)
        # ONE C call: ctrl/xfrc H2D, mj_step (PHYSICS_STEPS_PER_CONTROL_STEP = 1, simulation_parameters.py:82),
        # obs/reward epilogue, results D2H
        res = self.batch.env_step(self.ctrl, self.force_body, self.xfrc, self.previous_action, self.latest_action,
                                  self.noise_obs, self.max_simulation_time, 1, out=getattr(self, "_out", None))
        self.time = res["time"].copy()
        obs = self._observe(only, res)
        reward, term, trunc, selfc = self._last
        if self.reward_override is not None:
            reward = np.full(self.num_envs, float(self.reward_override))
        elif self.use_potential_rewards:
            r = reward - self.previous_reward
            self.previous_reward = reward.copy()
            reward = r
        return obs, reward, term, trunc, selfc

    def close(self):
        self.batch.close()
