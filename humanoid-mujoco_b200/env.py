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
seeded numpy Generator at reset and by a counter-based splitmix64 stream for the pushes (those streams
are unpinned in the reference: unseeded `random`, JAX absent here); Rotation math is fp64 instead of
JAX's default fp32.

Where the work happens: `step` is ONE C call (hmj_env_step_fused) — actions are written into a pinned host
block, the action-noise / delay FIFO / push scheduler prologue, the physics step, the observation / reward
epilogue and the observation delay FIFOs all run as CUDA kernels on per-env state that lives on the device,
and the results come back in the same pinned block.  `step_host` is the numpy restatement of that
bookkeeping (the reference's own control flow, vectorised) kept as the checker for the device kernels.
"""
from __future__ import annotations

from collections.abc import Sequence

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

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = x
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def rand_u01(seed: int, env: np.ndarray, counter: np.ndarray) -> np.ndarray:
    """hmj_rand_u01 (csrc/hmj_dev.cuh): counter-based uniform in [0, 1), identical bits on host and device."""
    key = (np.asarray(env, dtype=np.uint64) << np.uint64(32)) | np.asarray(counter, dtype=np.uint64)
    h = _splitmix64(np.uint64(seed & 0xFFFFFFFFFFFFFFFF) ^ _splitmix64(key))
    return (h >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


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


# draw offsets of k_env_reset (csrc/hmj_reset.cu), one block of 512 counters per generation of a pool row
_RD = dict(FRICTION=0, MASS=1, EXTRA_BODY=15, EXTRA_MASS=16, ARMATURE=17, MARGIN=29, RANGE=41, GAIN=65, FORCERANGE=77,
           HF_COARSE=101, HF_FINE=117, DELAY=181, N_ACT=187, N_SAME=211, N_GYRO=235, N_IMU=241, JOINT=249, Z=261, QUAT=262)


def device_reset_draws(model: np.ndarray, base: np.ndarray, seed: int, rows, gens, rf: float, timestep: float = TIMESTEP):
    """numpy restatement of k_env_reset (csrc/hmj_reset.cu), the checker of the device-side CPUEnv.reset draws
    (cpu_env.py:135-328): the episode that pool row `rows[i]` (GLOBAL row index: N + spare row) holds in its generation
    `gens[i]`.  Same counter-based stream, same IEEE operation order (the kernel uses __dmul_rn / __dadd_rn).  Returns
    (params hmj_params_t[k], qpos [k][19], act_delay [k], obs_delay [k][3], act_noise [k][12], obs_noise [k][31])."""
    rows = np.asarray(rows, dtype=np.uint64)
    gens = np.asarray(gens, dtype=np.uint64)
    k = rows.shape[0]
    key = int(_splitmix64(np.uint64(seed & 0xFFFFFFFFFFFFFFFF) ^ np.uint64(0x5EED5EEDC0FFEE11)))

    def u(off, count=1):
        c = (gens * np.uint64(512))[:, None] + np.uint64(off) + np.arange(count, dtype=np.uint64)[None, :]
        return rand_u01(key, np.broadcast_to(rows[:, None], c.shape), c)

    def lerp(lo, hi, x):
        return lo + (hi - lo) * x

    def normal(off, count):
        uu = u(off, 2 * count)
        return np.sqrt(-2.0 * np.log(1.0 - uu[:, 0::2])) * np.cos(2.0 * np.pi * uu[:, 1::2])

    p = np.repeat(base[:1], k).copy()
    p["timestep"] = timestep
    a = lerp(0.5, 1.0, u(_RD["FRICTION"]))
    mult = (1.0 - rf) + a * rf
    p["floor_friction"] = np.asarray(model["geom_friction"][0][0])[None, :] * mult
    mass = np.tile(np.asarray(model["body_mass"][0]), (k, 1))
    mass[:, 1:] = np.maximum(1e-5, mass[:, 1:] + lerp(-0.05 * rf, 0.05 * rf, u(_RD["MASS"], 14)))
    eb = np.minimum(1 + (14.0 * u(_RD["EXTRA_BODY"])[:, 0]).astype(np.int64), 14)
    mass[np.arange(k), eb] += (0.2 * rf) * u(_RD["EXTRA_MASS"])[:, 0]
    p["body_mass"] = mass
    arm = np.tile(np.asarray(model["dof_armature"][0]), (k, 1))
    arm[:, 6:] += (0.0005 * u(_RD["ARMATURE"], 12)) * rf
    p["dof_armature"] = arm
    p["jnt_stiffness"] = np.asarray(model["jnt_stiffness"][0])[None, :]
    margin = np.tile(np.asarray(model["jnt_margin"][0]), (k, 1))
    margin[:, 1:] += (0.05 * u(_RD["MARGIN"], 12)) * rf
    p["jnt_margin"] = margin
    rng_ = np.tile(np.asarray(model["jnt_range"][0]), (k, 1, 1))
    rng_[:, 1:, :] += (lerp(-0.1, 0.1, u(_RD["RANGE"], 24)) * rf).reshape(k, 12, 2)
    p["jnt_range"] = rng_
    p["act_gainprm0"] = 2.0 + lerp(-0.5, 0.5, u(_RD["GAIN"], 12)) * rf
    p["act_forcerange"] = np.asarray(model["act_forcerange"][0])[None] + (lerp(-0.05, 0.05, u(_RD["FORCERANGE"], 24)) * rf).reshape(k, 12, 2)
    coarse = u(_RD["HF_COARSE"], 16).reshape(k, 4, 4)
    up = np.repeat(np.repeat(coarse, 2, axis=1), 2, axis=2)
    pad = np.pad(up, ((0, 0), (1, 1), (1, 1)), mode="edge")
    kk = np.array([0.25, 0.5, 0.25])
    sm = np.zeros((k, 8, 8))
    for i in range(3):
        for j in range(3):
            sm = sm + (kk[i] * kk[j]) * pad[:, i:i + 8, j:j + 8]
    sm = sm + 0.05 * u(_RD["HF_FINE"], 64).reshape(k, 8, 8)
    mn = sm.min(axis=(1, 2), keepdims=True)
    span = np.maximum(sm.max(axis=(1, 2), keepdims=True) - mn, 1e-12)
    p["hfield_data"] = ((sm - mn) / span).astype(np.float32).reshape(k, 64) * np.float32(0.1 * rf)
    dl = np.rint(lerp(0.01 * rf, 0.05 * rf, u(_RD["DELAY"], 6)) / timestep).astype(np.int32)
    act_delay = dl[:, 0]
    obs_delay = np.stack([dl[:, 1], dl[:, 2], dl[:, 4]], axis=1)
    d2r = np.pi / 180.0
    act_noise = rf * (0.5 * d2r) * normal(_RD["N_ACT"], 12)
    same = normal(_RD["N_SAME"], 12)
    obs_noise = np.concatenate([rf * (2.0 * d2r) * same, rf * (5.0 * d2r) * same, rf * (2.0 * d2r) * normal(_RD["N_GYRO"], 3),
                                rf * (5.0 * d2r) * normal(_RD["N_IMU"], 4)], axis=1)
    qpos = np.tile(np.asarray(model["qpos0"][0], dtype=np.float64), (k, 1))
    qpos[:, 7:] += lerp(-0.2 * rf, 0.2 * rf, u(_RD["JOINT"], 12))
    qpos[:, 0] = 0.0
    qpos[:, 1] = 0.0
    qpos[:, 2] = -0.6 + (0.1 * rf) * u(_RD["Z"])[:, 0]
    qpos[:, 3:7] = np.array(INITIAL_QUAT_STANDUP) + lerp(-0.1 * rf, 0.1 * rf, u(_RD["QUAT"], 4))
    return p, qpos, act_delay, obs_delay, act_noise, obs_noise


def local_gravity_host(quat: np.ndarray) -> np.ndarray:
    """inverseRotateVectors(q, (0, 0, -1)) (cpu_env.py:26-28) for rows of wxyz quaternions: Rotation.from_quat
    normalises, so this is minus the third row of R(q / |q|).  numpy mirror of `local_gravity` in csrc/hmj_aux.cu."""
    q = np.asarray(quat, dtype=np.float64)
    q = q / np.linalg.norm(q, axis=-1, keepdims=True)
    w, x, y, z = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    return -np.stack([2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z], axis=-1)


def observation_host(qpos, qvel, root_quat, noise, act_qposadr, act_dofadr) -> np.ndarray:
    """The undelayed observation of CPUEnv._get_obs (cpu_env.py:465-571) for a batch: noisy joint angles and
    velocities in JOINT_NAMES (= actuator) order, noisy world-frame angular velocity, local gravity of the noisy
    quaternion.  numpy mirror of the observation half of k_env_obs_reward."""
    nz = np.zeros((qpos.shape[0], 31)) if noise is None else np.asarray(noise)
    return np.concatenate([qpos[:, act_qposadr] + nz[:, 0:12], qvel[:, act_dofadr] + nz[:, 12:24],
                           qvel[:, 3:6] + nz[:, 24:27], local_gravity_host(root_quat + nz[:, 27:31])], axis=1)


def standup_reward_host(qvel, root_quat, root_z, joint_torques, prev_action, latest_action, self_collision, time,
                        max_sim_time):
    """standupReward (reward_functions.py:247-374) for a batch, weights and joint pairs as selected there.
    `prev_action` / `latest_action` are the env's raw actions (the /(pi/2) of cpu_env.py:601-602 happens here).
    Returns (reward, terminated, truncated).  numpy mirror of the reward half of k_env_obs_reward; pinned against the
    reference's own function by tests/golden/env_golden.npz."""
    se = lambda x: np.exp(-x / 0.5)  # noqa: E731  scaled_exp
    g = local_gravity_host(root_quat)
    lat = np.asarray(latest_action) / (np.pi / 2)
    prv = np.asarray(prev_action) / (np.pi / 2)
    r = 5.0 * se(qvel[:, 0] ** 2 + qvel[:, 1] ** 2)
    r = r + 10.0 * se(g[:, 0] ** 2 + g[:, 1] ** 2 + (g[:, 2] + 1.0) ** 2)
    r = r + np.interp(root_z, [-0.6, -0.375], [0.0, 15.0])
    r = r + 2.5 * se(np.clip(np.abs(joint_torques) - 1.0, 0, np.inf) ** 2).sum(axis=1) / 12.0
    r = r + 2.0 * se(((lat - prv) * 500.0) ** 2).sum(axis=1) / 12.0
    r = r + 0.5 * se(lat ** 2).sum(axis=1) / 12.0
    # JOINT_NAMES order: R sh pitch, R sh roll, R elbow, L sh pitch, L sh roll, L elbow, L hip roll, L hip pitch, L knee,
    # R hip roll, R hip pitch, R knee; elbows equal, the other pairs opposite
    sym = se((lat[:, 5] - lat[:, 2]) ** 2)
    for i, j in ((6, 9), (7, 10), (8, 11), (3, 0), (4, 1)):
        sym = sym + se((lat[:, i] + lat[:, j]) ** 2)
    r = r + sym / 6.0
    r = r + np.where(self_collision, -20.0, 0.0)
    term = np.asarray(time) >= max_sim_time
    r = np.where(term, -100.0, r)
    trunc = (np.asarray(root_z) >= -0.375) & (np.max(np.abs(g[:, :2]), axis=1) < 0.7)
    return r, term, trunc


class SparseInfos(Sequence):
    """The per-env `info` dicts of one batched step as a read-only sequence of length N: only the envs whose episode
    ended carry an entry (`is_success`), built when somebody looks at it; everyone else shares one empty dict (building
    N dicts per step would cost more host time than the device step saves at 4096 envs)."""

    __slots__ = ("_n", "_ids", "_success", "_entries")
    _EMPTY: dict = {}

    def __init__(self, n, entries=None, ids=None, success=None):
        self._n = int(n)
        self._entries = entries
        self._ids, self._success = ids, success

    def _build(self):
        if self._entries is None:
            self._entries = {int(i): {"is_success": bool(s)} for i, s in zip(self._ids, self._success)}
        return self._entries

    def __len__(self):
        return self._n

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(self._n))]
        i = int(i)
        if i < 0:
            i += self._n
        if not 0 <= i < self._n:
            raise IndexError(i)
        return self._build().get(i, self._EMPTY)


class BatchedHumanoidEnv:
    metadata = {"render_modes": [], "render_fps": 30}

    # per-env bookkeeping that lives on the device (csrc/hmj_dev.cuh: hmj_envstate_t) <-> host mirror attribute
    _DEVICE_STATE = (("env_act_noise", "noise_action"), ("env_obs_noise", "noise_obs"), ("env_act_delay", "action_delay"),
                     ("env_obs_delay", "obs_delay"), ("env_act_count", "action_count"), ("env_obs_count", "obs_count"),
                     ("env_prev_action", "previous_action"), ("env_latest_action", "latest_action"),
                     ("env_force_sched", "force_sched"), ("env_force_body", "force_body"),
                     ("env_force_draws", "force_draws"), ("ctrl", "ctrl"), ("xfrc", "xfrc"))

    def __init__(self, n_envs: int, randomization_factor: float = 0.0, use_potential_rewards: bool = False,
                 max_simulation_time_override=None, reward_override=None, device: int = 0, seed: int = 0):
        self.num_envs = int(n_envs)
        self.randomization_factor = float(randomization_factor)
        self.use_potential_rewards = bool(use_potential_rewards)
        self.reward_override = reward_override
        self.timestep = TIMESTEP
        mst = MAX_SIM_TIME_STANDUP if max_simulation_time_override is None else max_simulation_time_override
        self.max_simulation_time = np.inf if mst < 0 else float(mst)
        self.seed = int(seed)
        self.rng = np.random.default_rng(seed)
        self.batch = HumanoidBatch(self.num_envs, device)
        self.device_reset = self.batch.spare_rows > 0   # CPUEnv.reset on the device (episode pool); False: numpy path
        self.auto_reset = False                         # DummyVecEnv semantics inside step() (B200VecEnv switches it on)
        self._pool_cfg = None
        self.last_terminal = (np.zeros(0, dtype=np.int64), np.zeros((0, 30)))
        self.io = self.batch.env_io()
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
        self.action_delay = np.zeros(n, dtype=np.int32)
        self.obs_delay = np.zeros((n, 3), dtype=np.int32)  # joint, gyro, gravity
        self.action_hist = np.zeros((n, MAX_DELAY_STEPS + 1, NU))
        self.obs_hist = np.zeros((n, MAX_DELAY_STEPS + 1, 30))
        self.action_count = np.zeros(n, dtype=np.int32)
        self.obs_count = np.zeros(n, dtype=np.int32)
        self.previous_reward = np.zeros(n)
        self.force_sched = np.zeros((n, 5))  # start, duration, magnitude, dir x, dir y
        self.force_body = np.zeros(n, dtype=np.int32)
        self.force_draws = np.zeros(n, dtype=np.int32)
        self.xfrc = np.zeros((n, 6))
        self.torso_quat = np.zeros((n, 4))
        self.time = np.zeros(n)
        self._empty_info = {}
        self._no_infos = SparseInfos(self.num_envs, {})
        self._all_false = np.zeros(self.num_envs, dtype=bool)
        self._all_false.flags.writeable = False

    # ------------------------------------------------------------------ host mirror <-> device state
    def _push_state(self):
        for dev, host in self._DEVICE_STATE:
            self.batch.set(dev, getattr(self, host))

    def _pull_state(self):
        """Bring the host mirrors up to date (before a partial reset, or to continue with `step_host`)."""
        for dev, host in self._DEVICE_STATE:
            setattr(self, host, self.batch.get(dev))
        n = self.num_envs
        self.action_hist = self.batch.get_raw("env_act_hist", np.float64, (n, MAX_DELAY_STEPS + 1, NU))
        self.obs_hist = self.batch.get_raw("env_obs_hist", np.float64, (n, MAX_DELAY_STEPS + 1, 30))
        self.time = self.batch.get("time")

    # ------------------------------------------------------------------ randomisers (cpu_env.py:135-362)
    def _randomize_model(self, idx):
        self.params[idx] = randomize_params(self.base_params[idx], self.rng, self.randomization_factor, self.timestep)

    def _randomize_delays(self, idx):
        rf, rng, k = self.randomization_factor, self.rng, len(idx)
        draw = lambda: np.rint(rng.uniform(MIN_DELAY * rf, MAX_DELAY * rf, k) / self.timestep).astype(np.int32)  # noqa: E731
        self.action_delay[idx] = draw()
        self.obs_delay[idx, 0] = draw()  # joint angles + velocities
        self.obs_delay[idx, 1] = draw()  # gyro
        draw()                           # torso_local_velocity_delay (drawn, unused by the observation)
        self.obs_delay[idx, 2] = draw()  # gravity vector
        draw()                           # contact_sensor_delay (unused)
        # the FIFOs need no clearing: a slot is always written before the first read that can reach it
        self.action_count[idx] = 0
        self.obs_count[idx] = 0

    def _start_state(self, idx):
        return standup_start_state(self.model["qpos0"][0], self.rng, self.randomization_factor, len(idx))

    # ------------------------------------------------------------------ reset (cpu_env.py:374-416)
    def _pool_settings(self):
        return (bool(self.auto_reset), self.randomization_factor, self.max_simulation_time, self.seed, self.timestep)

    def _configure_pool(self):
        cur = self._pool_settings()
        if cur != self._pool_cfg:
            self.batch.env_pool_config(*cur)
            self._pool_cfg = cur

    def reset(self, seed=None, options=None, env_ids=None):
        """CPUEnv.reset for all envs (or `env_ids`).  With `device_reset` (default) the domain randomisation, delays, noise
        vectors, start pose, probe step and the rejection loop run on the device (episode pool, csrc/hmj_reset.cu): no
        bystander is touched and nothing but the new observations crosses the bus.  `reset_host` is the numpy path with
        full state round trips, kept as the checker."""
        if not self.device_reset:
            return self.reset_host(seed, options, env_ids)
        if seed is not None:
            self.seed = int(seed)
            self.rng = np.random.default_rng(seed)
        self._configure_pool()
        ids = None if env_ids is None else np.asarray(env_ids, dtype=np.int32)
        self.batch.env_reset(ids)
        io = self.io
        self.time = io["time"]
        self.torso_quat = io["torso_quat"]
        if ids is None:
            self.previous_reward[:] = 0.0
            return io["obs"].copy(), {}
        self.previous_reward[ids] = 0.0
        return io["obs"][ids].copy(), {}

    def reset_host(self, seed=None, options=None, env_ids=None):
        if seed is not None:
            self.seed = int(seed)
            self.rng = np.random.default_rng(seed)
        idx = np.arange(self.num_envs) if env_ids is None else np.asarray(env_ids, dtype=np.int64)
        if env_ids is not None:
            self._pull_state()
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
            self.force_sched[pending] = 0.0; self.force_body[pending] = 0; self.xfrc[pending] = 0.0
            self.previous_reward[pending] = 0.0
            self._push_state()
            # first step with action=None; reject starts that touch anything or end the episode
            mask = np.zeros(self.num_envs, dtype=bool); mask[pending] = True
            _obs, _, term, trunc, _ = self._step_device(None, only=mask)
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
                obs2 = self._observe_device(accepted)  # reset returns a SECOND _get_obs() (cpu_env.py:416)
                obs_out[accepted] = obs2[accepted]
            pending = np.nonzero(bad)[0]
            if len(pending):
                self._pull_state()
        if len(pending):
            raise RuntimeError(f"reset: {len(pending)} environments still in contact after 100 attempts")
        return (obs_out if env_ids is None else obs_out[idx]), {}

    def reset_to(self, params: np.ndarray, qpos: np.ndarray, qvel=None):
        """Start every episode from GIVEN model parameters and poses (a workload built elsewhere, e.g. bench.py): the
        per-episode draws that do not touch the physics (delays, constant noise vectors) and the trackers are
        initialised as in `reset`; no DR of the model, no first-step probe."""
        n, rf = self.num_envs, self.randomization_factor
        ids = np.arange(n)
        self.params[:] = params
        self.batch.set_params(self.params)
        self._randomize_delays(ids)
        self.noise_action[:] = rf * np.deg2rad(JOINT_ACTION_NOISE_STDDEV) * self.rng.standard_normal((n, NU))
        same12 = self.rng.standard_normal((n, NU))
        self.noise_obs[:, 0:12] = rf * np.deg2rad(JOINT_ANGLE_NOISE_STDDEV) * same12
        self.noise_obs[:, 12:24] = rf * np.deg2rad(JOINT_VELOCITY_NOISE_STDDEV) * same12
        self.noise_obs[:, 24:27] = rf * np.deg2rad(GYRO_NOISE_STDDEV) * self.rng.standard_normal((n, 3))
        self.noise_obs[:, 27:31] = rf * np.deg2rad(IMU_NOISE_STDDEV) * self.rng.standard_normal((n, 4))
        self.batch.set_state(qpos, np.zeros((n, NV)) if qvel is None else qvel, np.zeros(n), np.zeros((n, NV)))
        self.ctrl[:] = 0.0
        self.previous_action[:] = 0.0; self.latest_action[:] = 0.0
        self.force_sched[:] = 0.0; self.force_body[:] = 0; self.xfrc[:] = 0.0
        self.previous_reward[:] = 0.0
        self._push_state()

    # ------------------------------------------------------------------ step (cpu_env.py:676-693)
    def step(self, action=None):
        """One control step of every environment.  Returns views of the pinned result block (valid until the next
        call): obs [N][30], reward [N], terminated [N], truncated [N], infos (a sequence of N dicts, `is_success` on episode end: SparseInfos)."""
        if self.device_reset:
            self._configure_pool()
        obs, reward, term, trunc, _ = self._step_device(action, only=None)
        if term is self._all_false and trunc is self._all_false:
            return obs, reward, term, trunc, self._no_infos      # nothing ended: shared (read-only) flags and infos
        ids = np.nonzero(term | trunc)[0]
        infos = SparseInfos(self.num_envs, ids=ids, success=trunc[ids])
        if self.auto_reset and self.device_reset:
            # the ended envs already run their next episode (obs rows = its first observation); the terminal
            # observations come back in the pool's result block (first 64) or are fetched (a mass ending)
            tid, tobs = self.batch.env_terminal()
            self.last_terminal = (tid, tobs)
            if self.use_potential_rewards:
                self.previous_reward[ids] = 0.0
        return obs, reward, term, trunc, infos

    def _step_device(self, action, only):
        io = self.io
        act = None
        if action is not None:
            # copied to the device from where it lies (no host staging copy); pinned arrays (hmj.pinned_empty) go by DMA
            act = np.ascontiguousarray(action, dtype=np.float64)
            if act.size != self.num_envs * NU:
                raise ValueError(f"action must hold {self.num_envs} x {NU} values")
        if only is not None:
            io["mask"][...] = only
        # ONE C call: action H2D, prologue, mj_step (PHYSICS_STEPS_PER_CONTROL_STEP = 1, simulation_parameters.py:82),
        # obs/reward epilogue + delay FIFOs, results D2H
        self.batch.env_step_fused(act, only is not None, self.randomization_factor, self.max_simulation_time, self.seed, 1)
        self.time = io["time"]
        self.torso_quat = io["torso_quat"]
        if not io["flag_bytes"].any():    # almost every step: no flag of any kind raised (one scan instead of three)
            af = self._all_false
            return self._finish(io["obs"], io["reward"], af, af, af, only)
        return self._finish(io["obs"], io["reward"], self._flags(io["terminated"]), self._flags(io["truncated"]),
                            self._flags(io["self_collision"]), only)

    def _flags(self, raw):
        """int32 device flags -> bool array; the all-clear case (almost every step) hands out one shared read-only array."""
        return raw.astype(bool) if raw.any() else self._all_false

    def _finish(self, obs, reward, term, trunc, selfc, only=None):
        """reward_override / potential rewards (cpu_env.py:606-614).  `only`: the envs this step belongs to (reset's probe
        step of a subset) — the potential-reward memory of the others must not move."""
        if self.reward_override is not None:
            reward = np.full(self.num_envs, float(self.reward_override))
        elif self.use_potential_rewards:
            r = reward - self.previous_reward
            if only is None:
                self.previous_reward = reward.copy()
            else:
                self.previous_reward = np.where(only, reward, self.previous_reward)
            reward = r
        return obs, reward, term, trunc, selfc

    def _observe_device(self, only):
        """A second _get_obs() on the current state (reset's return value): epilogue + delay FIFOs, no physics."""
        io = self.io
        io["mask"][...] = only
        self.batch.env_step_fused(False, True, self.randomization_factor, self.max_simulation_time, self.seed, -1)
        return io["obs"].copy()

    # ------------------------------------------------------------------ numpy restatement of the device bookkeeping
    # (the reference's control flow, vectorised; the checker for k_env_pre / k_env_post — tests/test_gpu_env.py)
    def step_host(self, action=None):
        """Same contract as `step`, bookkeeping in numpy on the host mirrors (call `_pull_state()` first when the
        device path ran before).  The device bookkeeping arrays are NOT updated by this path."""
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
        rf = self.randomization_factor
        sel = np.ones(self.num_envs, dtype=bool) if only is None else only
        t = self.time
        fs = self.force_sched
        renew = sel & (t >= fs[:, 0] + fs[:, 1])
        if renew.any():
            ids = np.nonzero(renew)[0]
            c = self.force_draws[ids].astype(np.uint64) * np.uint64(8)
            u = [rand_u01(self.seed, ids, c + np.uint64(j)) for j in range(6)]
            fs[ids, 0] = t[ids] + (MIN_EXTERNAL_FORCE_INTERVAL + (MAX_EXTERNAL_FORCE_INTERVAL - MIN_EXTERNAL_FORCE_INTERVAL) * u[0])
            fs[ids, 1] = MIN_EXTERNAL_FORCE_DURATION + 0.10 * u[1]
            fs[ids, 2] = 5.0 * rf + (15.0 * rf - 5.0 * rf) * u[2]
            dx, dy = -1.0 + 2.0 * u[3], -1.0 + 2.0 * u[4]
            nrm = np.sqrt(dx * dx + dy * dy)
            zero = nrm == 0
            nrm = np.where(zero, 1.0, nrm)
            fs[ids, 3] = np.where(zero, 1.0, dx / nrm)
            fs[ids, 4] = np.where(zero, 0.0, dy / nrm)
            self.xfrc[ids, 0:2] = 0.0
            self.force_body[ids] = np.minimum(1 + (14.0 * u[5]).astype(np.int32), 14)
            self.force_draws[ids] += 1
        active = sel & (t > fs[:, 0]) & (t < fs[:, 0] + fs[:, 1])
        ids = np.nonzero(active)[0]
        self.xfrc[ids, 0:2] = fs[ids, 3:5] * fs[ids, 2:3]

    def _observe(self, only=None, res=None):
        """_get_obs + _get_reward; `res` = results of hmj_env_step (undelayed observation from the epilogue kernel)."""
        obs_now, reward, root_quat = res["obs"], res["reward"], res["root_quat"]
        term, trunc, selfc = res["terminated"].astype(bool), res["truncated"].astype(bool), res["self_collision"].astype(bool)
        sel = np.ones(self.num_envs, dtype=bool) if only is None else only
        ids = np.nonzero(sel)[0]
        self.torso_quat = np.array(self.torso_quat)
        self.torso_quat[ids] = root_quat[ids] + self.noise_obs[ids, 27:31]
        # delay FIFOs (joint angles+velocities share a delay; gyro; gravity), cpu_env.py:135-187,528-547
        slot = self.obs_count[ids] % (MAX_DELAY_STEPS + 1)
        self.obs_hist[ids, slot] = obs_now[ids]
        out = obs_now.copy()
        init = np.zeros(30); init[29] = -1.0
        for grp, cols in ((0, slice(0, 24)), (1, slice(24, 27)), (2, slice(27, 30))):
            d = self.obs_delay[ids, grp]
            src = (self.obs_count[ids] - d) % (MAX_DELAY_STEPS + 1)
            val = self.obs_hist[ids, src][:, cols]
            out[ids, cols] = np.where((self.obs_count[ids] >= d)[:, None], val, init[cols][None, :])
        self.obs_count[ids] += 1
        return out, reward, term, trunc, selfc

    def _step_impl(self, action, only):
        self._apply_action(action, only)
        self._apply_external_forces(only)
        res = self.batch.env_step(self.ctrl, self.force_body, self.xfrc, self.previous_action, self.latest_action,
                                  self.noise_obs, self.max_simulation_time, 1)
        self.time = res["time"].copy()
        obs, reward, term, trunc, selfc = self._observe(only, res)
        return self._finish(obs, reward, term, trunc, selfc, only)

    def close(self):
        self.batch.close()
