"""HumanoidBatch — thin object wrapper over the hmj_batch_* C ABI (include/hmj_b200.h).

Plays the role of (MjModel, MjData) x N in the reference (simulation/cpu_env.py:86,397): state in,
`step`/`forward`, state and derived quantities out.  numpy in / numpy out; device pointers are
exposed for zero-copy torch interop.
"""
from __future__ import annotations

import ctypes as C
import os
import re

import numpy as np

from ._lib import check, load_library

_HERE = os.path.dirname(os.path.abspath(__file__))
_HEADER = os.path.join(os.path.dirname(_HERE), "include", "hmj_model.h")
MODEL_PATH = os.path.join(_HERE, "assets", "humanoid_model.bin")

_CTYPES = {"int32_t": np.int32, "double": np.float64, "float": np.float32}


def _parse_header(path):
    """numpy dtypes of the POD structs in include/hmj_model.h (C alignment)."""
    src = re.sub(r"/\*.*?\*/", "", open(path).read(), flags=re.S)
    defines, types, structs = {}, dict(_CTYPES), {}
    for m in re.finditer(r"#define\s+(HMJ_\w+)\s+(.+)", src):
        try:
            defines[m.group(1)] = int(eval(m.group(2).strip(), {}, dict(defines)))  # noqa: S307 (own header)
        except Exception:
            pass
    for sm in re.finditer(r"typedef struct \w+ \{(.*?)\} (\w+);", src, re.S):
        fields = []
        for stmt in sm.group(1).split(";"):
            stmt = stmt.strip()
            if not stmt:
                continue
            m = re.match(r"(\w+)\s+(.*)", stmt, re.S)
            for d in m.group(2).split(","):
                dm = re.match(r"(\w+)((?:\[[^\]]+\])*)$", d.strip())
                dims = tuple(int(eval(x, {}, dict(defines))) for x in re.findall(r"\[([^\]]+)\]", dm.group(2)))  # noqa: S307
                fields.append((dm.group(1), types[m.group(1)], dims) if dims else (dm.group(1), types[m.group(1)]))
        structs[sm.group(2)] = types[sm.group(2)] = np.dtype(fields, align=True)
    return structs, defines


_STRUCTS, DEFINES = _parse_header(_HEADER)
MODEL_DTYPE = _STRUCTS["hmj_model_t"]
PARAMS_DTYPE = _STRUCTS["hmj_params_t"]
CONTACT_DTYPE = _STRUCTS["hmj_contact_t"]

NQ, NV, NU = DEFINES["HMJ_NQ"], DEFINES["HMJ_NV"], DEFINES["HMJ_NU"]
MAXCON = DEFINES["HMJ_MAXCON"]

_ARRAY_SHAPES = {
    "qpos": (NQ,), "qvel": (NV,), "time": (), "warmstart": (NV,), "ctrl": (NU,), "xfrc": (6,),
    "geom_xpos": (DEFINES["HMJ_NGEOM"], 3), "geom_xmat": (DEFINES["HMJ_NGEOM"], 9), "cdof": (NV, 6), "com": (3,),
    "qM": (DEFINES["HMJ_NM"],), "qfrc_smooth": (NV,), "qacc_smooth": (NV,), "qacc": (NV,), "qfrc_constraint": (NV,),
    "sens": (16,), "residual": (DEFINES["HMJ_NRESIDUAL"],), "cost_sum": (), "cost_last": (), "trace": (3,), "returns": (),
}
_ARRAY_SHAPES.update({
    "env_act_noise": (NU,), "env_obs_noise": (31,), "env_prev_action": (NU,), "env_latest_action": (NU,),
    "env_force_sched": (5,), "env_torso_quat": (4,),
})
_INT_ARRAYS = {"ncon", "nefc", "niter", "status", "xfrc_body", "env_act_delay", "env_act_count", "env_obs_count",
               "env_force_body", "env_force_draws"}
_INT_SHAPES = {"env_obs_delay": (3,)}


def load_model(path: str = MODEL_PATH) -> np.ndarray:
    raw = np.fromfile(path, dtype=np.uint8)
    if raw.size != MODEL_DTYPE.itemsize:
        raise ValueError(f"model blob is {raw.size} bytes, expected sizeof(hmj_model_t) = {MODEL_DTYPE.itemsize}")
    m = raw.view(MODEL_DTYPE)
    if int(m["magic"][0]) != DEFINES["HMJ_MAGIC"] or int(m["version"][0]) != DEFINES["HMJ_VERSION"]:
        raise ValueError("model blob magic/version mismatch: re-run tools/compile_model.py")
    return m.copy()


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def default_params(model: np.ndarray, n: int = 1) -> np.ndarray:
    """hmj_params_t x n initialised from the model (the un-randomised MjModel fields)."""
    lib = load_library()
    p = np.zeros(1, dtype=PARAMS_DTYPE)
    check(lib.hmj_default_params(_p(model), _p(p)))
    return np.repeat(p, n)


class HumanoidBatch:
    """N lock-step humanoid simulations on one GPU."""

    def __init__(self, n_envs: int, device: int = 0, model: np.ndarray | None = None):
        self.lib = load_library()
        self.model = load_model() if model is None else model
        self.n = int(n_envs)
        self.device = int(device)
        h = C.c_void_p()
        check(self.lib.hmj_batch_create(_p(self.model), self.n, self.device, C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self.lib.hmj_batch_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- parameters / state
    def default_params(self) -> np.ndarray:
        return default_params(self.model, self.n)

    def set_params(self, params: np.ndarray, first: int = 0):
        params = np.ascontiguousarray(params, dtype=PARAMS_DTYPE)
        check(self.lib.hmj_batch_set_params(self._h, _p(params), first, params.shape[0]))

    def get_params(self) -> np.ndarray:
        out = np.zeros(self.n, dtype=PARAMS_DTYPE)
        check(self.lib.hmj_batch_get_params(self._h, _p(out), 0, self.n))
        return out

    def _f64(self, a, shape):
        if a is None:
            return None
        a = np.ascontiguousarray(a, dtype=np.float64)
        if a.shape != (self.n,) + shape:
            a = np.ascontiguousarray(np.broadcast_to(a, (self.n,) + shape))
        return a

    def set_state(self, qpos=None, qvel=None, time=None, warmstart=None):
        a = [self._f64(qpos, (NQ,)), self._f64(qvel, (NV,)), self._f64(time, ()), self._f64(warmstart, (NV,))]
        check(self.lib.hmj_batch_set_state(self._h, *[_p(x) for x in a]))

    def get_state(self):
        qpos, qvel = np.zeros((self.n, NQ)), np.zeros((self.n, NV))
        time, ws = np.zeros(self.n), np.zeros((self.n, NV))
        check(self.lib.hmj_batch_get_state(self._h, _p(qpos), _p(qvel), _p(time), _p(ws)))
        return qpos, qvel, time, ws

    def set_ctrl(self, ctrl):
        check(self.lib.hmj_batch_set_ctrl(self._h, _p(self._f64(ctrl, (NU,)))))

    def set_xfrc(self, body, wrench):
        body = np.ascontiguousarray(np.broadcast_to(np.asarray(body, dtype=np.int32), (self.n,)))
        check(self.lib.hmj_batch_set_xfrc(self._h, _p(body), _p(self._f64(wrench, (6,)))))

    # ---- stepping
    def step(self, ctrl=None, nsteps: int = 1):
        check(self.lib.hmj_batch_step(self._h, _p(self._f64(ctrl, (NU,))), int(nsteps)))

    def step_dev(self, ctrl_ptr: int | None = None, nsteps: int = 1):
        check(self.lib.hmj_batch_step_dev(self._h, C.c_void_p(ctrl_ptr) if ctrl_ptr else None, int(nsteps)))

    def forward(self):
        check(self.lib.hmj_batch_forward(self._h))

    def run_stage(self, stage: int):
        check(self.lib.hmj_batch_run_stage(self._h, int(stage)))

    def sync(self):
        check(self.lib.hmj_batch_sync(self._h))

    # ---- results
    def get(self, name: str) -> np.ndarray:
        if name in _INT_ARRAYS:
            out = np.zeros(self.n, dtype=np.int32)
        elif name in _INT_SHAPES:
            out = np.zeros((self.n,) + _INT_SHAPES[name], dtype=np.int32)
        else:
            out = np.zeros((self.n,) + _ARRAY_SHAPES[name])
        check(self.lib.hmj_batch_get_array(self._h, name.encode(), _p(out), out.nbytes))
        return out

    def get_raw(self, name: str, dtype, shape) -> np.ndarray:
        out = np.zeros(shape, dtype=dtype)
        check(self.lib.hmj_batch_get_array(self._h, name.encode(), _p(out), out.nbytes))
        return out

    def set_raw(self, name: str, value: np.ndarray):
        """Upload a named device array from a host array of exactly its size (e.g. "contact": CONTACT_DTYPE [n][MAXCON])."""
        a = np.ascontiguousarray(value)
        check(self.lib.hmj_batch_set_array(self._h, name.encode(), _p(a), a.nbytes))

    def set(self, name: str, value):
        if name in _INT_ARRAYS:
            a = np.ascontiguousarray(np.broadcast_to(np.asarray(value, dtype=np.int32), (self.n,)))
        elif name in _INT_SHAPES:
            a = np.ascontiguousarray(np.broadcast_to(np.asarray(value, dtype=np.int32), (self.n,) + _INT_SHAPES[name]))
        else:
            a = self._f64(value, _ARRAY_SHAPES[name])
        check(self.lib.hmj_batch_set_array(self._h, name.encode(), _p(a), a.nbytes))

    def contacts(self):
        ncon = np.zeros(self.n, dtype=np.int32)
        con = np.zeros((self.n, MAXCON), dtype=CONTACT_DTYPE)
        check(self.lib.hmj_batch_get_contacts(self._h, _p(ncon), _p(con)))
        return ncon, con

    def device_ptr(self, name: str) -> int:
        p = self.lib.hmj_batch_device_ptr(self._h, name.encode())
        if not p:
            raise KeyError(name)
        return int(p)

    @property
    def stream(self) -> int:
        return int(self.lib.hmj_batch_stream(self._h) or 0)

    @property
    def kernel_launches(self) -> int:
        return int(self.lib.hmj_batch_kernel_launches(self._h))

    # ---- env epilogue
    def obs_reward(self, prev_action, latest_action, noise=None, max_sim_time=np.inf):
        obs, rew = np.zeros((self.n, 30)), np.zeros(self.n)
        term, trunc, selfc = (np.zeros(self.n, dtype=np.int32) for _ in range(3))
        check(self.lib.hmj_env_obs_reward(
            self._h, _p(self._f64(prev_action, (NU,))), _p(self._f64(latest_action, (NU,))), _p(self._f64(noise, (31,))),
            float(max_sim_time), _p(obs), _p(rew), _p(term), _p(trunc), _p(selfc)))
        return obs, rew, term.astype(bool), trunc.astype(bool), selfc.astype(bool)

    def obs_reward_noalloc(self, prev_action, latest_action, noise=None, max_sim_time=np.inf):
        """`obs_reward` through hmj_env_obs_reward_noalloc: device staging in the library's static array (batches of at
        most `hmj_env_obs_reward_noalloc_max_envs()` envs), no allocation inside the C call."""
        obs, rew = np.zeros((self.n, 30)), np.zeros(self.n)
        term, trunc, selfc = (np.zeros(self.n, dtype=np.int32) for _ in range(3))
        mst = 1e300 if not np.isfinite(max_sim_time) else float(max_sim_time)
        check(self.lib.hmj_env_obs_reward_noalloc(
            self._h, _p(self._f64(prev_action, (NU,))), _p(self._f64(latest_action, (NU,))), _p(self._f64(noise, (31,))),
            mst, _p(obs), _p(rew), _p(term), _p(trunc), _p(selfc)))
        return obs, rew, term.astype(bool), trunc.astype(bool), selfc.astype(bool)

    def env_step(self, ctrl, xfrc_body, xfrc, prev_action, latest_action, noise=None, max_sim_time=np.inf, nsteps=1, out=None):
        """CPUEnv.step for all envs in one C call (hmj_env_step).  `out` may hold preallocated (pinned) result arrays."""
        n = self.n
        if out is None:
            out = {"obs": np.zeros((n, 30)), "reward": np.zeros(n), "terminated": np.zeros(n, dtype=np.int32),
                   "truncated": np.zeros(n, dtype=np.int32), "self_collision": np.zeros(n, dtype=np.int32),
                   "time": np.zeros(n), "root_quat": np.zeros((n, 4))}
        body = None if xfrc_body is None else np.ascontiguousarray(np.broadcast_to(np.asarray(xfrc_body, dtype=np.int32), (n,)))
        check(self.lib.hmj_env_step(
            self._h, _p(self._f64(ctrl, (NU,))), _p(body), _p(self._f64(xfrc, (6,))), _p(self._f64(prev_action, (NU,))),
            _p(self._f64(latest_action, (NU,))), _p(self._f64(noise, (31,))), float(max_sim_time), int(nsteps),
            _p(out["obs"]), _p(out["reward"]), _p(out["terminated"]), _p(out["truncated"]), _p(out["self_collision"]),
            _p(out["time"]), _p(out["root_quat"])))
        return out

    def env_io(self):
        """numpy views of the pinned host IO block of the fused control step (hmj_env_io)."""
        if getattr(self, "_io", None) is None:
            ptrs = [C.c_void_p() for _ in range(9)]
            check(self.lib.hmj_env_io(self._h, *[C.byref(p) for p in ptrs]))
            n = self.n
            spec = [("action", np.float64, (n, NU)), ("mask", np.int32, (n,)), ("obs", np.float64, (n, 30)),
                    ("reward", np.float64, (n,)), ("time", np.float64, (n,)), ("torso_quat", np.float64, (n, 4)),
                    ("terminated", np.int32, (n,)), ("truncated", np.int32, (n,)), ("self_collision", np.int32, (n,))]
            io = {}
            for p, (name, dt, shape) in zip(ptrs, spec):
                cnt = int(np.prod(shape))
                ct = C.c_double if dt is np.float64 else C.c_int32
                io[name] = np.ctypeslib.as_array((ct * cnt).from_address(p.value)).reshape(shape)
            # terminated | truncated | self_collision are adjacent in the block (hmj_env_io): one byte view over all three
            io["flag_bytes"] = np.ctypeslib.as_array((C.c_bool * (12 * n)).from_address(ptrs[6].value))
            self._io = io
        return self._io

    def env_step_fused(self, action, has_mask: bool, rf: float, max_sim_time: float, seed: int, nsteps: int = 1):
        """CPUEnv.step with device-resident bookkeeping (hmj_env_step_fused[_from]); results land in env_io().
        `action`: True = the controls are in env_io()["action"], False / None = replay data.ctrl, or a C-contiguous
        float64 [N][12] host array that is copied to the device straight from where it lies (pinned memory, e.g.
        `pinned_empty`, makes that a plain DMA)."""
        mst = 1e300 if not np.isfinite(max_sim_time) else float(max_sim_time)
        if isinstance(action, np.ndarray):
            if action.dtype != np.float64 or not action.flags.c_contiguous or action.size != self.n * NU:
                raise ValueError("env_step_fused: action must be a C-contiguous float64 [N][12] array")
            ptr = action.ctypes.data
        elif action:
            ptr = self.env_io()["action"].ctypes.data
        else:
            ptr = None
        check(self.lib.hmj_env_step_fused_from(self._h, ptr, int(has_mask), float(rf), mst, int(seed) & (2**64 - 1), int(nsteps)))

    # ---- episode pool: CPUEnv.reset / auto-reset on the device
    @property
    def spare_rows(self) -> int:
        return int(self.lib.hmj_batch_spare_rows(self._h))

    def env_pool_config(self, enable: bool, rf: float, max_sim_time: float, seed: int, timestep: float):
        mst = 1e300 if not np.isfinite(max_sim_time) else float(max_sim_time)
        check(self.lib.hmj_env_pool_config(self._h, int(bool(enable)), float(rf), mst, int(seed) & (2**64 - 1), float(timestep)))

    def env_reset(self, ids=None):
        """CPUEnv.reset of the listed envs (None: all) on the device; their first observations land in env_io()."""
        if ids is None:
            check(self.lib.hmj_env_reset(self._h, None, 0))
        else:
            a = np.ascontiguousarray(ids, dtype=np.int32)
            check(self.lib.hmj_env_reset(self._h, a.ctypes.data, int(a.size)))

    def env_terminal(self):
        """(env ids, terminal observations [count][30]) of the episodes that ended in the last pooled step (copies)."""
        ids, obs, cnt = C.c_void_p(), C.c_void_p(), C.c_int32(0)
        check(self.lib.hmj_env_terminal(self._h, C.byref(ids), C.byref(obs), C.byref(cnt)))
        k = cnt.value
        if k == 0:
            return np.zeros(0, dtype=np.int64), np.zeros((0, 30))
        i = np.ctypeslib.as_array((C.c_int32 * k).from_address(ids.value)).astype(np.int64)
        o = np.ctypeslib.as_array((C.c_double * (k * 30)).from_address(obs.value)).reshape(k, 30).copy()
        return i, o

    def env_reset_io(self):
        """Views of the pool's per-step result block: counts[8], ended_ids[64], term_obs[64][30] (hmj_env_reset_io)."""
        if getattr(self, "_rio", None) is None:
            ptrs = [C.c_void_p() for _ in range(3)]
            check(self.lib.hmj_env_reset_io(self._h, *[C.byref(p) for p in ptrs]))
            self._rio = {
                "counts": np.ctypeslib.as_array((C.c_int32 * 8).from_address(ptrs[0].value)),
                "ended_ids": np.ctypeslib.as_array((C.c_int32 * 64).from_address(ptrs[1].value)),
                "term_obs": np.ctypeslib.as_array((C.c_double * (64 * 30)).from_address(ptrs[2].value)).reshape(64, 30)}
        return self._rio

    # ---- per-kernel timing
    def profile(self, enable: bool):
        check(self.lib.hmj_batch_profile(self._h, int(bool(enable))))

    def profile_read(self):
        a, b, c, n = C.c_double(0), C.c_double(0), C.c_double(0), C.c_int(0)
        check(self.lib.hmj_batch_profile_read(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(n)))
        return {"smooth_ms": a.value, "collide_ms": b.value, "solve_ms": c.value, "steps": n.value}

    # ---- planner rollouts
    def plan_rollouts(self, state, time, spline_times, spline_values, horizon, interpolation=2, record=False):
        state = np.ascontiguousarray(state, dtype=np.float64)
        st = np.ascontiguousarray(spline_times, dtype=np.float64)
        sv = np.ascontiguousarray(spline_values, dtype=np.float64)
        assert sv.shape == (self.n, st.shape[0], NU), sv.shape
        ret = np.zeros(self.n)
        fail = np.zeros(self.n, dtype=np.int32)
        check(self.lib.hmj_plan_rollouts(self._h, _p(state), float(time), _p(st), _p(sv), st.shape[0], int(interpolation),
                                         int(horizon), int(bool(record)), _p(ret), _p(fail)))
        return ret, fail.astype(bool)

    def plan_rollouts_noisy(self, state, time, spline_times, spline_values, horizon, interpolation=2, xfrc_std=0.1, xfrc_rate=0.1,
                            xfrc_noise=None, seed=0, first_global=0):
        """Trajectory::NoisyRollout for every sample (hmj_plan_rollouts_noisy): OU wrench on every body."""
        state = np.ascontiguousarray(state, dtype=np.float64)
        st = np.ascontiguousarray(spline_times, dtype=np.float64)
        sv = np.ascontiguousarray(spline_values, dtype=np.float64)
        assert sv.shape == (self.n, st.shape[0], NU), sv.shape
        nz = None
        if xfrc_noise is not None:
            nz = np.ascontiguousarray(xfrc_noise, dtype=np.float64)
            assert nz.shape == (self.n, max(horizon - 1, 0), 90), nz.shape
        ret = np.zeros(self.n)
        fail = np.zeros(self.n, dtype=np.int32)
        check(self.lib.hmj_plan_rollouts_noisy(self._h, _p(state), float(time), _p(st), _p(sv), st.shape[0], int(interpolation),
                                               int(horizon), float(xfrc_std), float(xfrc_rate), _p(nz) if nz is not None else None,
                                               int(seed) & (2**64 - 1), int(first_global), _p(ret), _p(fail)))
        return ret, fail.astype(bool)

    def plan_iterate(self, state, time, spline_times, nominal_values, horizon, interpolation=2, noise=None, noise_std=0.0,
                     noise_std2=0.0, seed=0, iteration=0, first_global=0, record_dev: int | None = None, noise_std_vec=None):
        """Device half of one OptimizePolicy iteration (hmj_plan_iterate): enqueue only."""
        state = np.ascontiguousarray(state, dtype=np.float64)
        st = np.ascontiguousarray(spline_times, dtype=np.float64)
        nv = np.ascontiguousarray(nominal_values, dtype=np.float64)
        assert nv.shape == (st.shape[0], NU), nv.shape
        nz = None
        if noise is not None:
            nz = np.ascontiguousarray(noise, dtype=np.float64)
            assert nz.shape == (self.n, st.shape[0], NU), nz.shape
        sv = None
        if noise_std_vec is not None:
            sv = np.ascontiguousarray(noise_std_vec, dtype=np.float64)
            assert sv.size == st.shape[0] * NU, sv.shape
        check(self.lib.hmj_plan_iterate_ex(self._h, _p(state), float(time), _p(st), _p(nv), st.shape[0], int(interpolation),
                                           int(horizon), _p(nz) if nz is not None else None, float(noise_std), float(noise_std2),
                                           _p(sv) if sv is not None else None, int(seed) & (2**64 - 1), int(iteration),
                                           int(first_global), record_dev))

    def plan_select(self, nnode: int, gathered_dev: int | None = None, nranks: int = 1) -> np.ndarray:
        """Global winner record {cost, global index, nominal return, spline values [nnode][12]} (hmj_plan_select)."""
        rec = np.zeros(self.lib.hmj_plan_record_size(int(nnode)))
        check(self.lib.hmj_plan_select(self._h, gathered_dev, int(nranks), int(nnode), _p(rec)))
        return rec

    def plan_population(self, nnode: int, values: bool = True):
        ret = np.zeros(self.n)
        fail = np.zeros(self.n, dtype=np.int32)
        vals = np.zeros((self.n, int(nnode), NU)) if values else None
        check(self.lib.hmj_plan_population(self._h, int(nnode), _p(ret), _p(fail), _p(vals) if values else None))
        return ret, fail.astype(bool), vals

    @property
    def host_launches(self) -> int:
        return int(self.lib.hmj_batch_host_launches(self._h))

    def inverse(self, qacc=None) -> np.ndarray:
        """mj_inverse for every env at its current state / ctrl (hmj_batch_inverse): qfrc_inverse [N][18]."""
        qa = None if qacc is None else self._f64(qacc, (NV,))
        out = np.zeros((self.n, NV))
        check(self.lib.hmj_batch_inverse(self._h, _p(qa) if qa is not None else None, _p(out)))
        return out

    def fd_transition(self, x, u, time, warmstart=None, eps: float = 1e-6, centered: bool = False):
        """mjd_transitionFD for T knots in one batched step (hmj_fd_transition): returns A [T][36][36], B [T][36][12],
        C [T][28][36], D [T][28][12].  Needs n_envs >= 97 T; overwrites the batch state."""
        x = np.ascontiguousarray(x, dtype=np.float64)
        u = np.ascontiguousarray(u, dtype=np.float64)
        t = np.ascontiguousarray(time, dtype=np.float64)
        T = x.shape[0]
        assert x.shape == (T, DEFINES["HMJ_NSTATE"]) and u.shape == (T, NU) and t.shape == (T,)
        ws = None if warmstart is None else np.ascontiguousarray(warmstart, dtype=np.float64)
        nr, ndx = DEFINES["HMJ_NRESIDUAL"], 2 * NV
        A, B, Cm, D = np.zeros((T, ndx, ndx)), np.zeros((T, ndx, NU)), np.zeros((T, nr, ndx)), np.zeros((T, nr, NU))
        check(self.lib.hmj_fd_transition(self._h, T, _p(x), _p(u), _p(t), _p(ws) if ws is not None else None, float(eps),
                                         int(bool(centered)), _p(A), _p(B), _p(Cm), _p(D)))
        return A, B, Cm, D

    def plan_argmin(self):
        cost, idx = C.c_double(0), C.c_int32(0)
        check(self.lib.hmj_plan_argmin(self._h, C.byref(cost), C.byref(idx)))
        return cost.value, idx.value

    def plan_fetch(self, sample: int, horizon: int):
        ns, nr = DEFINES["HMJ_NSTATE"], DEFINES["HMJ_NRESIDUAL"]
        out = {"states": np.zeros((horizon, ns)), "actions": np.zeros((horizon, NU)), "times": np.zeros(horizon),
               "residual": np.zeros((horizon, nr)), "costs": np.zeros(horizon), "trace": np.zeros((horizon, 3))}
        check(self.lib.hmj_plan_fetch(self._h, int(sample), int(horizon), _p(out["states"]), _p(out["actions"]), _p(out["times"]),
                                      _p(out["residual"]), _p(out["costs"]), _p(out["trace"])))
        return out


def pinned_empty(shape, dtype=np.float64):
    """A numpy array in page-locked host memory (hmj_pinned_alloc): step inputs placed here reach the device by DMA
    with no staging copy.  The memory is released when the array (and every view of it) is garbage collected."""
    import weakref

    from ._lib import load_library

    lib = load_library()
    dt = np.dtype(dtype)
    count = int(np.prod(shape))
    p = C.c_void_p()
    check(lib.hmj_pinned_alloc(max(count * dt.itemsize, 1), C.byref(p)))
    raw = (C.c_char * max(count * dt.itemsize, 1)).from_address(p.value)
    arr = np.frombuffer(raw, dtype=dt, count=count).reshape(shape)
    weakref.finalize(raw, lib.hmj_pinned_free, p.value)
    return arr
