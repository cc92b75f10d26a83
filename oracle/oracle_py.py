"""ctypes binding of the CPU oracle (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  The product package never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(_HERE)
sys.path.insert(0, os.path.join(REPO, "tools"))
from model_layout import (  # noqa: E402
    DEFAULT_MODEL_PATH, DEFINES, MODEL_DTYPE, PARAMS_DTYPE, load_model, oracle_dtypes,
)

LIB_PATH = os.path.join(_HERE, "_build", "libhmj_oracle.so")
_structs, ORC_DEFINES = oracle_dtypes()
DATA_DTYPE = _structs["orc_data_t"]


def _host_key() -> str:
    """The oracle is built -march=native: identify the host CPU so a library built elsewhere is rebuilt here."""
    import hashlib
    try:
        flags = next(ln for ln in open("/proc/cpuinfo") if ln.startswith("flags"))
    except Exception:
        flags = "unknown"
    return hashlib.sha1(flags.encode()).hexdigest()


def build(force: bool = False) -> str:
    stamp = os.path.join(_HERE, "_build", "host.key")
    key = _host_key()
    have = open(stamp).read().strip() if os.path.exists(stamp) else ""
    if force or not os.path.exists(LIB_PATH) or have != key:
        subprocess.check_call(["make", "-B", "-C", _HERE, "_build/libhmj_oracle.so"], stdout=subprocess.DEVNULL)
        with open(stamp, "w") as f:
            f.write(key)
    else:
        subprocess.check_call(["make", "-C", _HERE, "_build/libhmj_oracle.so"], stdout=subprocess.DEVNULL)
    return LIB_PATH


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    """Single-environment oracle: owns a model blob, a params struct and an orc_data_t."""

    def __init__(self, model_path: str = DEFAULT_MODEL_PATH):
        self.lib = C.CDLL(build())
        self.model = load_model(model_path).copy()
        assert self.lib.orc_sizeof_model() == MODEL_DTYPE.itemsize
        assert self.lib.orc_sizeof_params() == PARAMS_DTYPE.itemsize
        assert self.lib.orc_sizeof_data() == DATA_DTYPE.itemsize, (self.lib.orc_sizeof_data(), DATA_DTYPE.itemsize)
        self.params = np.zeros(1, dtype=PARAMS_DTYPE)
        self.data = np.zeros(1, dtype=DATA_DTYPE)
        self.lib.orc_default_params(_ptr(self.model), _ptr(self.params))
        self.lib.orc_reset_data(_ptr(self.model), _ptr(self.data))
        self.lib.orc_norm.restype = C.c_double
        self.lib.orc_cost_value.restype = C.c_double

    # ---- convenience accessors
    @property
    def d(self):
        return self.data[0]

    @property
    def p(self):
        return self.params[0]

    def reset(self):
        self.lib.orc_reset_data(_ptr(self.model), _ptr(self.data))

    def set_state(self, qpos, qvel, ctrl=None, warmstart=None, time=None):
        self.data["qpos"][0] = qpos
        self.data["qvel"][0] = qvel
        if ctrl is not None:
            self.data["ctrl"][0] = ctrl
        if warmstart is not None:
            self.data["qacc_warmstart"][0] = warmstart
        if time is not None:
            self.data["time"][0] = time

    def call(self, name: str):
        getattr(self.lib, name)(_ptr(self.model), _ptr(self.params), _ptr(self.data))

    def forward(self):
        self.call("orc_forward")

    def step(self, n: int = 1):
        return self.lib.orc_step_n(_ptr(self.model), _ptr(self.params), _ptr(self.data), int(n))

    def contacts(self):
        n = int(self.d["ncon"])
        return self.d["contact"][:n]

    # ---- mjpc pieces
    def spline_sample(self, times, values, interpolation, t):
        times = np.ascontiguousarray(times, dtype=np.float64)
        values = np.ascontiguousarray(values, dtype=np.float64)
        n = times.shape[0]
        dim = values.shape[1] if n else 1
        out = np.zeros(dim)
        self.lib.orc_spline_sample(_ptr(times), _ptr(values), n, dim, int(interpolation), C.c_double(t), _ptr(out))
        return out

    def norm(self, x, params, ntype):
        x = np.ascontiguousarray(x, dtype=np.float64)
        prm = np.zeros(3)
        prm[: len(params)] = params
        return float(self.lib.orc_norm(_ptr(x), _ptr(prm), x.size, int(ntype)))

    def cost_value(self, residual):
        r = np.ascontiguousarray(residual, dtype=np.float64)
        return float(self.lib.orc_cost_value(_ptr(self.model), _ptr(r)))

    def rollout(self, state, time, spline_times, spline_values, horizon, interpolation=2):
        ns, nu, nr = DEFINES["HMJ_NSTATE"], DEFINES["HMJ_NU"], DEFINES["HMJ_NRESIDUAL"]
        state = np.ascontiguousarray(state, dtype=np.float64)
        st = np.ascontiguousarray(spline_times, dtype=np.float64)
        sv = np.ascontiguousarray(spline_values, dtype=np.float64)
        out = {
            "states": np.zeros((horizon, ns)), "actions": np.zeros((horizon, nu)), "times": np.zeros(horizon),
            "residual": np.zeros((horizon, nr)), "costs": np.zeros(horizon), "trace": np.zeros((horizon, 3)),
        }
        tot = C.c_double(0)
        fail = self.lib.orc_rollout(
            _ptr(self.model), _ptr(self.params), _ptr(state), C.c_double(time), _ptr(st), _ptr(sv), st.shape[0],
            int(interpolation), int(horizon), _ptr(out["states"]), _ptr(out["actions"]), _ptr(out["times"]),
            _ptr(out["residual"]), _ptr(out["costs"]), _ptr(out["trace"]), C.byref(tot))
        out["total_return"] = tot.value
        out["failure"] = bool(fail)
        return out

    def rollout_noisy(self, state, time, spline_times, spline_values, horizon, interpolation, xfrc_std, xfrc_rate, noise):
        """Trajectory::NoisyRollout with injected unit normals `noise` [horizon-1][90]: (total_return, failure)."""
        state = np.ascontiguousarray(state, dtype=np.float64)
        st = np.ascontiguousarray(spline_times, dtype=np.float64)
        sv = np.ascontiguousarray(spline_values, dtype=np.float64)
        nz = np.ascontiguousarray(noise, dtype=np.float64)
        tot = C.c_double(0)
        fail = self.lib.orc_rollout_noisy(_ptr(self.model), _ptr(self.params), _ptr(state), C.c_double(time), _ptr(st), _ptr(sv),
                                          st.shape[0], int(interpolation), int(horizon), C.c_double(xfrc_std),
                                          C.c_double(xfrc_rate), _ptr(nz), C.byref(tot))
        return tot.value, bool(fail)

    def inverse(self, qacc=None):
        """mj_inverse at the current state for the given acceleration (zeros by default): qfrc_inverse [18]."""
        nv = DEFINES["HMJ_NV"]
        qa = np.zeros(nv) if qacc is None else np.ascontiguousarray(qacc, dtype=np.float64)
        out = np.zeros(nv)
        self.lib.orc_inverse(_ptr(self.model), _ptr(self.params), _ptr(self.data), _ptr(qa), _ptr(out))
        return out

    def transition_fd(self, x, u, time, warmstart=None, eps=1e-6, centered=False):
        """mjd_transitionFD for one knot (hmj_oracle_fd.c): A [36][36], B [36][12], C [28][36], D [28][12]."""
        nv, nu, nr = DEFINES["HMJ_NV"], DEFINES["HMJ_NU"], DEFINES["HMJ_NRESIDUAL"]
        x = np.ascontiguousarray(x, dtype=np.float64)
        u = np.ascontiguousarray(u, dtype=np.float64)
        ws = None if warmstart is None else np.ascontiguousarray(warmstart, dtype=np.float64)
        A, B, Cm, D = np.zeros((2 * nv, 2 * nv)), np.zeros((2 * nv, nu)), np.zeros((nr, 2 * nv)), np.zeros((nr, nu))
        self.lib.orc_transition_fd(_ptr(self.model), _ptr(self.params), _ptr(x), _ptr(u), C.c_double(time),
                                   _ptr(ws) if ws is not None else None, C.c_double(eps), int(bool(centered)),
                                   _ptr(A), _ptr(B), _ptr(Cm), _ptr(D))
        return A, B, Cm, D


class OracleBatch:
    """N oracle environments stepped by the pthread harness (hmj_oracle_batch.c): CPU-baseline timing and
    bulk generation of oracle outputs.  TEST INFRASTRUCTURE ONLY."""

    def __init__(self, n: int, model_path: str = DEFAULT_MODEL_PATH):
        self.lib = C.CDLL(build())
        self.lib.orc_step_batch.restype = C.c_double
        self.lib.orc_rollout_batch.restype = C.c_double
        self.n = int(n)
        self.model = load_model(model_path).copy()
        self.params = np.zeros(self.n, dtype=PARAMS_DTYPE)
        p0 = np.zeros(1, dtype=PARAMS_DTYPE)
        self.lib.orc_default_params(_ptr(self.model), _ptr(p0))
        self.params[:] = p0[0]
        self.data = np.zeros(self.n, dtype=DATA_DTYPE)
        for i in range(self.n):
            self.lib.orc_reset_data(_ptr(self.model), C.c_void_p(self.data.ctypes.data + i * DATA_DTYPE.itemsize))

    def set_state(self, qpos, qvel, warmstart=None, time=None):
        self.data["qpos"][:] = qpos
        self.data["qvel"][:] = qvel
        if warmstart is not None:
            self.data["qacc_warmstart"][:] = warmstart
        if time is not None:
            self.data["time"][:] = time

    def step(self, ctrl, nsteps: int, nthreads: int) -> float:
        """ctrl: [S][n][12] (cycled over steps), [n][12] or None (keep data.ctrl).  Returns wall seconds."""
        cs = 0
        if ctrl is not None:
            ctrl = np.ascontiguousarray(ctrl, dtype=np.float64)
            if ctrl.ndim == 2:
                ctrl = ctrl[None]
            assert ctrl.shape[1:] == (self.n, DEFINES["HMJ_NU"]), ctrl.shape
            cs = ctrl.shape[0]
        return float(self.lib.orc_step_batch(_ptr(self.model), _ptr(self.params), _ptr(self.data), self.n,
                                             _ptr(ctrl) if ctrl is not None else None, cs, int(nsteps), int(nthreads)))

    def rollouts(self, state, time, spline_times, spline_values, horizon, interpolation=2, nthreads=1, params=None):
        state = np.ascontiguousarray(state, dtype=np.float64)
        st = np.ascontiguousarray(spline_times, dtype=np.float64)
        sv = np.ascontiguousarray(spline_values, dtype=np.float64)
        n = sv.shape[0]
        ret = np.zeros(n)
        fail = np.zeros(n, dtype=np.int32)
        p = self.params[:1] if params is None else params
        secs = float(self.lib.orc_rollout_batch(
            _ptr(self.model), _ptr(p), _ptr(state), C.c_double(time), _ptr(st), _ptr(sv), n, st.shape[0],
            int(interpolation), int(horizon), _ptr(ret), _ptr(fail), int(nthreads)))
        return ret, fail.astype(bool), secs
