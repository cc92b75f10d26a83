import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for sub in ("", "tools", "oracle"):
    p = os.path.join(REPO, sub)
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle_py import Oracle

    return Oracle()


def standup_state(model, rng=None, rf=0.0):
    """CPUEnv.reset start pose (simulation/cpu_env.py:282-328, standup reward selected)."""
    import numpy as np

    q = np.array(model["qpos0"][0], dtype=np.float64)
    q[2] = -0.6
    q[3:7] = [-0.5, -0.5, 0.5, 0.5]
    if rng is not None and rf > 0:
        q[7:] += rng.uniform(-0.2 * rf, 0.2 * rf, 12)
        q[2] += rng.uniform(0, 0.1 * rf)
        q[3:7] += rng.uniform(-0.1 * rf, 0.1 * rf, 4)
    return q


def gpu_available():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def make_scenarios(oracle, n, seed=0, settle=(0, 400), rf=1.0, ctrl_scale=1.0):
    """Seeded contact-rich test states produced by the ORACLE (checker): each env starts from the CPUEnv
    standup pose with DR'd parameters, is rolled `settle` steps under random controls, and the resulting
    (qpos, qvel, warmstart, ctrl, params) is the test input for both implementations."""
    import numpy as np
    from model_layout import PARAMS_DTYPE

    rng = np.random.default_rng(seed)
    qpos, qvel, ws, ctrl = np.zeros((n, 19)), np.zeros((n, 18)), np.zeros((n, 18)), np.zeros((n, 12))
    params = np.zeros(n, dtype=PARAMS_DTYPE)
    for i in range(n):
        oracle.reset()
        oracle.lib.orc_default_params(oracle.model.ctypes.data, oracle.params.ctypes.data)
        p = oracle.params
        p["hfield_data"][0] = (rng.uniform(0, 1, 64) * 0.1 * rf).astype(np.float32)
        p["act_gainprm0"][0] = 2 + rng.uniform(-0.5, 0.5, 12) * rf
        p["body_mass"][0][1:] = np.maximum(1e-5, p["body_mass"][0][1:] + rng.uniform(-0.05, 0.05, 14) * rf)
        p["dof_armature"][0][6:] += rng.uniform(0, 5e-4, 12) * rf
        p["jnt_margin"][0][1:] += rng.uniform(0, 0.05, 12) * rf
        p["jnt_range"][0][1:] += rng.uniform(-0.1, 0.1, (12, 2)) * rf
        p["act_forcerange"][0] += rng.uniform(-0.05, 0.05, (12, 2)) * rf
        p["floor_friction"][0] *= (1 - rf) + rng.uniform(0.5, 1.0) * rf
        oracle.set_state(standup_state(oracle.model, rng, 1.0), rng.uniform(-0.5, 0.5, 18))
        nsteps = int(rng.integers(settle[0], settle[1] + 1))
        for _ in range(nsteps):
            oracle.data["ctrl"][0] = rng.uniform(-1, 1, 12) * ctrl_scale
            oracle.step(1)
        assert oracle.d["status"] == 0
        qpos[i], qvel[i], ws[i] = oracle.d["qpos"], oracle.d["qvel"], oracle.d["qacc_warmstart"]
        ctrl[i] = rng.uniform(-1, 1, 12) * ctrl_scale
        params[i] = oracle.params[0]
    return qpos, qvel, ws, ctrl, params


def oracle_run(oracle, qpos, qvel, ws, ctrl, params, nsteps, fn="step"):
    """Run the oracle on one scenario; returns a copy of its orc_data_t."""
    oracle.reset()
    oracle.params[0] = params
    oracle.set_state(qpos, qvel, ctrl, ws)
    if fn == "forward":
        oracle.forward()
    else:
        oracle.step(nsteps)
    return oracle.data.copy()[0]
