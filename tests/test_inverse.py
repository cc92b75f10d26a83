"""mj_inverse (controllers/mpc_utils.py:36-55): the oracle's restatement against the forward
dynamics it inverts (CPU), the batched GPU call against the oracle, and the height sweep of the MPC workaround."""
import numpy as np
import pytest

from conftest import make_scenarios


@pytest.fixture(autouse=True)
def _restore_oracle_params(oracle):
    """The oracle fixture is shared by the whole session: put the default parameters back after each test here."""
    yield
    oracle.lib.orc_default_params(oracle.model.ctypes.data, oracle.params.ctypes.data)
    oracle.reset()


def test_oracle_inverse_inverts_forward(oracle):
    """mj_inverse(mj_forward(x, u).qacc) returns the applied generalized forces: qfrc_actuator (+ J' xfrc), to the
    solver's tolerance."""
    qpos, qvel, ws, ctrl, params = make_scenarios(oracle, 6, seed=5, settle=(50, 200))
    for i in range(6):
        oracle.reset()
        oracle.params[0] = params[i]
        oracle.set_state(qpos[i], qvel[i], ctrl[i])
        oracle.data["qacc_warmstart"][0] = ws[i]
        oracle.forward()
        qacc = oracle.d["qacc"].copy()
        act = oracle.d["qfrc_actuator"].copy()
        inv = oracle.inverse(qacc)
        scale = max(1.0, np.max(np.abs(oracle.d["qfrc_bias"])))
        assert np.max(np.abs(inv - act)) < 1e-5 * scale, (i, np.max(np.abs(inv - act)))


@pytest.mark.gpu
def test_gpu_inverse_matches_oracle(oracle):
    import humanoid_mujoco_b200 as hmj

    n = 24
    a = make_scenarios(oracle, n // 2, seed=9, settle=(0, 300))                       # thrashing: few contacts
    c = make_scenarios(oracle, n // 2, seed=10, settle=(200, 300), ctrl_scale=0.05)    # lying on the floor: contact-rich
    qpos, qvel, ws, ctrl, params = (np.concatenate([x, y]) for x, y in zip(a, c))
    rng = np.random.default_rng(1)
    qacc = rng.normal(0, 5.0, (n, 18))
    qacc[: n // 3] = 0.0                                   # the MPC workaround's case
    b = hmj.HumanoidBatch(n, device=0)
    b.set_params(params)
    b.set_state(qpos, qvel, np.zeros(n), ws)
    b.set_ctrl(ctrl)
    got = b.inverse(qacc)
    ncon = b.get("ncon")
    for i in range(n):
        oracle.reset()
        oracle.params[0] = params[i]
        oracle.set_state(qpos[i], qvel[i], ctrl[i])
        ref = oracle.inverse(qacc[i])
        assert int(oracle.d["ncon"]) == int(ncon[i])
        err = np.max(np.abs(got[i] - ref)) / max(1.0, np.max(np.abs(ref)))
        assert err < 1e-9, (i, err)
    assert (ncon > 0).sum() >= n // 3 and (ncon == 0).any()   # constrained and unconstrained envs both covered
    b.close()


@pytest.mark.gpu
def test_height_sweep(oracle):
    """find_ideal_distance_to_ground: 200 candidate heights in one batched mj_inverse; the vertical force equals the
    oracle's at sampled offsets and the chosen offset is the oracle's arg-min."""
    import humanoid_mujoco_b200 as hmj
    from humanoid_mujoco_b200.mpc_utils import N_OFFSETS, find_ideal_distance_to_ground

    b = hmj.HumanoidBatch(N_OFFSETS, device=0)
    base = np.array(b.model["qpos0"][0], dtype=np.float64)
    base[2] = -0.45                                       # standing height above the (flat default) floor, roughly
    idx = np.array(b.model["act_qposadr"][0])
    joints = base[idx] + 0.05
    quat = np.array([1.0, 0.0, 0.0, 0.0])
    best, vertical = find_ideal_distance_to_ground(b, base, joints, quat, idx)
    offsets = np.linspace(-0.1, 0.1, N_OFFSETS)
    ref = np.zeros(N_OFFSETS)
    for k in range(N_OFFSETS):
        q = base.copy(); q[3:7] = quat; q[idx] = joints; q[2] += offsets[k]
        oracle.reset()
        oracle.lib.orc_default_params(oracle.model.ctypes.data, oracle.params.ctypes.data)
        oracle.set_state(q, np.zeros(18), np.zeros(12))
        ref[k] = oracle.inverse(None)[2]
    assert np.max(np.abs(vertical - ref) / np.maximum(np.abs(ref), 1.0)) < 1e-9
    assert best == offsets[int(np.argmin(np.abs(ref)))]
    b.close()
