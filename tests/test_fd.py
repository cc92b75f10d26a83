"""Finite-differenced transition matrices (mjpc/planners/model_derivatives.cc:51-96 ->
mjd_transitionFD): the oracle's restatement against analytic facts of the integrator (CPU), and the batched GPU
implementation (hmj_fd_transition: all knots x all nudged copies in ONE step of the batch) against the oracle."""
import numpy as np
import pytest


@pytest.fixture(autouse=True)
def _restore_oracle_params(oracle):
    """The oracle fixture is shared by the whole session: put the default parameters back after each test here."""
    yield
    oracle.lib.orc_default_params(oracle.model.ctypes.data, oracle.params.ctypes.data)
    oracle.reset()


def plan_state(oracle, nsteps=50, timestep=0.02):
    oracle.reset()
    oracle.lib.orc_default_params(oracle.model.ctypes.data, oracle.params.ctypes.data)
    oracle.params["timestep"][0] = timestep
    oracle.set_state(np.array(oracle.model["qpos0"][0]), np.zeros(18), np.zeros(12))
    oracle.step(nsteps)
    return np.concatenate([oracle.d["qpos"], oracle.d["qvel"]]).copy(), float(oracle.d["time"])


def test_oracle_fd_free_flight_structure(oracle):
    """In free flight (no contact, no active limit) with semi-implicit Euler: q' = q (+) h v', so the position rows of A
    and B are h times the velocity rows (+ identity for dq/dq on the translational and hinge coordinates), and a forward
    and a centred difference agree to O(eps)."""
    oracle.reset()
    oracle.lib.orc_default_params(oracle.model.ctypes.data, oracle.params.ctypes.data)
    h = 0.002
    oracle.params["timestep"][0] = h
    rng = np.random.default_rng(0)
    q = np.array(oracle.model["qpos0"][0], dtype=np.float64)
    q[2] = 0.5                                        # well above the floor
    q[7:] += rng.uniform(-0.1, 0.1, 12)
    v = rng.uniform(-0.3, 0.3, 18)
    u = rng.uniform(-0.2, 0.2, 12)
    x = np.concatenate([q, v])
    A, B, Cm, D = oracle.transition_fd(x, u, 0.0, eps=1e-6, centered=True)
    A1, B1, _, _ = oracle.transition_fd(x, u, 0.0, eps=1e-6, centered=False)
    assert np.max(np.abs(A - A1)) < 1e-3 and np.max(np.abs(B - B1)) < 1e-4
    lin = [0, 1, 2] + list(range(6, 18))              # coordinates where (+) is plain addition
    eye = np.eye(18)
    assert np.max(np.abs(B[lin] - h * B[18:][lin])) < 1e-5
    assert np.max(np.abs(A[np.ix_(lin, range(18, 36))] - h * A[np.ix_([18 + i for i in lin], range(18, 36))])) < 1e-5
    assert np.max(np.abs(A[np.ix_(lin, lin)] - eye[np.ix_(lin, lin)] - h * A[np.ix_([18 + i for i in lin], lin)])) < 1e-5
    # residuals 4..15 are the joint velocities, 16..27 the controls (stand.cc:40-99): exact unit blocks of C and D
    assert np.max(np.abs(Cm[4:16, 24:36] - np.eye(12))) < 1e-6
    assert np.max(np.abs(D[16:28] - np.eye(12))) < 1e-6
    # a control at its upper limit is differenced backward, outside both limits its column is zero
    u2 = u.copy(); u2[3] = oracle.model["act_ctrlrange"][0][3][1]
    _, B2, _, D2 = oracle.transition_fd(x, u2, 0.0, eps=1e-6, centered=False)
    assert np.max(np.abs(B2[:, 3])) > 0 and abs(D2[19, 3] - 1.0) < 1e-6


@pytest.mark.gpu
def test_gpu_fd_matches_oracle(oracle):
    import humanoid_mujoco_b200 as hmj

    state, time = plan_state(oracle)
    # knots along a nominal rollout under a small spline policy
    rng = np.random.default_rng(3)
    T = 5
    times = np.array([time, time + 0.17, time + 0.34])
    vals = rng.normal(0, 0.05, (3, 12))
    r = oracle.rollout(state, time, times, vals, T + 1, 2)
    x, u, tt = r["states"][:T], r["actions"][:T].copy(), r["times"][:T]
    u[2, 5] = oracle.model["act_ctrlrange"][0][5][0]          # one control at its lower limit: forward nudge only
    b = hmj.HumanoidBatch(97 * T, device=0)
    p = b.default_params(); p["timestep"] = 0.02; b.set_params(p)
    ws = rng.normal(0, 0.1, (T, 18))
    for centered in (False, True):
        A, B, Cm, D = b.fd_transition(x, u, tt, warmstart=ws, eps=1e-6, centered=centered)
        for t in range(T):
            Ar, Br, Cr, Dr = oracle.transition_fd(x[t], u[t], tt[t], warmstart=ws[t], eps=1e-6, centered=centered)
            for got, ref, name in ((A[t], Ar, "A"), (B[t], Br, "B"), (Cm[t], Cr, "C"), (D[t], Dr, "D")):
                err = np.max(np.abs(got - ref) / np.maximum(np.abs(ref), 1.0))
                assert err < 1e-5, (centered, t, name, err)
    with pytest.raises(hmj.HmjError):
        hmj.HumanoidBatch(96, device=0).fd_transition(x[:1], u[:1], tt[:1])
    b.close()
