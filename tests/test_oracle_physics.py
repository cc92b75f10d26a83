"""Self-consistency checks of the mj_step restatement (the reference pins nothing on the
humanoid path): independent re-derivations inside the test must agree with
the oracle's stage outputs."""
import ctypes as C

import numpy as np
import pytest

from conftest import standup_state

NV, NBODY = 18, 15


def dense_M(o):
    m, d = o.model, o.d
    M = np.zeros((NV, NV))
    for i in range(NV):
        adr = m["dof_Madr"][0][i]
        j = i
        while j >= 0:
            M[i, j] = M[j, i] = d["qM"][adr]
            adr += 1
            j = m["dof_parentid"][0][j]
    return M


def body_jac(o, body, point):
    jp, jr = np.zeros((3, NV)), np.zeros((3, NV))
    pt = np.ascontiguousarray(point, dtype=np.float64)
    o.lib.orc_jac(o.model.ctypes.data_as(C.c_void_p), o.data.ctypes.data_as(C.c_void_p),
                  jp.ctypes.data_as(C.c_void_p), jr.ctypes.data_as(C.c_void_p), pt.ctypes.data_as(C.c_void_p), body)
    return jp, jr


def random_state(o, seed):
    rng = np.random.default_rng(seed)
    q = standup_state(o.model, rng, 1.0)
    q[2] += 0.3  # in the air
    v = rng.uniform(-1, 1, NV)
    return q, v, rng


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_mass_matrix_matches_jacobian_sum(oracle, seed):
    o = oracle
    q, v, _ = random_state(o, seed)
    o.reset(); o.set_state(q, v); o.forward()
    M = dense_M(o)
    M2 = np.diag(o.p["dof_armature"])
    for b in range(1, NBODY):
        jp, jr = body_jac(o, b, o.d["xipos"][b])
        R = o.d["ximat"][b].reshape(3, 3)
        Iw = R @ np.diag(o.model["body_inertia"][0][b]) @ R.T
        M2 = M2 + o.p["body_mass"][b] * jp.T @ jp + jr.T @ Iw @ jr
    assert np.allclose(M, M2, rtol=1e-11, atol=1e-13)
    assert np.all(np.linalg.eigvalsh(M) > 0)
    # LDL solve consistency
    assert np.allclose(M @ o.d["qacc_smooth"], o.d["qfrc_smooth"], rtol=1e-10, atol=1e-11)


def test_bias_is_gravity_at_rest(oracle):
    o = oracle
    q, _, _ = random_state(o, 3)
    o.reset(); o.set_state(q, np.zeros(NV)); o.forward()
    g = np.zeros(NV)
    for b in range(1, NBODY):
        jp, _ = body_jac(o, b, o.d["xipos"][b])
        g += jp.T @ (o.p["body_mass"][b] * np.array([0, 0, 9.81]))
    assert np.allclose(o.d["qfrc_bias"], g, rtol=1e-10, atol=1e-12)


def test_bias_matches_finite_difference_of_momentum(oracle):
    """qfrc_bias - gravity = C(q,v)v: check via d/dt(M v) - 0.5 d/dq (v'Mv) on a hinge dof."""
    o = oracle
    q, v, _ = random_state(o, 4)
    v[:6] = 0  # keep it simple: base fixed velocity zero
    o.reset(); o.set_state(q, v); o.forward()
    bias = o.d["qfrc_bias"].copy()
    o.set_state(q, np.zeros(NV)); o.forward()
    grav = o.d["qfrc_bias"].copy()
    cor = bias - grav
    eps = 1e-6

    def Mq(qq):
        o.set_state(qq, v); o.forward()
        return dense_M(o)

    # Lagrange: c = Mdot v - 0.5 * grad_q (v' M v); hinges only (qpos[7+i] <-> dof 6+i)
    Mdot = np.zeros((NV, NV))
    grad = np.zeros(NV)
    for i in range(12):
        dq = np.zeros(19); dq[7 + i] = eps
        Mp, Mm = Mq(q + dq), Mq(q - dq)
        dM = (Mp - Mm) / (2 * eps)
        Mdot += dM * v[6 + i]
        grad[6 + i] = v @ dM @ v
    c = Mdot @ v - 0.5 * grad
    assert np.allclose(cor[6:], c[6:], rtol=1e-5, atol=1e-7)


def settle(o, n=600):
    o.reset()
    o.set_state(standup_state(o.model), np.zeros(NV))
    o.step(n)


def test_resting_contact_supports_weight_and_kkt(oracle):
    o = oracle
    settle(o)
    d = o.d
    assert d["status"] == 0 and d["ncon"] >= 3
    nefc = int(d["nefc"])
    J = d["efc_J"][:nefc]
    f = d["efc_force"][:nefc]
    # all floor contacts: geom1 == 3 (hfield before mesh), normals ~ +z, penetrating
    for c in o.contacts():
        assert c["geom1"] == 3 and c["frame"][2] > 0.99 and c["dist"] < 0
    # KKT of the primal problem: M a - tau = J' f, f = -D * min(0, J a - aref), f >= 0
    M = dense_M(o)
    grad = M @ d["qacc"] - d["qfrc_smooth"] - J.T @ f
    assert np.max(np.abs(grad)) < 1e-6
    jar = J @ d["qacc"] - d["efc_aref"][:nefc]
    assert np.allclose(f, -d["efc_D"][:nefc] * np.minimum(jar, 0), rtol=1e-9, atol=1e-12)
    assert np.all(f >= 0)
    # vertical force balance at (near) rest: contact forces carry the weight
    total_mass = o.p["body_mass"].sum()
    fz = (J.T @ f)[2]
    assert fz == pytest.approx(total_mass * 9.81, rel=0.05)


def test_energy_dissipates_without_actuation(oracle):
    o = oracle
    o.reset()
    o.set_state(standup_state(o.model), np.zeros(NV))

    def energy():
        o.forward()
        M = dense_M(o)
        ke = 0.5 * o.d["qvel"] @ M @ o.d["qvel"]
        pe = sum(o.p["body_mass"][b] * 9.81 * o.d["xipos"][b][2] for b in range(1, NBODY))
        return ke + pe

    e0 = energy()
    o.step(400)
    e1 = energy()
    assert e1 < e0 and o.d["status"] == 0


def test_joint_limit_rows(oracle):
    o = oracle
    q = standup_state(o.model); q[2] = 0.5
    q[7] = 0.2  # right_shoulder_roll range [-1.5708, 0] -> upper limit violated
    o.reset(); o.set_state(q, np.zeros(NV)); o.forward()
    d = o.d
    assert d["nefc"] >= 1
    rows = [(int(d["efc_id"][i]), d["efc_pos"][i], d["efc_J"][i][6]) for i in range(d["nefc"]) if d["efc_type"][i] == 3]
    assert any(j == 1 and pos == pytest.approx(-0.2) and jac == -1.0 for j, pos, jac in rows)
    assert d["qacc"][6] < d["qacc_smooth"][6]  # limit force pushes the joint back


def test_bad_state_resets_and_flags(oracle):
    o = oracle
    o.reset()
    q = standup_state(o.model); q[0] = np.nan
    o.set_state(q, np.zeros(NV))
    st = o.step(1)
    assert st & 1  # HMJ_STATUS_BADQPOS
    assert np.all(np.isfinite(o.d["qpos"]))
