"""Oracle pins for the mjpc pieces: golden vectors produced by the reference's own
norm.cc / spline.cc (tests/golden/make_mjpc_golden.py) and the known answers of
mjpc/test/spline/spline_test.cc:40-159, mjpc/test/tasks/task_test.cc:76-95."""
import ctypes as C
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
G = np.load(os.path.join(HERE, "golden", "mjpc_golden.npz"))


def test_spline_golden(oracle):
    worst = 0.0
    for k in range(G["spline_meta"].shape[0]):
        n, dim, interp = G["spline_meta"][k]
        out = oracle.spline_sample(G["spline_times"][k, :n], G["spline_values"][k, :n, :dim], interp, G["spline_query"][k])
        worst = max(worst, np.max(np.abs(out - G["spline_out"][k, :dim])))
    assert worst == 0.0  # same expression order as spline.cc -> bit-exact


def test_norm_golden(oracle):
    for k in range(G["norm_meta"].shape[0]):
        ntype, n = G["norm_meta"][k]
        y = oracle.norm(G["norm_in"][k, 3:3 + n], G["norm_in"][k, :3], ntype)
        assert y == pytest.approx(G["norm_out"][k], rel=1e-14, abs=1e-300)


def test_live_reference_lib_if_present(oracle):
    path = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "libmjpc_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref not built (reference tree absent)")
    lib = C.CDLL(path)
    lib.ref_norm.restype = C.c_double
    rng = np.random.default_rng(5)
    for _ in range(200):
        n = int(rng.integers(1, 6))
        times = np.cumsum(rng.uniform(0.02, 0.3, n))
        vals = rng.uniform(-1, 1, (n, 12))
        t = rng.uniform(times[0] - 0.1, times[-1] + 0.1)
        ref = np.zeros(12)
        lib.ref_spline_sample(times.ctypes.data_as(C.c_void_p), vals.ctypes.data_as(C.c_void_p), n, 12, 2,
                              C.c_double(t), ref.ctypes.data_as(C.c_void_p))
        assert np.array_equal(ref, oracle.spline_sample(times, vals, 2, t))


# ---- spline_test.cc known answers
def test_spline_empty(oracle):
    assert np.all(oracle.spline_sample(np.zeros(0), np.zeros((0, 10)), 2, 2.0) == 0.0)


@pytest.mark.parametrize("interp", [0, 1, 2])
def test_spline_one_two_nodes(oracle, interp):
    for t in (0.0, 2.0, 4.0):
        assert list(oracle.spline_sample([1.0], [[1.0, 2.0]], interp, t)) == [1.0, 2.0]
    tt, vv = [1.0, 2.0], [[1.0, 2.0], [3.0, 4.0]]
    assert list(oracle.spline_sample(tt, vv, interp, 0)) == [1.0, 2.0]
    assert list(oracle.spline_sample(tt, vv, interp, 1)) == [1.0, 2.0]
    assert list(oracle.spline_sample(tt, vv, interp, 2)) == [3.0, 4.0]
    assert list(oracle.spline_sample(tt, vv, interp, 3)) == [3.0, 4.0]


def test_spline_zero_linear_cubic(oracle):
    tt, vv = [1.0, 2.0], [[1.0, 2.0], [3.0, 4.0]]
    assert list(oracle.spline_sample(tt, vv, 0, 1.5)) == [1.0, 2.0]
    assert list(oracle.spline_sample(tt, vv, 1, 1.5)) == [2.0, 3.0]
    assert list(oracle.spline_sample(tt, vv, 2, 1.5)) == [2.0, 3.0]
    t4, v4 = [0.0, 1.0, 2.0, 3.0], [[1.0, 2.0], [1.0, 2.0], [3.0, 4.0], [3.0, 4.0]]
    assert list(oracle.spline_sample(t4, v4, 2, 1.5)) == [2.0, 3.0]
    for x in np.arange(0.0, 1.0001, 0.125):  # y = -x^3 + 2x^2 on nodes (-1,1),(0,0),(1,1)
        y = oracle.spline_sample([-1.0, 0.0, 1.0], [[1.0], [0.0], [1.0]], 2, x)[0]
        assert y == pytest.approx(-x ** 3 + 2 * x ** 2, abs=1e-15)


# ---- cost aggregation (task_test.cc:76-95 pattern on the humanoid_cap terms, task.xml:23-27)
def test_cost_terms_humanoid_cap(oracle):
    rng = np.random.default_rng(0)
    r = rng.uniform(-0.5, 0.5, 28)
    c = 100.0 * (np.sqrt(r[0] ** 2 + 0.1 ** 2) - 0.1)
    c += 50.0 * (np.sqrt(r[1] ** 2 + 0.1 ** 2) - 0.1)
    c += 10.0 * 0.5 * np.dot(r[2:4], r[2:4])
    c += 0.01 * 0.5 * np.dot(r[4:16], r[4:16])
    c += 0.025 * np.sum(0.3 ** 2 * (np.cosh(r[16:28] / 0.3) - 1.0))
    assert oracle.cost_value(r) == pytest.approx(c, rel=1e-13)


def test_cost_risk_transform(oracle):
    r = np.array([1e-3, 2e-3, 3e-3, 4e-3] + [0.0] * 24)
    c = oracle.cost_value(r)
    oracle.model["task_risk"] = 0.2
    try:
        assert oracle.cost_value(r) == pytest.approx((np.exp(0.2 * c) - 1.0) / 0.2, abs=1e-12)
    finally:
        oracle.model["task_risk"] = 0.0
