"""Host-side logic that needs no GPU: the counter-based push stream, DR ranges (cpu_env.py:189-280 /
simulation_parameters.py:5-48), start poses (cpu_env.py:282-328)."""
import numpy as np

import humanoid_mujoco_b200  # noqa: F401  (registers the package alias)
from humanoid_mujoco_b200 import env as E


def _u01_exact(seed, env, c):
    M = (1 << 64) - 1

    def sm(x):
        x = (x + 0x9E3779B97F4A7C15) & M
        z = x
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
        return z ^ (z >> 31)

    return (sm((seed & M) ^ sm((env << 32) | c)) >> 11) / 9007199254740992.0


def test_rand_u01_matches_exact_integer_arithmetic():
    envs = np.array([0, 1, 77, 4095, 1 << 20])
    ctrs = np.array([0, 5, 8, 123456, (1 << 31) + 3])
    for seed in (0, 5, 2**63 + 11):
        got = E.rand_u01(seed, envs, ctrs)
        want = [_u01_exact(seed, int(e), int(c)) for e, c in zip(envs, ctrs)]
        assert np.array_equal(got, np.array(want))
    u = E.rand_u01(1, np.arange(20000), np.zeros(20000, dtype=np.int64))
    assert 0.0 <= u.min() and u.max() < 1.0 and abs(u.mean() - 0.5) < 0.01


def test_randomize_params_ranges():
    from humanoid_mujoco_b200.batch import PARAMS_DTYPE, load_model

    m = load_model()
    base = np.zeros(256, dtype=PARAMS_DTYPE)
    for f in ("body_mass", "dof_armature", "jnt_range", "jnt_margin", "jnt_stiffness", "act_gainprm0", "act_forcerange"):
        base[f] = m[f][0]
    base["floor_friction"] = m["geom_friction"][0][0]
    rng = np.random.default_rng(0)
    p = E.randomize_params(base, rng, 1.0)
    assert np.all(p["floor_friction"][:, 0] <= base["floor_friction"][:, 0] + 1e-15)
    assert np.all(p["floor_friction"][:, 0] >= 0.5 * base["floor_friction"][:, 0] - 1e-15)
    assert np.all(np.abs(p["act_gainprm0"] - 2.0) <= 0.5)
    assert np.all(np.abs(p["jnt_range"][:, 1:] - base["jnt_range"][:, 1:]) <= 0.1 + 1e-15)
    assert np.all(p["hfield_data"] >= 0) and np.all(p["hfield_data"] <= np.float32(0.1) + 1e-7)
    assert np.all(p["body_mass"][:, 1:] > 0)
    p0 = E.randomize_params(base, rng, 0.0)
    assert np.array_equal(p0["body_mass"][:, 2:], base["body_mass"][:, 2:]) and np.all(p0["hfield_data"] == 0)
    q = E.standup_start_state(m["qpos0"][0], rng, 1.0, 64)
    assert np.all((q[:, 2] >= -0.6) & (q[:, 2] <= -0.5)) and np.all(np.abs(q[:, 7:] - m["qpos0"][0][7:]) <= 0.2 + 1e-15)


def test_support_cells_contain_the_whole_hull_argmax():
    """Direction-cell candidate lists (hmj_support_cells, used by every support call of the collision kernels): for any
    direction, the first maximum over the cell's list must be the first maximum over the whole hull (that is what makes
    the contacts bit-identical to the exhaustive scan of the oracle).  Checked with the kernels' own cell look-up
    (face by exact comparison, position on the face in float32) on random, axis-aligned, face-normal and cell-border
    directions."""
    import ctypes as C
    from humanoid_mujoco_b200._lib import load_library, check
    from humanoid_mujoco_b200.batch import load_model

    lib = load_library()
    model = load_model()
    m = model[0]
    nmesh = 7
    grid, entries = C.c_int(0), C.c_int(0)
    start = np.zeros(nmesh * (6 * 8 * 8 + 2), dtype=np.uint16)
    vert = np.zeros(16384, dtype=np.uint16)
    check(lib.hmj_support_cells(model.ctypes.data, start.ctypes.data, start.size, vert.ctypes.data, vert.size,
                                C.byref(grid), C.byref(entries)))
    G = grid.value
    assert G == 8 and 0 < entries.value <= vert.size
    start = start.reshape(nmesh, 6 * G * G + 2)

    def cells(d):
        a = np.abs(d)
        ax = np.argmax(a, axis=1)                     # first maximum: lowest axis on ties, as the kernel
        rows = np.arange(len(d))
        am = a[rows, ax]
        dm, da, db = d[rows, ax], d[rows, (ax + 1) % 3], d[rows, (ax + 2) % 3]
        inv = np.float32(1.0) / am.astype(np.float32)
        u, v = da.astype(np.float32) * inv, db.astype(np.float32) * inv
        half = np.float32(0.5 * G)
        iu = np.clip(((u + np.float32(1.0)) * half).astype(np.int32), 0, G - 1)
        iv = np.clip(((v + np.float32(1.0)) * half).astype(np.int32), 0, G - 1)
        return ((ax << 1) | (dm < 0)) * (G * G) + iu * G + iv

    rng = np.random.default_rng(0)
    dirs = [rng.normal(size=(20000, 3)), np.vstack([np.eye(3), -np.eye(3)]),
            rng.normal(size=(2000, 3)) * np.array([1.0, 1e-9, 1e-9]), 1e-20 * rng.normal(size=(100, 3))]
    # directions on the cell borders (and a hair to either side), all faces
    t = -1.0 + 2.0 * np.arange(G + 1) / G
    for eps in (0.0, 1e-7, -1e-7, 1e-4, -1e-4):
        uu, vv = np.meshgrid(t + eps, rng.uniform(-1, 1, 5))
        for ax in range(3):
            for sgn in (1.0, -1.0):
                d = np.zeros((uu.size, 3))
                d[:, ax], d[:, (ax + 1) % 3], d[:, (ax + 2) % 3] = sgn, uu.ravel(), vv.ravel()
                dirs.append(d)
                dirs.append(d[:, [0, 1, 2]] * 3.7)
    total_len = 0
    for k in range(nmesh):
        adr, num = int(m["mesh_vertadr"][k]), int(m["mesh_vertnum"][k])
        P = m["mesh_vert"][adr:adr + num].astype(np.float64)
        # face normals of the hull: the directions where many vertices tie or nearly tie
        tri = rng.integers(0, num, (3000, 3))
        nrm = np.cross(P[tri[:, 1]] - P[tri[:, 0]], P[tri[:, 2]] - P[tri[:, 0]])
        nrm = nrm[np.linalg.norm(nrm, axis=1) > 1e-12]
        for d in dirs + [nrm, -nrm]:
            d = d[np.abs(d).max(axis=1) > 1e-30]
            c = cells(d)
            # the kernel's expression, term by term (no fused multiply-add)
            val = (d[:, 0:1] * P[:, 0][None, :] + d[:, 1:2] * P[:, 1][None, :]) + d[:, 2:3] * P[:, 2][None, :]
            want = np.argmax(val, axis=1)             # first maximum over the hull
            for j in range(len(d)):
                lst = vert[start[k, c[j]]:start[k, c[j] + 1]]
                assert len(lst) > 0 and np.all(np.diff(lst.astype(np.int32)) > 0)
                got = lst[np.argmax(val[j, lst])]
                assert got == want[j], (k, j, d[j], got, want[j])
        lens = np.diff(start[k, :6 * G * G + 1].astype(np.int32))
        assert lens.min() >= 1
        total_len += lens.sum()
    assert total_len == entries.value
