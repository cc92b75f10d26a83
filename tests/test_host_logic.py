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
