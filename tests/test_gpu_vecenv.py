"""The DummyVecEnv seam (rl/train.py:134-151, benchmark.py:37-39): auto-reset semantics of B200VecEnv."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hmj():
    import humanoid_mujoco_b200 as h

    return h


def test_spaces_and_attrs(hmj):
    v = hmj.B200VecEnv(4, randomization_factor=0.0, seed=1)
    assert v.num_envs == 4 and v.observation_space.shape == (30,) and v.action_space.shape == (12,)
    assert np.all(v.action_space.low == -1) and np.all(v.observation_space.high == 10)
    v.set_attr("randomization_factor", 0.5)           # randomization_adaptation_callback.py:53
    assert v.get_attr("randomization_factor") == [0.5] * 4
    with pytest.raises(ValueError):
        v.set_attr("randomization_factor", 1.0, indices=[0])
    obs = v.reset()
    assert obs.shape == (4, 30) and np.all(np.isfinite(obs))
    v.close()


def test_steps_match_the_plain_env_until_an_episode_ends(hmj):
    n = 8
    v = hmj.B200VecEnv(n, randomization_factor=1.0, seed=7)
    e = hmj.BatchedHumanoidEnv(n, randomization_factor=1.0, seed=7)
    o1 = v.reset()
    o2, _ = e.reset()
    assert np.array_equal(o1, o2)
    rng = np.random.default_rng(0)
    for _ in range(20):
        a = rng.uniform(-1, 1, (n, 12))
        ov, rv, dv, iv = v.step(a)
        oe, re_, te, tre, _ = e.step(a)
        if (te | tre).any():
            break
        assert np.array_equal(ov, oe) and np.array_equal(rv, re_) and not dv.any()
        assert all(i["TimeLimit.truncated"] is False and "terminal_observation" not in i for i in iv)
    v.close(); e.close()


def test_auto_reset_on_time_limit(hmj):
    """max_simulation_time_override = 0.01 s: `terminated` fires once data.time >= 0.01 (cpu_env.py:586-616 -> reward -100);
    the same call resets those envs and returns the first observation of the new episode."""
    n = 16
    v = hmj.B200VecEnv(n, randomization_factor=1.0, max_simulation_time_override=0.01, seed=3)
    v.reset()
    rng = np.random.default_rng(1)
    ended_once = np.zeros(n, dtype=bool)
    for k in range(12):
        obs, rew, dones, infos = v.step(rng.uniform(-1, 1, (n, 12)))
        assert obs.shape == (n, 30) and np.all(np.isfinite(obs))
        for i in range(n):
            if dones[i]:
                ended_once[i] = True
                assert "terminal_observation" in infos[i] and infos[i]["terminal_observation"].shape == (30,)
                assert "is_success" in infos[i]
                if not infos[i]["TimeLimit.truncated"]:
                    assert rew[i] == -100.0
                # new episode: reset ran its own first step, so the clock is one timestep in
                assert v.env.time[i] <= 2 * v.env.timestep + 1e-12
            else:
                assert "terminal_observation" not in infos[i]
        if dones.any():
            assert np.all(v.env.time[~dones] > 2 * v.env.timestep) or k < 2
    assert ended_once.all()
    v.close()
