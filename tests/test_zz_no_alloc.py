"""hmj_env_obs_reward_noalloc (allocation-free variant: device staging in a static array of the library) against
hmj_env_obs_reward (staging block allocated on first use), through the C-ABI.  Both run the same kernel on the same
state, so the bar is the one of the other epilogue tests: bit-exact doubles, equal flags.  The file sorts last on
purpose; it must also pass alone in a fresh process."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N = 96


def _batch_on_the_floor(hmj, n, seed):
    """n robots dropped on the bumpy floor and shaken a little: contacts, torques and tilted torsos in the state."""
    rng = np.random.default_rng(seed)
    b = hmj.HumanoidBatch(n)
    qpos = np.tile(b.model["qpos0"][0], (n, 1))
    qpos[:, 2] += rng.uniform(0.0, 0.05, n)
    qpos[:, 7:] += rng.uniform(-0.3, 0.3, (n, 12))
    z = np.zeros((n, 18))
    b.set_state(qpos, z, np.zeros(n), z)
    b.step(np.zeros((n, 12)), 150)
    for _ in range(5):
        b.step(rng.uniform(-1, 1, (n, 12)), 1)
    return b, rng


def _same(a, c):
    for x, y in zip(a, c):
        assert x.dtype == y.dtype and np.array_equal(x, y)


def test_noalloc_matches_allocating_entry_point_twice():
    import humanoid_mujoco_b200 as hmj

    b, rng = _batch_on_the_floor(hmj, N, 11)
    assert b.lib.hmj_env_obs_reward_noalloc_max_envs() >= N
    # call 1: noise on, no time limit
    prev, lat, noise = rng.uniform(-1.5, 1.5, (N, 12)), rng.uniform(-1.5, 1.5, (N, 12)), rng.normal(0, 0.05, (N, 31))
    got1 = b.obs_reward_noalloc(prev, lat, noise)
    want1 = b.obs_reward(prev, lat, noise)
    _same(got1, want1)
    assert np.all(np.isfinite(got1[0])) and np.all(np.isfinite(got1[1])) and np.abs(got1[0]).sum() > 0
    # call 2, right behind it on another state: no noise (NULL), a time limit that has passed (terminated, reward -100);
    # nothing of call 1 may survive in the shared array
    b.step(rng.uniform(-1, 1, (N, 12)), 3)
    prev2, lat2 = lat, rng.uniform(-1.5, 1.5, (N, 12))
    got2 = b.obs_reward_noalloc(prev2, lat2, None, max_sim_time=0.1)
    want2 = b.obs_reward(prev2, lat2, None, max_sim_time=0.1)
    _same(got2, want2)
    assert got2[2].all() and np.all(got2[1] == -100.0)
    assert not np.array_equal(got1[0], got2[0])
    # and the first inputs again give the first answer's flags / shapes on the new state, equal to the allocating path
    _same(b.obs_reward_noalloc(prev, lat, noise), b.obs_reward(prev, lat, noise))
    b.close()


def test_noalloc_rejects_batches_beyond_the_static_scratch():
    import humanoid_mujoco_b200 as hmj
    from humanoid_mujoco_b200._lib import HmjError

    limit = hmj.HumanoidBatch(1).lib.hmj_env_obs_reward_noalloc_max_envs()
    assert limit == 4096
    n = limit + 1
    b = hmj.HumanoidBatch(n)
    a = np.zeros((n, 12))
    with pytest.raises(HmjError, match="exceeds the static scratch"):
        b.obs_reward_noalloc(a, a, None)
    b.close()
