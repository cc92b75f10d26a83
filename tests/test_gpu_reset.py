"""CPUEnv.reset and DummyVecEnv's auto-reset on the device (cpu_env.py:135-416, rl/train.py:134-151):
the episode pool of csrc/hmj_reset.cu against its numpy restatement (env.device_reset_draws), the oracle's first step,
and an undisturbed twin."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hmj():
    import humanoid_mujoco_b200 as h

    return h


def test_device_reset_draws_match_numpy_and_oracle(hmj, oracle):
    from humanoid_mujoco_b200.env import device_reset_draws

    n, seed = 96, 5
    env = hmj.BatchedHumanoidEnv(n, randomization_factor=1.0, seed=seed)
    assert env.device_reset and env.batch.spare_rows >= 32
    obs, _ = env.reset()
    assert obs.shape == (n, 30) and np.all(np.isfinite(obs))
    b = env.batch
    src = b.get_raw("env_episode_src", np.int32, (n,))
    rows, gens = src & 0xFFFF, src >> 16
    assert len(set(zip(rows.tolist(), gens.tolist()))) == n          # every env got its own episode
    assert gens.max() >= 1                                            # the rejection rule made some rows draw again
    p_ref, q_ref, ad, od, an, on = device_reset_draws(b.model, b.default_params(), seed, n + rows, gens, 1.0)
    p_dev = b.get_params()
    for f in ("timestep", "body_mass", "dof_armature", "jnt_range", "jnt_margin", "jnt_stiffness", "act_gainprm0",
              "act_forcerange", "floor_friction"):
        assert np.array_equal(p_dev[f], p_ref[f]), f                 # same IEEE operation sequence on both sides
    assert np.max(np.abs(p_dev["hfield_data"] - p_ref["hfield_data"])) < 1e-7
    assert np.array_equal(b.get("env_act_delay"), ad) and np.array_equal(b.get("env_obs_delay"), od)
    assert ad.min() >= 5 and ad.max() <= 25                           # 10..50 ms at rf = 1, dt = 2 ms
    assert np.max(np.abs(b.get("env_act_noise") - an)) < 1e-15 and np.max(np.abs(b.get("env_obs_noise") - on)) < 1e-14
    # reset accepts only contact-free first steps, and leaves every env one probe step into its episode
    assert np.all(b.get("ncon") == 0)
    q, v, t, _ = b.get_state()
    assert np.allclose(t, 0.002)
    assert np.array_equal(b.get("env_act_count"), np.ones(n, dtype=np.int32))
    assert np.array_equal(b.get("env_obs_count"), 2 * np.ones(n, dtype=np.int32))
    # the state after the probe step = the ORACLE's first step from the drawn pose with the drawn parameters
    for i in range(0, n, 7):
        oracle.reset()
        oracle.params[0] = p_ref[i]
        oracle.set_state(q_ref[i], np.zeros(18), np.zeros(12))
        oracle.step(1)
        assert int(oracle.d["ncon"]) == 0
        assert np.max(np.abs(oracle.d["qpos"] - q[i])) < 1e-9 and np.max(np.abs(oracle.d["qvel"] - v[i])) < 1e-8
    # reproducible: same seed, same episodes
    env2 = hmj.BatchedHumanoidEnv(n, randomization_factor=1.0, seed=seed)
    obs2, _ = env2.reset()
    assert np.array_equal(obs, obs2)
    assert np.array_equal(env2.batch.get_raw("env_episode_src", np.int32, (n,)), src)
    env.close(); env2.close()


def test_partial_reset_touches_nobody_else(hmj):
    n = 64
    env = hmj.BatchedHumanoidEnv(n, randomization_factor=1.0, seed=2)
    env.reset()
    rng = np.random.default_rng(0)
    for _ in range(5):
        env.step(rng.uniform(-1, 1, (n, 12)))
    names = ("qpos", "qvel", "time", "warmstart", "ctrl", "env_act_noise", "env_obs_noise", "env_act_delay", "env_act_count",
             "env_obs_count", "env_prev_action", "env_latest_action", "env_force_sched")
    before = {k: env.batch.get(k).copy() for k in names}
    p0 = env.batch.get_params().copy()
    ids = np.array([3, 17, 40])
    obs, _ = env.reset(env_ids=ids)
    assert obs.shape == (3, 30)
    keep = np.ones(n, dtype=bool); keep[ids] = False
    for k in names:
        now = env.batch.get(k)
        assert np.array_equal(now[keep], before[k][keep]), k
    assert np.array_equal(env.batch.get_params()[keep], p0[keep])
    assert np.allclose(env.batch.get("time")[ids], 0.002) and np.all(env.batch.get("ncon")[ids] == 0)
    assert not np.array_equal(env.batch.get_params()[ids], p0[ids])
    env.close()


def test_auto_reset_against_an_undisturbed_twin(hmj):
    """B200VecEnv (auto-reset inside the device step) next to a plain env without it, same seed and actions.  Envs are made
    to time out at different steps; at each ending the VecEnv must report the twin's observation as terminal_observation,
    the twin's reward / flags, a fresh first observation for the ended env — and leave every other env bit-identical."""
    n = 48
    v = hmj.B200VecEnv(n, randomization_factor=1.0, max_simulation_time_override=0.1, use_potential_rewards=True, seed=11)
    e = hmj.BatchedHumanoidEnv(n, randomization_factor=1.0, max_simulation_time_override=0.1, use_potential_rewards=True, seed=11)
    o1 = v.reset()
    o2, _ = e.reset()
    assert np.array_equal(o1, o2)
    # stagger the clocks: env i has already lived i % 12 extra steps, so endings spread over a dozen steps
    for env in (v.env, e):
        t = env.batch.get("time")
        t += (np.arange(n) % 12) * 0.002
        env.batch.set("time", t)
    alive = np.ones(n, dtype=bool)           # twin rows still comparable (the twin keeps stepping its ended episodes)
    rng = np.random.default_rng(3)
    endings = 0
    for step in range(70):
        a = rng.uniform(-1, 1, (n, 12))
        ov, rv, dv, iv = v.step(a)
        oe, re_, te, tre, _ = e.step(a)
        de = te | tre
        assert np.array_equal(dv[alive], de[alive]), step
        assert np.array_equal(rv[alive], re_[alive]), step
        run = alive & ~dv
        assert np.array_equal(ov[run], oe[run]), step          # bystanders: untouched by anybody's reset
        for i in np.nonzero(alive & dv)[0]:
            endings += 1
            assert np.array_equal(iv[i]["terminal_observation"], oe[i])
            assert iv[i]["TimeLimit.truncated"] == bool(tre[i] and not te[i])
            assert not np.array_equal(ov[i], oe[i])
            assert abs(v.env.time[i] - 0.002) < 1e-12          # new episode, one probe step in
        alive &= ~dv
        if not alive.any():
            break
    assert endings == n
    # second-generation episodes keep running and ending: the pool keeps up over a long run
    total = 0
    for step in range(200):
        ov, rv, dv, iv = v.step(rng.uniform(-1, 1, (n, 12)))
        total += int(dv.sum())
        assert np.all(np.isfinite(ov))
        for i in np.nonzero(dv)[0]:
            assert iv[i]["terminal_observation"].shape == (30,)
    assert total >= 3 * n
    v.close(); e.close()


def test_mass_ending_larger_than_the_pool(hmj):
    """All envs time out in the same step (more than the pool holds): the call still returns every env reset."""
    n = 300
    v = hmj.B200VecEnv(n, randomization_factor=1.0, max_simulation_time_override=0.02, seed=4)
    assert v.env.batch.spare_rows < n
    v.reset()
    rng = np.random.default_rng(1)
    saw = 0
    for step in range(12):
        obs, rew, dones, infos = v.step(rng.uniform(-1, 1, (n, 12)))
        if dones.all():
            saw += 1
            assert np.all(np.abs(v.env.time - 0.002) < 1e-12)
            assert np.all(rew == -100.0)
            assert all(infos[i]["terminal_observation"].shape == (30,) for i in range(n))
            assert np.all(np.isfinite(obs))
    assert saw >= 1
    v.close()
