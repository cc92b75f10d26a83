"""BatchedHumanoidEnv on the GPU: the CPUEnv contract, and the device-resident bookkeeping kernels
(k_env_pre / k_env_post: action noise + delay FIFO, push scheduler, observation delay FIFOs) against their
numpy restatement (`step_host`, the reference's control flow of simulation/cpu_env.py:465-693 vectorised).
Bars: bit-exact — both paths run the same physics kernels and the bookkeeping is copy / add / compare work."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N = 64


def make_env(hmj, seed=3, rf=1.0):
    env = hmj.BatchedHumanoidEnv(N, randomization_factor=rf, seed=seed)
    obs, info = env.reset(seed=seed)
    return env, obs


def test_reset_and_step_contract():
    import humanoid_mujoco_b200 as hmj

    env, obs = make_env(hmj)
    assert obs.shape == (N, 30) and info_ok(env)
    assert np.all(np.isfinite(obs))
    # reset accepts only contact-free starts (cpu_env.py:411-414)
    assert np.all(env.batch.get("ncon") == 0)
    a = np.random.default_rng(0).uniform(-1, 1, (N, 12))
    o, r, term, trunc, infos = env.step(a)
    assert o.shape == (N, 30) and r.shape == (N,) and term.shape == (N,) and trunc.shape == (N,) and len(infos) == N
    assert o.dtype == np.float64 and term.dtype == bool
    assert np.all(np.abs(env.time - 2 * 0.002) < 1e-12)  # reset's probe step + this one
    # partial reset leaves the other environments untouched
    q0 = env.batch.get_state()[0].copy()
    env.reset(env_ids=[1, 5])
    q1 = env.batch.get_state()[0]
    keep = np.ones(N, dtype=bool); keep[[1, 5]] = False
    assert np.array_equal(q0[keep], q1[keep]) and not np.array_equal(q0[~keep], q1[~keep])
    env.close()


def info_ok(env):
    return env.io["obs"].shape == (N, 30) and env.torso_quat.shape == (N, 4)


def test_device_bookkeeping_matches_host_restatement():
    import humanoid_mujoco_b200 as hmj

    dev, obs_d = make_env(hmj, seed=7)
    host, obs_h = make_env(hmj, seed=7)
    assert np.array_equal(obs_d, obs_h)
    # a push that starts almost immediately, so the scheduler, the active window and the renewal are all exercised
    for e in (dev, host):
        fs = e.batch.get("env_force_sched")
        fs[:, 0], fs[:, 1], fs[:, 2], fs[:, 3], fs[:, 4] = 0.006, 0.02, 10.0, 0.6, 0.8
        e.batch.set("env_force_sched", fs)
        e.batch.set("env_force_body", np.arange(N, dtype=np.int32) % 14 + 1)
    host._pull_state()
    assert host.action_delay.max() > 0 and host.obs_delay.max() > 0  # rf = 1: delays of 5..25 steps
    rng = np.random.default_rng(1)
    saw_push = False
    for step in range(60):
        a = rng.uniform(-1, 1, (N, 12))
        od, rd, td, ud, _ = dev.step(a)
        oh, rh, th, uh, _ = host.step_host(a)
        assert np.array_equal(od, oh), step
        assert np.array_equal(rd, rh), step
        assert np.array_equal(td, th) and np.array_equal(ud, uh), step
        assert np.array_equal(dev.torso_quat, host.torso_quat), step
        saw_push |= bool(np.any(dev.batch.get("xfrc")[:, :2] != 0))
    assert saw_push
    # bookkeeping state after 60 steps (incl. the renewed push schedule drawn from the splitmix64 stream)
    mirror = {h: getattr(host, h).copy() for _, h in host._DEVICE_STATE}
    dev._pull_state()
    for _, h in dev._DEVICE_STATE:
        assert np.array_equal(getattr(dev, h), mirror[h]), h
    assert np.all(dev.force_draws >= 1) and np.all(dev.force_sched[:, 0] > 1.0)
    dev.close(); host.close()


def test_step_none_replays_ctrl_and_potential_rewards():
    import humanoid_mujoco_b200 as hmj

    env = hmj.BatchedHumanoidEnv(N, randomization_factor=0.0, use_potential_rewards=True, seed=2)
    env.reset(seed=2)
    a = np.full((N, 12), 0.3)
    _, r1, _, _, _ = env.step(a)
    r1 = r1.copy()
    _, r2, _, _, _ = env.step(None)  # CPUEnv.step(None): data.ctrl goes through the FIFO again (cpu_env.py:660-662)
    assert np.allclose(env.batch.get("ctrl"), 0.3)  # rf = 0: no noise, no delay
    assert np.all(np.abs(r2) < np.abs(r1) + 50)      # potential reward = difference of consecutive rewards
    env.close()


def test_action_source_does_not_change_results():
    """The controls reach the device from wherever the caller keeps them (hmj_env_step_fused_from): a page-locked array
    (hmj.pinned_empty), a pageable array, a non-contiguous view and a float32 array must give bit-identical steps."""
    import humanoid_mujoco_b200 as hmj

    rng = np.random.default_rng(4)
    acts = rng.uniform(-1, 1, (6, N, 12)).astype(np.float32).astype(np.float64)   # exactly representable in float32
    runs = []
    for kind in ("pinned", "pageable", "strided", "float32"):
        env, _ = make_env(hmj, seed=9)
        pin = hmj.pinned_empty((N, 12))
        out = []
        for s in range(acts.shape[0]):
            if kind == "pinned":
                pin[...] = acts[s]
                a = pin
            elif kind == "pageable":
                a = acts[s].copy()
            elif kind == "strided":
                wide = np.zeros((N, 24))
                wide[:, ::2] = acts[s]
                a = wide[:, ::2]
            else:
                a = acts[s].astype(np.float32)
            o, r, term, trunc, infos = env.step(a)
            out.append((o.copy(), r.copy(), term.copy(), trunc.copy()))
        runs.append(out)
        env.close()
    for other in runs[1:]:
        for a, b in zip(runs[0], other):
            for x, y in zip(a, b):
                assert np.array_equal(x, y)
    with pytest.raises(ValueError):
        env2, _ = make_env(hmj, seed=9)
        env2.step(np.zeros((N, 11)))


def test_infos_are_sparse_and_flags_shared_when_nothing_ends():
    import humanoid_mujoco_b200 as hmj
    from humanoid_mujoco_b200.env import SparseInfos

    env, _ = make_env(hmj, seed=2, rf=0.0)
    o, r, term, trunc, infos = env.step(np.zeros((N, 12)))
    assert isinstance(infos, SparseInfos) and len(infos) == N and infos[0] == {} and infos[-1] == {}
    assert not term.any() and not trunc.any()
    with pytest.raises(ValueError):
        term[0] = True            # the all-clear flags are one shared read-only array
    env.max_simulation_time = 0.0  # every episode is over at the next step: terminated everywhere (cpu_env.py:684)
    o, r, term, trunc, infos = env.step(np.zeros((N, 12)))
    assert term.all() and all("is_success" in infos[i] for i in range(N))
    assert [i for i in infos][3] == infos[3] and infos[1:3] == [infos[1], infos[2]]
    env.close()
