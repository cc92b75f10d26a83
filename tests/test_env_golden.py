"""The CPUEnv reward / observation epilogue against fixtures produced by the REFERENCE'S OWN Python
(tests/golden/make_env_golden.py: simulation/reward_functions.py standupReward + cpu_env.py inverseRotateVectors,
imported unmodified with scipy's Rotation standing in for jax.scipy's).

  * not gpu: the numpy host mirror (env.py: standup_reward_host / observation_host) and the compiled model's
    actuator -> qpos / dof maps against the fixture;
  * gpu: k_env_obs_reward through the C ABI (hmj_env_obs_reward) on the fixture's states.
Bar: fp64 vs scipy fp64 — 1e-12 on reward / observation, flags exact."""
import os

import numpy as np
import pytest

TOL = 1e-12
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "env_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLD))


def test_model_joint_maps_follow_reference_joint_names(gold):
    """joint_qpos_idx / joint_dof_idx (cpu_env.py:111-122, JOINT_NAMES order) == the compiled model's actuator maps."""
    from model_layout import load_model

    m = load_model()
    assert np.array_equal(np.array(m["act_qposadr"][0]), gold["joint_qpos_idx"])
    assert np.array_equal(np.array(m["act_dofadr"][0]), gold["joint_dof_idx"])


def test_host_mirror_matches_reference_reward_and_observation(gold):
    from humanoid_mujoco_b200.env import local_gravity_host, observation_host, standup_reward_host

    jq, jd = gold["joint_qpos_idx"], gold["joint_dof_idx"]
    assert gold["reward"].shape[0] >= 2000
    g = local_gravity_host(gold["quat"])
    assert np.max(np.abs(g - gold["local_gravity"])) < TOL
    tau = (gold["qfrc_constraint"] + gold["qfrc_smooth"])[:, jd]
    r, term, trunc = standup_reward_host(gold["qvel"], gold["quat"], gold["z"], tau, gold["prev_action"], gold["latest_action"],
                                         gold["self_collision"], gold["time"], float(gold["max_time"]))
    assert np.array_equal(term, gold["terminated"]) and np.array_equal(trunc, gold["truncated"])
    assert np.max(np.abs(r - gold["reward"])) < TOL * 100   # rewards reach 35: relative 1e-12
    assert term.any() and (~term).any() and trunc.any() and gold["self_collision"].any()
    obs = observation_host(gold["qpos"], gold["qvel"], gold["quat"], gold["noise"], jq, jd)
    assert np.max(np.abs(obs - gold["obs"])) < TOL


@pytest.mark.gpu
def test_device_epilogue_matches_reference_reward_and_observation(gold):
    """k_env_obs_reward reads qpos/qvel/time, the one-step-stale xquat[1] / xpos[1].z ("sens"), qfrc_constraint +
    qfrc_smooth and the contact list; all of them are uploaded from the fixture, so the kernel sees exactly the values
    the reference function was called with."""
    import humanoid_mujoco_b200 as hmj
    from humanoid_mujoco_b200.batch import CONTACT_DTYPE, MAXCON

    n = gold["reward"].shape[0]
    b = hmj.HumanoidBatch(n, device=0)
    b.set_state(gold["qpos"], gold["qvel"], gold["time"], np.zeros((n, 18)))
    sens = np.zeros((n, 16))
    sens[:, 3:7] = gold["quat"]     # SENS_ROOT_XQUAT
    sens[:, 7] = gold["z"]          # SENS_ROOT_Z
    b.set("sens", sens)
    b.set("qfrc_constraint", gold["qfrc_constraint"])
    b.set("qfrc_smooth", gold["qfrc_smooth"])
    # contact lists: self-colliding envs hold a floor contact followed by a mesh-mesh one, the others 0..2 floor contacts
    con = np.zeros((n, MAXCON), dtype=CONTACT_DTYPE)
    ncon = (np.arange(n) % 3).astype(np.int32)
    con["g1"][:, :] = 0
    con["g2"][:, :] = 9
    selfc = gold["self_collision"]
    ncon[selfc] = 2
    con["g1"][selfc, 1] = 4
    con["g2"][selfc, 1] = 6
    b.set("ncon", ncon)
    b.set_raw("contact", con)
    obs, rew, term, trunc, sc = b.obs_reward(gold["prev_action"], gold["latest_action"], gold["noise"], float(gold["max_time"]))
    assert np.array_equal(term, gold["terminated"])
    assert np.array_equal(trunc, gold["truncated"])
    assert np.array_equal(sc, selfc)
    err_r = np.max(np.abs(rew - gold["reward"]))
    err_o = np.max(np.abs(obs - gold["obs"]))
    print("device epilogue vs reference: reward", err_r, "obs", err_o)
    assert err_r < TOL * 100
    assert err_o < TOL
    # noise = NULL must equal zero noise
    zero = gold["noise"][:, 0] == 0
    obs0 = b.obs_reward(gold["prev_action"], gold["latest_action"], None, float(gold["max_time"]))[0]
    assert np.max(np.abs(obs0[zero] - gold["obs"][zero])) < TOL
    b.close()
