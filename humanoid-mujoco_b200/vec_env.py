"""`B200VecEnv` — the `DummyVecEnv([lambda: CPUEnv(...)] * n)` seam of the reference (rl/train.py:134-151,
simulation/benchmark.py:37-39) over ONE batched device env.

stable_baselines3 is not a dependency of this repo (it is absent from the image), so this class restates the part of
SB3's `VecEnv` / `DummyVecEnv` contract the reference touches, with the same names and semantics:

  * `reset() -> obs [n, 30]`;
  * `step_async(actions)` / `step_wait()` / `step(actions) -> (obs, rewards, dones, infos)` with AUTO-RESET: an env whose
    episode ended (`terminated or truncated`) is reset inside the same call, its last observation is handed out as
    `infos[i]["terminal_observation"]`, `infos[i]["TimeLimit.truncated"] = truncated and not terminated`, and the returned
    row of `obs` is the first observation of the new episode (DummyVecEnv.step_wait);
  * `set_attr / get_attr / env_method` (rl/randomization_adaptation_callback.py:53 sets `randomization_factor`);
  * `num_envs`, `observation_space`, `action_space` (Box(-10, 10, (30,)) / Box(-1, 1, (12,)), cpu_env.py:56-63).

A maintainer who has SB3 installed subclasses `stable_baselines3.common.vec_env.VecEnv` with these methods unchanged.
"""
from __future__ import annotations

from types import SimpleNamespace

import numpy as np

from .env import BatchedHumanoidEnv


def _box(low, high, shape):
    return SimpleNamespace(low=np.full(shape, low, dtype=np.float64), high=np.full(shape, high, dtype=np.float64),
                           shape=tuple(shape), dtype=np.float64)


class B200VecEnv:
    def __init__(self, n_envs: int, randomization_factor: float = 0.0, use_potential_rewards: bool = False,
                 max_simulation_time_override=None, reward_override=None, device: int = 0, seed: int = 0):
        self.env = BatchedHumanoidEnv(n_envs, randomization_factor=randomization_factor,
                                      use_potential_rewards=use_potential_rewards,
                                      max_simulation_time_override=max_simulation_time_override,
                                      reward_override=reward_override, device=device, seed=seed)
        self.num_envs = int(n_envs)
        self.env.auto_reset = self.env.device_reset   # auto-reset inside the device step (episode pool)
        self.observation_space = _box(-10.0, 10.0, (30,))
        self.action_space = _box(-1.0, 1.0, (12,))
        self._actions = None
        self._running_info = {"TimeLimit.truncated": False}
        self.reset_infos = [{} for _ in range(self.num_envs)]

    # ------------------------------------------------------------------ VecEnv contract
    def reset(self):
        obs, _ = self.env.reset()
        self.reset_infos = [{} for _ in range(self.num_envs)]
        return obs.copy()

    def step_async(self, actions):
        self._actions = np.asarray(actions, dtype=np.float64).reshape(self.num_envs, 12)

    def step_wait(self):
        obs, reward, term, trunc, raw = self.env.step(self._actions)
        obs = np.array(obs, copy=True)
        reward = np.array(reward, copy=True)
        dones = np.asarray(term | trunc)
        # envs still running share one `{"TimeLimit.truncated": False}` (treat it as read-only); ended envs get their own dict
        infos = [self._running_info] * self.num_envs
        ended = np.nonzero(dones)[0]
        if len(ended):
            if self.env.auto_reset:   # the device already swapped in the next episodes: obs rows are their first observations
                tid, tobs = self.env.last_terminal
                where = {int(e): k for k, e in enumerate(tid)}
                for i in ended:
                    d = dict(raw[i])
                    d["TimeLimit.truncated"] = bool(trunc[i] and not term[i])
                    d["terminal_observation"] = tobs[where[int(i)]].copy()
                    infos[i] = d
            else:
                for i in ended:
                    d = dict(raw[i])
                    d["TimeLimit.truncated"] = bool(trunc[i] and not term[i])
                    d["terminal_observation"] = obs[i].copy()
                    infos[i] = d
                new_obs, _ = self.env.reset(env_ids=ended)
                obs[ended] = new_obs
        return obs, reward, dones, infos

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def close(self):
        self.env.close()

    # ------------------------------------------------------------------ attribute access (VecEnv.get_attr / set_attr / env_method)
    def get_attr(self, attr_name, indices=None):
        value = getattr(self.env, attr_name)
        ids = range(self.num_envs) if indices is None else np.atleast_1d(indices)
        if isinstance(value, np.ndarray) and value.shape[:1] == (self.num_envs,):
            return [value[i] for i in ids]
        return [value for _ in ids]

    def set_attr(self, attr_name, value, indices=None):
        """The batched env holds one value per attribute (the reference sets the same value on every env:
        randomization_adaptation_callback.py:53); a per-index subset is therefore refused instead of silently widened."""
        if indices is not None and len(np.atleast_1d(indices)) != self.num_envs:
            raise ValueError("B200VecEnv.set_attr: attributes are shared by all environments of the batch")
        setattr(self.env, attr_name, value)

    def env_method(self, method_name, *args, indices=None, **kwargs):
        out = getattr(self.env, method_name)(*args, **kwargs)
        ids = range(self.num_envs) if indices is None else np.atleast_1d(indices)
        return [out for _ in ids]

    def seed(self, seed=None):
        if seed is not None:
            self.env.seed = int(seed)
            self.env.rng = np.random.default_rng(seed)
        return [seed] * self.num_envs
