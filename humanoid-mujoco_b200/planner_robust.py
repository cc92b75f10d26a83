"""mjpc's RobustPlanner (mjpc/planners/robust/robust_planner.cc) over the device rollout path.

The reference wraps a delegate planner: the delegate proposes its `ncandidates` best candidates (`OptimizePolicyCandidates`,
sampling/planner.cc:113-187), every candidate is rolled out `nrepetitions` more times under an Ornstein-Uhlenbeck wrench on every
body (`Trajectory::NoisyRollout`, trajectory.cc:147-155), and the candidate with the best mean return over its valid noisy rollouts
becomes the policy (robust_planner.cc:96-151; a candidate none of whose rollouts is valid keeps the delegate's score).  Here the
delegate is `SamplingPlanner` (its iteration runs on the device), the top candidates come from the population read-back, and the
ncandidates x nrepetitions noisy rollouts are ONE batched call (hmj_plan_rollouts_noisy) on a second, small batch.
"""
from __future__ import annotations

import numpy as np

from .batch import NU, HumanoidBatch
from .planner import SamplingPlanner, TimeSpline


class RobustPlanner:
    def __init__(self, delegate: SamplingPlanner, ncandidates: int = 5, nrepetitions: int = 5, xfrc_std: float = 0.1,
                 xfrc_rate: float = 0.1):
        if delegate.world != 1:
            raise ValueError("RobustPlanner: the robustness pass is single-rank (the delegate may still be sharded upstream)")
        self.delegate_ = delegate
        self.ncandidates_, self.nrepetitions_ = int(ncandidates), int(nrepetitions)   # robust_planner.cc:52-57 defaults
        self.xfrc_std_, self.xfrc_rate_ = float(xfrc_std), float(xfrc_rate)
        self.batch = HumanoidBatch(self.ncandidates_ * self.nrepetitions_, delegate._device, delegate.model)
        p = self.batch.default_params()
        p["timestep"] = delegate.timestep
        self.batch.set_params(p)
        self.best_candidate, self.scores = -1, None

    def SetState(self, qpos, qvel, time):
        self.delegate_.SetState(qpos, qvel, time)

    def OptimizePolicy(self, horizon: int | None = None, noise=None, xfrc_noise=None):
        d = self.delegate_
        horizon = d.steps() if horizon is None else int(horizon)
        d.UpdateNominalPolicy(horizon)
        d.OptimizePolicyCandidates(self.ncandidates_, horizon, noise)
        ret, vals = d.total_return, d.candidate_values
        ncand = min(self.ncandidates_, ret.shape[0])
        order = np.argsort(ret, kind="stable")[:ncand]              # the delegate's partial_sort, ties to the lowest index
        self.candidates = order
        if ncand == 1:
            d._winner_values, d.winner = vals[order[0]].copy(), int(order[0])
            d.CopyCandidateToPolicy(0)
            self.best_candidate = 0
            return
        R = self.nrepetitions_
        cand_vals = np.repeat(vals[order], R, axis=0)               # trajectories_[R * i + j] rolls candidate i
        if cand_vals.shape[0] != self.batch.n:                      # fewer candidates than planned: pad with copies (ignored)
            pad = np.repeat(cand_vals[-1:], self.batch.n - cand_vals.shape[0], axis=0)
            cand_vals = np.concatenate([cand_vals, pad])
        nret, nfail = self.batch.plan_rollouts_noisy(d.state, d.time, d.candidate_times, cand_vals, horizon, d.interpolation_,
                                                     self.xfrc_std_, self.xfrc_rate_, xfrc_noise, seed=d.seed)
        self.noisy_return, self.noisy_failure = nret, nfail
        best, best_score, scores = -1, 0.0, []
        for c in range(ncand):
            mean_return, valid = float(ret[order[c]]), 0            # CandidateScore (robust_planner.cc:131)
            for j in range(R):
                if nfail[R * c + j]:
                    continue
                mean_return = (valid * mean_return + float(nret[R * c + j])) / (valid + 1)
                valid += 1
            scores.append(mean_return)
            if best == -1 or mean_return < best_score:
                best, best_score = c, mean_return
        self.best_candidate, self.scores = best, np.array(scores)
        d._winner_values, d.winner, d.winner_return = vals[order[best]].copy(), int(order[best]), float(ret[order[best]])
        d.CopyCandidateToPolicy(best)

    def ActionFromPolicy(self, time: float, use_previous: bool = False):
        return self.delegate_.ActionFromPolicy(time, use_previous)

    def close(self):
        self.batch.close()
