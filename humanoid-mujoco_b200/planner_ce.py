"""Cross-entropy planner of mjpc (mjpc/planners/cross_entropy/planner.cc) on the device rollout path.

Same class shape as `SamplingPlanner` (it shares Initialize / Allocate / SetState / UpdateNominalPolicy = the reference's
`ResamplePolicy`, :314-343); what differs is the sampling distribution and the update (:164-283):

  * every candidate is the resampled nominal plus Gaussian noise whose standard deviation per spline parameter is
    max(sqrt(variance[k]), std_min) — absolute, not scaled by the control range (AddNoiseToPolicy :346-378); the nominal
    itself is rolled out alongside (NominalTrajectory, :407-409): here it is global sample 0 of the batch;
  * the n_elite best candidates (std::partial_sort, :192-201) are averaged into the new nominal (:213-236) and the
    variance is refitted (:239-252) — with the reference's own quirk kept: the inner loop reads the BEST candidate's
    parameter for every elite, so variance[k] = n_elite (p_best[k] - p_avg[k])^2 / (n_elite - 1);
  * improvement = max(avg_return - best_return, 0) (:270-272).

Candidates, rollouts and returns live on the device (hmj_plan_iterate_ex: one CUDA-graph launch for all rollouts); the
population's returns and spline values are read back once per iteration for the elite selection.  With torch.distributed
initialised every rank contributes its local top-n_elite rows and all ranks merge the same global elite set.
"""
from __future__ import annotations

import numpy as np

from .batch import NU
from .planner import SamplingPlanner, TimeSpline


class CrossEntropyPlanner(SamplingPlanner):
    """`num_trajectory` = samples of THIS rank, including (on rank 0) the un-noised nominal at local index 0."""

    def Initialize(self):
        super().Initialize()
        self.std_initial_ = float(self.model["sampling_exploration"][0])   # planner.cc:55-57
        self.std_min_ = 0.1                                                # "std_min" default, :58
        self.n_elite_ = max(self.num_trajectory_ * self.world // 10, 2)    # :64-65, over the global population

    def Allocate(self):
        super().Allocate()
        self.variance = np.full(64 * NU, self.std_initial_ ** 2)          # Reset(): planner.cc:138-140
        self.avg_return = 0.0

    def Reset(self, horizon: int, initial_repeated_action=None):
        super().Reset(horizon, initial_repeated_action)
        if hasattr(self, "variance"):
            self.variance[:] = self.std_initial_ ** 2

    def ResamplePolicy(self, horizon: int):
        self.UpdateNominalPolicy(horizon)

    def OptimizePolicy(self, horizon: int | None = None, noise: np.ndarray | None = None):
        horizon = self.steps() if horizon is None else int(horizon)
        self.ResamplePolicy(horizon)
        times, nominal = self.policy.plan.arrays()
        P = times.shape[0]
        std = np.maximum(np.sqrt(self.variance[:P * NU]), self.std_min_)
        first_global = self.rank * self.num_trajectory_
        self.batch.plan_iterate(self.state, self.time, times, nominal, horizon, self.interpolation_, noise=noise,
                                seed=self.seed, iteration=self.iteration, first_global=first_global, noise_std_vec=std)
        self.iteration += 1
        self.candidate_times = times
        self._population, self._population_nnode, self._horizon = None, P, horizon
        ret, _fail, vals = self._fetch_population()
        # local elites (the nominal, global sample 0, is not a candidate)
        gidx = first_global + np.arange(self.num_trajectory_)
        cand = gidx != 0
        k = min(self.n_elite_, int(cand.sum()))
        order = np.lexsort((gidx[cand], ret[cand]))[:k]
        rows = np.concatenate([ret[cand][order, None], gidx[cand][order, None].astype(np.float64),
                               vals[cand][order].reshape(k, P * NU)], axis=1)
        nominal_return = float(ret[0]) if first_global == 0 else np.nan
        if self.world > 1:
            from .dist import gather_rows
            pad = np.full((self.n_elite_, rows.shape[1]), np.inf)
            pad[:k] = rows
            allrows = gather_rows(pad, self._device)
            allrows = allrows[np.isfinite(allrows[:, 0])]
            nominal_return = float(gather_rows(np.array([[nominal_return]]), self._device)[0, 0])
        else:
            allrows = rows
        n_elite = min(self.n_elite_, allrows.shape[0])
        sel = np.lexsort((allrows[:, 1], allrows[:, 0]))[:n_elite]
        elite = allrows[sel]
        params = elite[:, 2:]
        p_avg = params.sum(axis=0) * (1.0 / n_elite)            # mju_scl(.., 1 / n_elite)
        self.avg_return = float(elite[:, 0].sum() / n_elite)
        var = np.zeros(P * NU)
        diff = params[0] - p_avg                                 # the reference reads the BEST candidate's node for every elite
        for _ in range(n_elite):
            var += diff * diff / (n_elite - 1)
        self.variance[:P * NU] = var
        self.previous_policy.CopyFrom(self.policy)
        plan = TimeSpline(NU, self.interpolation_)
        for t, v in zip(times, p_avg.reshape(P, NU)):
            plan.AddNode(float(t), v)
        self.policy.plan = plan
        self.nominal_return = nominal_return
        self.winner = int(elite[0, 1])
        self.winner_return = float(elite[0, 0])
        self.elite_indices = elite[:, 1].astype(np.int64)
        self.improvement = max(self.avg_return - self.winner_return, 0.0)
