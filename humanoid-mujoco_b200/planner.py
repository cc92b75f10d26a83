"""Host side of mjpc's predictive-sampling planner with GPU rollouts.

Mirrors the reference classes and method names so a maintainer can map them one to one:

  TimeSpline      <- mjpc/spline/spline.{h,cc}        (Sample :103-156, DiscardBefore :164-188,
                                                       AddNode :203-238, Slope :259-277)
  SamplingPolicy  <- mjpc/planners/sampling/policy.cc (Action :52-59, Reset :43-50, CopyFrom :61-64)
  SamplingPlanner <- mjpc/planners/sampling/planner.cc (Initialize :40-75, OptimizePolicy :190-208,
                     UpdateNominalPolicy :236-310, AddNoiseToPolicy :313-339, Rollouts :342-380,
                     CopyCandidateToPolicy :525-534, ActionFromPolicy :225-233)

What changed: `Rollouts` no longer fans out over a ThreadPool (mjpc/threadpool.cc) — it packs the
candidate splines and makes ONE call into the C ABI (hmj_plan_rollouts), where each sample is a
warp of a CUDA grid and spline evaluation, residual and cost accumulation are fused into the
step kernels.  The static kMaxTrajectory = 128 cap (planners/planner.h:28) is gone: the sample
count is the batch size.  With torch.distributed initialised, samples are sharded over ranks and
the winner is agreed by an all-gather arg-min followed by a broadcast of the winning spline.
"""
from __future__ import annotations

import bisect

import numpy as np

from .batch import DEFINES, NU, HumanoidBatch

K_ZERO, K_LINEAR, K_CUBIC = 0, 1, 2
K_MAX_TRAJECTORY_HORIZON = 512  # mjpc/trajectory.h:27
K_MAX_RETURN = 1.0e6


class TimeSpline:
    """Time-indexed spline of `dim` values (zero / linear / cubic-Hermite with finite-difference slopes)."""

    def __init__(self, dim: int = 0, interpolation: int = K_ZERO):
        self.dim = int(dim)
        self.interpolation = int(interpolation)
        self.times: list[float] = []
        self.values: list[np.ndarray] = []

    # -- container API (subset used by the planner)
    def Size(self) -> int:
        return len(self.times)

    def Dim(self) -> int:
        return self.dim

    def SetInterpolation(self, interpolation: int):
        self.interpolation = int(interpolation)

    def Interpolation(self) -> int:
        return self.interpolation

    def Clear(self):
        self.times, self.values = [], []

    def Reserve(self, num_nodes: int):  # storage is dynamic; kept for interface parity
        return None

    def AddNode(self, time: float, values=None) -> np.ndarray:
        v = np.zeros(self.dim) if values is None else np.array(values, dtype=np.float64).reshape(self.dim)
        if not self.times or time > self.times[-1]:
            self.times.append(float(time)); self.values.append(v)
        elif time < self.times[0]:
            self.times.insert(0, float(time)); self.values.insert(0, v)
        else:
            raise ValueError("Adding nodes to the middle of the spline isn't supported.")
        return v

    def NodeAt(self, index: int):
        return self.times[index], self.values[index]

    def DiscardBefore(self, time: float) -> int:
        last = bisect.bisect_right(self.times, time)
        if last == 0:
            return 0
        keep = 1 if self.interpolation == K_CUBIC else 0
        last -= 1
        while last != 0 and keep:
            last -= 1
            keep -= 1
        del self.times[:last]
        del self.values[:last]
        return last

    def _slope(self, i: int) -> np.ndarray:
        t, v = self.times, self.values
        if i == 0:
            return (v[1] - v[0]) / (t[1] - t[0])
        if i == len(t) - 1:
            return (v[i] - v[i - 1]) / (t[i] - t[i - 1])
        return 0.5 * (v[i + 1] - v[i]) / (t[i + 1] - t[i]) + 0.5 * (v[i] - v[i - 1]) / (t[i] - t[i - 1])

    def Sample(self, time: float) -> np.ndarray:
        if not self.times:
            return np.zeros(self.dim)
        upper = bisect.bisect_right(self.times, time)
        if upper == len(self.times):
            return self.values[-1].copy()
        if upper == 0:
            return self.values[0].copy()
        lower = upper - 1
        t = (time - self.times[lower]) / (self.times[upper] - self.times[lower])
        lo, up = self.values[lower], self.values[upper]
        if self.interpolation == K_ZERO:
            return lo.copy()
        if self.interpolation == K_LINEAR:
            return lo * (1 - t) + up * t
        dt = self.times[upper] - self.times[lower]
        c0 = 2.0 * t * t * t - 3.0 * t * t + 1.0
        c1 = (t * t * t - 2.0 * t * t + t) * dt
        c2 = -2.0 * t * t * t + 3 * t * t
        c3 = (t * t * t - t * t) * dt
        return c0 * lo + c1 * self._slope(lower) + c2 * up + c3 * self._slope(upper)

    def copy(self) -> "TimeSpline":
        s = TimeSpline(self.dim, self.interpolation)
        s.times = list(self.times)
        s.values = [v.copy() for v in self.values]
        return s

    def arrays(self):
        return np.array(self.times, dtype=np.float64), np.array(self.values, dtype=np.float64).reshape(len(self.times), self.dim)


class SamplingPolicy:
    def __init__(self, ctrlrange: np.ndarray, num_spline_points: int = 3, interpolation: int = K_CUBIC):
        self.ctrlrange = np.asarray(ctrlrange, dtype=np.float64)
        self.num_spline_points = int(num_spline_points)
        self.plan = TimeSpline(NU, interpolation)

    def Reset(self, horizon: int = 0, initial_repeated_action=None):
        self.plan.Clear()
        if initial_repeated_action is not None:
            self.plan.AddNode(0.0, initial_repeated_action)

    def Action(self, time: float) -> np.ndarray:
        a = self.plan.Sample(time)
        return np.clip(a, self.ctrlrange[:, 0], self.ctrlrange[:, 1])  # Clamp, utilities.cc:94-98

    def CopyFrom(self, policy: "SamplingPolicy", horizon: int = 0):
        self.plan = policy.plan.copy()
        self.num_spline_points = policy.num_spline_points


class SamplingPlanner:
    """Predictive sampling.  `num_trajectory` is the number of samples evaluated by THIS rank."""

    def __init__(self, num_trajectory: int | None = None, device: int = 0, seed: int = 0, model=None):
        self.batch = None
        self._device = device
        self._model_override = model
        self._requested = num_trajectory
        self.rank, self.world = 0, 1
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                self.rank, self.world = dist.get_rank(), dist.get_world_size()
        except Exception:
            pass
        # every rank draws its own candidates: the per-rank noise stream is (seed, rank), so equal seeds on all ranks
        # do not duplicate samples
        self.rng = np.random.default_rng([int(seed), self.rank])
        self.seed, self.iteration = int(seed), 0    # device noise stream: (seed, iteration, GLOBAL sample index)
        self._rec_bufs = None
        self._population, self._population_nnode = None, None
        self.Initialize()
        self.Allocate()
        self.Reset(0)

    # -- Initialize / Allocate / Reset ------------------------------------------------------------
    def Initialize(self):
        from .batch import load_model
        m = self._model_override if self._model_override is not None else load_model()
        self.model = m
        self.noise_exploration = [float(m["sampling_exploration"][0]), 0.0]
        self.num_trajectory_ = int(self._requested or m["sampling_trajectories"][0])
        self.interpolation_ = K_CUBIC  # sampling_representation default (planner.cc:65)
        self.sliding_plan_ = 0
        self.num_spline_points = int(m["sampling_spline_points"][0])
        self.timestep = float(m["agent_timestep"][0])   # Agent::PlanIteration writes opt.timestep (agent.cc:275)
        self.agent_horizon = float(m["agent_horizon"][0])
        self.ctrlrange = np.array(m["act_ctrlrange"][0], dtype=np.float64)
        self.winner = 0

    def Allocate(self):
        self.batch = HumanoidBatch(self.num_trajectory_, self._device, self.model)
        p = self.batch.default_params()
        p["timestep"] = self.timestep
        self.batch.set_params(p)
        self.state = np.zeros(DEFINES["HMJ_NSTATE"])
        self.time = 0.0
        self.policy = SamplingPolicy(self.ctrlrange, self.num_spline_points, self.interpolation_)
        self.previous_policy = SamplingPolicy(self.ctrlrange, self.num_spline_points, self.interpolation_)
        self.total_return = np.zeros(self.num_trajectory_)
        self.failure = np.zeros(self.num_trajectory_, dtype=bool)
        self.candidate_values = None
        self.winner = -1

    def Reset(self, horizon: int, initial_repeated_action=None):
        self.state[:] = 0.0
        self.time = 0.0
        self.policy.Reset(horizon, initial_repeated_action)
        self.previous_policy.Reset(horizon, initial_repeated_action)
        self.improvement = 0.0
        self.winner = 0

    def SetState(self, qpos, qvel, time: float):
        self.state[: DEFINES["HMJ_NQ"]] = qpos
        self.state[DEFINES["HMJ_NQ"]:] = qvel
        self.time = float(time)

    def steps(self) -> int:
        """Agent::PlanIteration: steps_ = max(min(horizon/timestep + 1, 512), 1) with int truncation (agent.cc:279-280)."""
        return int(max(min(self.agent_horizon / self.timestep + 1, K_MAX_TRAJECTORY_HORIZON), 1))

    # -- one planning iteration -------------------------------------------------------------------
    def OptimizePolicy(self, horizon: int | None = None, noise: np.ndarray | None = None):
        horizon = self.steps() if horizon is None else int(horizon)
        self.UpdateNominalPolicy(horizon)
        self.OptimizePolicyCandidates(1, horizon, noise)
        self.CopyCandidateToPolicy(0)
        best_return = self.nominal_return
        self.improvement = max(best_return - self.winner_return, 0.0)

    def UpdateNominalPolicy(self, horizon: int):
        """Resample the current plan onto a fresh grid of spline points starting now (non-sliding plan)."""
        P = self.num_spline_points
        nominal_time = self.time
        time_horizon = (horizon - 1) * self.timestep
        if self.interpolation_ == K_ZERO:
            time_shift = max(time_horizon / P, 1.0e-5)
        else:
            time_shift = max(time_horizon / (P - 1), 1.0e-5)
        scratch = TimeSpline(NU, self.interpolation_)
        for _ in range(P):
            scratch.AddNode(nominal_time, self.policy.Action(nominal_time))
            nominal_time += time_shift
        self.policy.plan = scratch

    def AddNoiseToPolicy(self, values: np.ndarray, noise: np.ndarray | None):
        """values [n][P][12] (copies of the nominal); sample 0 of the GLOBAL population stays un-noised."""
        n, P, _ = values.shape
        scale = 0.5 * (self.ctrlrange[:, 1] - self.ctrlrange[:, 0])
        if noise is None:
            std = np.full(n, self.noise_exploration[0])
            if self.noise_exploration[1] > 0:
                std = np.where(self.rng.random(n) < 0.2, self.noise_exploration[1], std)
            noise = self.rng.standard_normal((n, P, NU)) * (scale[None, None, :] * std[:, None, None])
        noise = np.array(noise, dtype=np.float64, copy=True)
        if self.rank == 0:
            noise[0] = 0.0
        out = values + noise
        np.clip(out, self.ctrlrange[:, 0], self.ctrlrange[:, 1], out=out)
        return out

    def Rollouts(self, horizon: int, noise: np.ndarray | None = None, record: bool = False):
        times, nominal = self.policy.plan.arrays()
        values = np.broadcast_to(nominal, (self.num_trajectory_,) + nominal.shape).copy()
        values = self.AddNoiseToPolicy(values, noise)
        self.candidate_times, self.candidate_values = times, values
        ret, fail = self.batch.plan_rollouts(self.state, self.time, times, values, horizon, self.interpolation_, record)
        self.total_return, self.failure = ret, fail   # host copies (the setters switch off the lazy device read-back)
        self._horizon = horizon

    def OptimizePolicyCandidates(self, ncandidates: int, horizon: int, noise: np.ndarray | None = None):
        """Candidates, rollouts and arg-min on the device (hmj_plan_iterate: one CUDA-graph launch for all rollouts), then
        the winner exchange.  `noise` [n][P][12] injects the perturbations (parity tests); None draws them on the device
        from the counter-based stream (seed, iteration, global sample index)."""
        times, nominal = self.policy.plan.arrays()
        P = times.shape[0]
        first_global = self.rank * self.num_trajectory_
        rec_local = rec_all = None
        if self.world > 1:
            import torch
            import torch.distributed as dist
            if dist.get_backend() == "nccl":
                R = 3 + P * NU
                if self._rec_bufs is None or self._rec_bufs[0].numel() != R:
                    dev = torch.device("cuda", self._device)
                    self._rec_bufs = (torch.zeros(R, dtype=torch.float64, device=dev),
                                      torch.zeros((self.world, R), dtype=torch.float64, device=dev))
                rec_local, rec_all = self._rec_bufs
        self.batch.plan_iterate(self.state, self.time, times, nominal, horizon, self.interpolation_, noise=noise,
                                noise_std=self.noise_exploration[0], noise_std2=self.noise_exploration[1], seed=self.seed,
                                iteration=self.iteration, first_global=first_global,
                                record_dev=rec_local.data_ptr() if rec_local is not None else None)
        self.iteration += 1
        self.candidate_times = times
        self._population = None          # fetched on demand (total_return / failure / candidate_values)
        self._population_nnode = P
        self._horizon = horizon
        if rec_all is not None:
            from .dist import exchange_records
            rec = exchange_records(None, self.batch, rec_local, rec_all)
        else:
            rec = self.batch.plan_select(P)
            if self.world > 1:
                from .dist import exchange_records
                rec = exchange_records(rec)
        self.winner_return = float(rec[0])
        self.winner = int(rec[1])
        self.nominal_return = float(rec[2])
        self._winner_values = rec[3:].reshape(P, NU).copy()
        return 1

    # population of the last iteration, read back only when somebody asks (the reference keeps trajectory[i].total_return)
    def _fetch_population(self):
        if self._population is None:
            self._population = self.batch.plan_population(self._population_nnode)
        return self._population

    @property
    def total_return(self):
        return self._total_return if self._population_nnode is None else self._fetch_population()[0]

    @total_return.setter
    def total_return(self, v):
        self._total_return, self._population_nnode = v, None

    @property
    def failure(self):
        return self._failure if self._population_nnode is None else self._fetch_population()[1]

    @failure.setter
    def failure(self, v):
        self._failure = v

    @property
    def candidate_values(self):
        return self._candidate_values if self._population_nnode is None else self._fetch_population()[2]

    @candidate_values.setter
    def candidate_values(self, v):
        self._candidate_values = v

    def CopyCandidateToPolicy(self, candidate: int = 0):
        self.previous_policy.CopyFrom(self.policy)
        plan = TimeSpline(NU, self.interpolation_)
        for t, v in zip(self.candidate_times, self._winner_values):
            plan.AddNode(float(t), v)
        self.policy.plan = plan

    def ActionFromPolicy(self, time: float, use_previous: bool = False) -> np.ndarray:
        return (self.previous_policy if use_previous else self.policy).Action(time)

    def BestTrajectory(self):
        """Full buffers of the local winner (requires a recorded rollout)."""
        if self.winner < 0:
            return None
        local = self.winner - self.rank * self.num_trajectory_
        if not (0 <= local < self.num_trajectory_):
            return None
        return self.batch.plan_fetch(local, self._horizon)
