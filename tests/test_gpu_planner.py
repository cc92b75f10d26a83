"""GPU parity of the predictive-sampling rollouts against the CPU oracle.

Bar (BASELINE.json north_star): the planner's chosen sample index is identical whenever cost gaps exceed
1e-9; per-sample returns and recorded trajectories agree within the 10-step / 1e-6 trajectory tolerance
scaled to the horizon (a rollout is H-1 chained mj_steps).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL_RETURN = 1e-6     # relative, per-sample total_return
TOL_TRAJ = 1e-6       # absolute on states of an H=18 rollout
GAP = 1e-9            # winner must be identical when the two best returns differ by more than this


@pytest.fixture(scope="module")
def hmj():
    import humanoid_mujoco_b200 as h

    return h


def settled_task_state(oracle, timestep=0.02, nsteps=50):
    """C2/C4 start: qpos0, qvel = 0, 50 zero-control steps at the agent timestep."""
    oracle.reset()
    oracle.lib.orc_default_params(oracle.model.ctypes.data, oracle.params.ctypes.data)
    oracle.params["timestep"][0] = timestep
    oracle.set_state(np.array(oracle.model["qpos0"][0]), np.zeros(18), np.zeros(12))
    oracle.step(nsteps)
    assert oracle.d["status"] == 0
    return np.concatenate([oracle.d["qpos"], oracle.d["qvel"]]).copy(), float(oracle.d["time"])


def oracle_returns(oracle, state, time, times, values, horizon, interp):
    ret = np.zeros(values.shape[0])
    fail = np.zeros(values.shape[0], dtype=bool)
    for i in range(values.shape[0]):
        r = oracle.rollout(state, time, times, values[i], horizon, interp)
        ret[i], fail[i] = r["total_return"], r["failure"]
    return ret, fail


def argmin_with_gap(ret):
    order = np.argsort(ret, kind="stable")
    gap = ret[order[1]] - ret[order[0]] if ret.size > 1 else np.inf
    return int(order[0]), float(gap)


def test_c2_rollouts_match_oracle(hmj, oracle):
    """Config C2: 10 rollouts x 0.35 s horizon (H = 18), cubic 3-node spline, sigma = 0.05."""
    state, time = settled_task_state(oracle)
    pl = hmj.SamplingPlanner(num_trajectory=10, device=0, seed=0)
    assert pl.steps() == 18 and pl.num_spline_points == 3
    pl.SetState(state[:19], state[19:], time)
    H = pl.steps()
    pl.UpdateNominalPolicy(H)
    pl.Rollouts(H, record=True)
    times, values = pl.candidate_times, pl.candidate_values
    assert np.array_equal(values[0], pl.policy.plan.arrays()[1])  # sample 0 = un-noised nominal (planner.cc:361)
    ref, ref_fail = oracle_returns(oracle, state, time, times, values, H, pl.interpolation_)
    assert not ref_fail.any() and not pl.failure.any()
    rel = np.max(np.abs(pl.total_return - ref) / np.maximum(np.abs(ref), 1.0))
    print("C2 return rel err", rel)
    assert rel < TOL_RETURN
    cost, idx = pl.batch.plan_argmin()
    want, gap = argmin_with_gap(ref)
    if gap > GAP:
        assert idx == want
    assert cost == pl.total_return[idx]
    # recorded buffers of every sample (Trajectory fields, trajectory.cc:100-210)
    for i in (0, idx, 9):
        got = pl.batch.plan_fetch(i, H)
        r = oracle.rollout(state, time, times, values[i], H, pl.interpolation_)
        assert np.max(np.abs(got["states"] - r["states"])) < TOL_TRAJ
        assert np.max(np.abs(got["actions"] - r["actions"])) < 1e-12
        assert np.array_equal(got["actions"][H - 1], got["actions"][H - 2])  # trajectory.cc:190-192
        assert np.max(np.abs(got["times"] - r["times"])) < 1e-12
        assert np.max(np.abs(got["residual"] - r["residual"])) < 1e-5
        assert np.max(np.abs(got["costs"] - r["costs"]) / np.maximum(np.abs(r["costs"]), 1.0)) < TOL_RETURN
        assert np.max(np.abs(got["trace"] - r["trace"])) < TOL_TRAJ


def test_planner_iterations_choose_the_oracle_winner(hmj, oracle):
    """Five OptimizePolicy iterations with injected noise; a host mirror whose rollouts run on the oracle must pick
    the same winner each iteration (when the gap allows) and end with the same nominal spline."""
    state, time = settled_task_state(oracle)
    n = 24
    pl = hmj.SamplingPlanner(num_trajectory=n, device=0, seed=3)
    pl.SetState(state[:19], state[19:], time)
    H = pl.steps()
    rng = np.random.default_rng(7)
    scale = 0.5 * (pl.ctrlrange[:, 1] - pl.ctrlrange[:, 0])
    ref_policy = hmj.SamplingPolicy(pl.ctrlrange, pl.num_spline_points, pl.interpolation_)
    ref_policy.Reset(H)
    decided = 0
    for it in range(5):
        noise = rng.standard_normal((n, pl.num_spline_points, 12)) * scale * 0.05
        pl.OptimizePolicy(H, noise=noise)
        # mirror: same host logic, oracle rollouts
        P = pl.num_spline_points
        shift = max((H - 1) * pl.timestep / (P - 1), 1e-5)
        # nominal_time accumulates by repeated addition in the reference (planner.cc:262-287)
        acc, tt = time, []
        for _ in range(P):
            tt.append(acc)
            acc += shift
        node_t = np.array(tt)
        nominal = np.stack([ref_policy.Action(t) for t in node_t])
        vals = np.broadcast_to(nominal, (n,) + nominal.shape).copy() + noise
        vals[0] = nominal
        np.clip(vals, pl.ctrlrange[:, 0], pl.ctrlrange[:, 1], out=vals)
        ref, _ = oracle_returns(oracle, state, time, node_t, vals, H, pl.interpolation_)
        want, gap = argmin_with_gap(ref)
        # A rollout is 17 chained contact-rich steps: a make/break-contact event turns a last-bit difference into a
        # visible one for an occasional sample (tools/dev/rollout_fidelity.py: the error is bimodal, ~1e-14 or > 1e-6).
        # The bar is therefore: the typical sample agrees to 1e-9, at most one sample per iteration is off, and the
        # winner is the oracle's whenever every sample agrees and the gap decides.
        rel = np.abs(pl.total_return - ref) / np.maximum(np.abs(ref), 1.0)
        agree = rel < TOL_RETURN
        assert np.median(rel) < 1e-9 and agree.sum() >= n - 1, (it, rel.max(), int((~agree).sum()))
        if gap > GAP and agree.all():
            decided += 1
            assert pl.winner == want, (it, pl.winner, want, gap)
            assert np.array_equal(pl.policy.plan.arrays()[1], vals[want])
        # the mirror follows the device's choice so that later iterations stay comparable
        plan = hmj.TimeSpline(12, pl.interpolation_)
        for t, v in zip(node_t, vals[pl.winner]):
            plan.AddNode(float(t), v)
        ref_policy.plan = plan
        # advance the plant one agent step with the planned action, as the agent loop does
        a = pl.ActionFromPolicy(time)
        assert np.max(np.abs(a - ref_policy.Action(time))) < 1e-12
        oracle.reset()
        oracle.params["timestep"][0] = pl.timestep
        oracle.set_state(state[:19], state[19:], a, time=time)
        oracle.step(1)
        state = np.concatenate([oracle.d["qpos"], oracle.d["qvel"]]).copy()
        time = float(oracle.d["time"])
        pl.SetState(state[:19], state[19:], time)
    assert decided >= 3


def test_c4_slice_properties(hmj, oracle):
    """Per-GPU slice of C4 (1024 samples x H = 100): properties that hold at full size without a full oracle run.
    Duplicated splines give bit-identical returns wherever they sit in the batch (no cross-sample coupling), sample 0
    is the nominal, a 16-sample subset matches the oracle, and the device argmin is the first minimum."""
    state, time = settled_task_state(oracle)
    n, H = 1024, 100
    pl = hmj.SamplingPlanner(num_trajectory=n, device=0, seed=5)
    pl.SetState(state[:19], state[19:], time)
    pl.UpdateNominalPolicy(H)
    rng = np.random.default_rng(9)
    scale = 0.5 * (pl.ctrlrange[:, 1] - pl.ctrlrange[:, 0])
    noise = rng.standard_normal((n, 3, 12)) * scale * 0.05
    noise[511] = 0.0                          # -> samples 511 and 512 are copies of nominal sample 0
    noise[512:] = noise[:512][::-1]          # second half mirrors the first: sample 512+k == sample 511-k
    pl.Rollouts(H, noise=noise)
    ret = pl.total_return.copy()
    assert np.array_equal(ret[1:511], ret[513:1023][::-1])   # sample 512+k == sample 511-k (1023 pairs with the un-noised 0)
    assert ret[0] == ret[511] == ret[512]
    pick = np.array([0, 1, 2, 3, 100, 200, 300, 400, 500, 511, 600, 700, 800, 900, 1000, 1023])
    ref, ref_fail = oracle_returns(oracle, state, time, pl.candidate_times, pl.candidate_values[pick], H, pl.interpolation_)
    good = ~ref_fail
    rel = np.max(np.abs(ret[pick][good] - ref[good]) / np.maximum(np.abs(ref[good]), 1.0))
    print("C4 slice (H=100) return rel err", rel, "failures", int(ref_fail.sum()))
    # 99 chained contact-rich steps at dt = 0.02: make/break-contact events amplify the 1e-9 one-step difference, so the
    # full-horizon return is only loosely comparable; the tight bar is applied to the 10-step prefix below.
    assert rel < 2e-2
    assert np.array_equal(pl.failure[pick], ref_fail)
    pl.Rollouts(11, noise=noise)    # same 1024 splines, 10 steps + final forward
    ref10, _ = oracle_returns(oracle, state, time, pl.candidate_times, pl.candidate_values[pick], 11, pl.interpolation_)
    rel10 = np.max(np.abs(pl.total_return[pick] - ref10) / np.maximum(np.abs(ref10), 1.0))
    print("C4 slice (10 steps) return rel err", rel10)
    assert rel10 < TOL_RETURN
    pl.Rollouts(H, noise=noise)
    assert np.array_equal(pl.total_return, ret)   # re-running the same population reproduces the same bits
    cost, idx = pl.batch.plan_argmin()
    assert idx == int(np.argmin(ret)) and cost == ret[idx]


def test_rollout_edge_cases(hmj, oracle):
    """H = 1 (forward only), zero-order and linear interpolation, failing rollout -> kMaxReturn-style failure flag."""
    state, time = settled_task_state(oracle)
    pl = hmj.SamplingPlanner(num_trajectory=4, device=0, seed=1)
    pl.SetState(state[:19], state[19:], time)
    for interp in (0, 1, 2):
        pl.interpolation_ = interp
        pl.policy.plan.SetInterpolation(interp)
        for H in (1, 2, 7):
            pl.UpdateNominalPolicy(max(H, 2))
            pl.Rollouts(H)
            ref, ref_fail = oracle_returns(oracle, state, time, pl.candidate_times, pl.candidate_values, H, interp)
            assert np.array_equal(pl.failure, ref_fail)
            assert np.max(np.abs(pl.total_return - ref) / np.maximum(np.abs(ref), 1.0)) < TOL_RETURN, (interp, H)
    # divergent state: huge velocity makes mj_checkAcc / checkPos fire -> failure in both
    bad = state.copy()
    bad[19:] = 1e9
    pl.SetState(bad[:19], bad[19:], time)
    pl.UpdateNominalPolicy(5)
    pl.Rollouts(5)
    ref, ref_fail = oracle_returns(oracle, bad, time, pl.candidate_times, pl.candidate_values, 5, 2)
    assert np.array_equal(pl.failure, ref_fail)


def test_device_iteration_equals_host_built_candidates(hmj, oracle):
    """hmj_plan_iterate (candidates = clamp(nominal + noise) built on the device, rollouts replayed as one CUDA graph,
    device arg-min + winner record) against the host-built path (`Rollouts` + `plan_argmin`): same bits."""
    state, time = settled_task_state(oracle)
    n = 64
    pl = hmj.SamplingPlanner(num_trajectory=n, device=0, seed=11)
    pl.SetState(state[:19], state[19:], time)
    H = pl.steps()
    rng = np.random.default_rng(21)
    scale = 0.5 * (pl.ctrlrange[:, 1] - pl.ctrlrange[:, 0])
    for it in range(3):     # iteration 0 runs un-captured, 1 captures the graph, 2 replays it
        pl.UpdateNominalPolicy(H)
        noise = rng.standard_normal((n, 3, 12)) * scale * (0.05 if it < 2 else 5.0)   # last one saturates the clamp
        pl.Rollouts(H, noise=noise)
        ret_host, vals_host = pl.total_return.copy(), pl.candidate_values.copy()
        cost_h, idx_h = pl.batch.plan_argmin()
        l0 = pl.batch.host_launches
        pl.OptimizePolicyCandidates(1, H, noise=noise)
        host_launches = pl.batch.host_launches - l0
        assert np.array_equal(pl.candidate_values, vals_host)
        assert np.array_equal(pl.total_return, ret_host)
        assert pl.winner == idx_h and pl.winner_return == cost_h and pl.nominal_return == ret_host[0]
        assert np.array_equal(pl._winner_values, vals_host[idx_h])
        if it > 0:
            assert host_launches <= 4, host_launches     # candidates, ONE graph launch for all rollouts, winner


def test_device_noise_stream(hmj, oracle):
    """Device-drawn exploration noise (counter-based, Box-Muller): sample 0 is the nominal, values are clamped, the
    perturbations have the requested scale, differ between iterations and do not depend on how samples are sharded."""
    state, time = settled_task_state(oracle)
    n = 512
    pl = hmj.SamplingPlanner(num_trajectory=n, device=0, seed=5)
    pl.SetState(state[:19], state[19:], time)
    pl.policy.Reset(0, np.zeros(12))
    pl.UpdateNominalPolicy(3)
    _, nominal = pl.policy.plan.arrays()
    pl.OptimizePolicyCandidates(1, 3)
    v0 = pl.candidate_values.copy()
    assert np.array_equal(v0[0], nominal)
    assert v0.min() >= pl.ctrlrange[:, 0].min() and v0.max() <= pl.ctrlrange[:, 1].max()
    scale = 0.5 * (pl.ctrlrange[:, 1] - pl.ctrlrange[:, 0])
    z = (v0[1:] - nominal) / (scale * pl.noise_exploration[0])
    assert abs(z.mean()) < 0.02 and abs(z.std() - 1.0) < 0.02, (z.mean(), z.std())
    pl.policy.Reset(0, np.zeros(12))
    pl.UpdateNominalPolicy(3)
    pl.OptimizePolicyCandidates(1, 3)
    assert not np.array_equal(pl.candidate_values[1:], v0[1:])           # next iteration, new draw
    # a second planner owning global samples [256, 512) draws exactly the second half of the first iteration
    p2 = hmj.SamplingPlanner(num_trajectory=256, device=0, seed=5)
    p2.rank, p2.world = 1, 1          # shard offset only; no process group in this test
    p2.SetState(state[:19], state[19:], time)
    p2.policy.Reset(0, np.zeros(12))
    p2.UpdateNominalPolicy(3)
    p2.OptimizePolicyCandidates(1, 3)
    assert np.array_equal(p2.candidate_values, v0[256:])


def test_device_select_matches_host_select(hmj):
    """k_plan_select (the reduction behind the all-gather) against dist.select_winner on fabricated records: ties go to
    the lowest global index, NaN never wins, the nominal return is rank 0's."""
    import torch

    from humanoid_mujoco_b200.dist import select_winner

    b = hmj.HumanoidBatch(4, device=0)
    P, R = 3, 39
    rng = np.random.default_rng(3)
    for case in range(6):
        W = (2, 8, 8, 8, 33, 3)[case]
        recs = rng.standard_normal((W, R))
        recs[:, 0] = rng.uniform(1, 2, W)
        recs[:, 1] = np.arange(W) * 1024 + rng.integers(0, 1024, W)
        if case == 2:
            recs[5, 0] = recs[2, 0] = 0.5
        if case == 3:
            recs[0, 0] = np.nan
            recs[4, 0] = np.nan
        if case == 5:
            recs[:, 0] = np.nan
        g = torch.tensor(recs, dtype=torch.float64, device="cuda:0")
        torch.cuda.synchronize()
        got = b.plan_select(P, g.data_ptr(), W)
        want = select_winner(recs)
        assert np.array_equal(got, want, equal_nan=True), case
    b.close()


def test_cross_entropy_planner_matches_mirror(hmj, oracle):
    """CrossEntropyPlanner (mjpc/planners/cross_entropy/planner.cc:164-283): three iterations with injected noise against
    a host mirror whose rollouts run on the oracle — same elite set, elite mean as the new nominal, the reference's
    variance formula (best candidate's parameter for every elite), improvement = avg - best."""
    state, time = settled_task_state(oracle)
    n = 41                                     # nominal + 40 candidates -> n_elite = 4
    pl = hmj.CrossEntropyPlanner(num_trajectory=n, device=0, seed=3)
    assert pl.n_elite_ == 4 and pl.std_min_ == 0.1
    pl.SetState(state[:19], state[19:], time)
    H = pl.steps()
    rng = np.random.default_rng(17)
    variance = np.full(36, pl.std_initial_ ** 2)
    ref_policy = hmj.SamplingPolicy(pl.ctrlrange, pl.num_spline_points, pl.interpolation_)
    ref_policy.Reset(H)
    compared = 0
    for it in range(3):
        std = np.maximum(np.sqrt(variance), pl.std_min_)
        noise = rng.standard_normal((n, 3, 12)) * std.reshape(3, 12)
        pl.OptimizePolicy(H, noise=noise)
        # mirror
        P = 3
        shift = max((H - 1) * pl.timestep / (P - 1), 1e-5)
        acc, tt = time, []
        for _ in range(P):
            tt.append(acc)
            acc += shift
        node_t = np.array(tt)
        nominal = np.stack([ref_policy.Action(t) for t in node_t])
        vals = nominal[None] + noise
        vals[0] = nominal
        np.clip(vals, pl.ctrlrange[:, 0], pl.ctrlrange[:, 1], out=vals)
        assert np.array_equal(pl.candidate_values, vals)
        ref, _ = oracle_returns(oracle, state, time, node_t, vals, H, pl.interpolation_)
        rel = np.abs(pl.total_return - ref) / np.maximum(np.abs(ref), 1.0)
        # std_min = 0.1 rad of absolute noise makes these rollouts rougher than predictive sampling's (5 % of the range):
        # the bimodal error of tools/dev/rollout_fidelity.py (1e-14 or a contact event) shows up in a few samples more
        assert np.median(rel) < 1e-9 and (rel < TOL_RETURN).sum() >= n - 3
        order = 1 + np.argsort(ref[1:], kind="stable")[:4]
        gaps = np.diff(np.sort(ref[1:])[:5])
        # the elite set is the oracle's whenever the samples that decide it (the five best on either side) agree
        decisive = np.union1d(order, 1 + np.argsort(pl.total_return[1:], kind="stable")[:5])
        if (rel[decisive] < 1e-9).all() and gaps.min() > GAP:
            compared += 1
            assert np.array_equal(pl.elite_indices, order)
        el = pl.elite_indices                       # follow the device's elite set so later iterations stay comparable
        p_avg = vals[el].reshape(4, 36).sum(axis=0) * 0.25
        diff = vals[el[0]].reshape(36) - p_avg
        var = np.zeros(36)
        for _ in range(4):
            var += diff * diff / 3
        assert np.max(np.abs(pl.policy.plan.arrays()[1].reshape(36) - p_avg)) < 1e-15
        assert np.max(np.abs(pl.variance[:36] - var)) < 1e-18
        assert abs(pl.avg_return - pl.total_return[el].sum() / 4) < 1e-12
        assert pl.improvement == max(pl.avg_return - pl.total_return[el[0]], 0.0)
        assert pl.nominal_return == pl.total_return[0]
        variance = var
        plan = hmj.TimeSpline(12, pl.interpolation_)
        for t, v in zip(node_t, p_avg.reshape(3, 12)):
            plan.AddNode(float(t), v)
        ref_policy.plan = plan
    assert compared >= 2
    # device-drawn noise uses the per-parameter std (no control-range scaling)
    pl.variance[:36] = 0.04
    pl.OptimizePolicy(H)
    z = (pl.candidate_values[1:] - pl.candidate_values[0]).reshape(-1, 36)
    inside = np.abs(pl.candidate_values[1:]).max() < pl.ctrlrange[:, 1].max()
    assert 0.15 < z.std() < 0.25 if inside else True


def test_noisy_rollouts_and_robust_planner(hmj, oracle):
    """Trajectory::NoisyRollout (OU wrench on every body, trajectory.cc:147-155) against the oracle with injected normals,
    then RobustPlanner.OptimizePolicy (robust_planner.cc:96-151): the candidate with the best mean noisy return wins."""
    state, time = settled_task_state(oracle)
    n, H = 24, 18
    pl = hmj.SamplingPlanner(num_trajectory=n, device=0, seed=6)
    pl.SetState(state[:19], state[19:], time)
    rp = hmj.RobustPlanner(pl, ncandidates=4, nrepetitions=3, xfrc_std=0.5, xfrc_rate=0.1)
    rng = np.random.default_rng(8)
    scale = 0.5 * (pl.ctrlrange[:, 1] - pl.ctrlrange[:, 0])
    noise = rng.standard_normal((n, 3, 12)) * scale * 0.05
    xnoise = rng.standard_normal((12, H - 1, 90))
    rp.OptimizePolicy(H, noise=noise, xfrc_noise=xnoise)
    # the four candidates are the delegate's four best; each noisy rollout equals the oracle's
    ret = pl.total_return
    assert np.array_equal(rp.candidates, np.argsort(ret, kind="stable")[:4])
    ref = np.zeros(12)
    for k in range(12):
        c = rp.candidates[k // 3]
        ref[k], fail = oracle.rollout_noisy(state, time, pl.candidate_times, pl.candidate_values[c], H, pl.interpolation_,
                                            0.5, 0.1, xnoise[k])
        assert not fail
    rel = np.abs(rp.noisy_return - ref) / np.maximum(np.abs(ref), 1.0)
    # 17 steps of a pushed robot in contact amplify rounding: a rollout may cross one of the reference algorithm's own
    # discontinuities (profiles/r2_step_discontinuity_probe.txt).  Bar: the bulk agrees to 1e-9, and
    # every rollout that does not agree to TOL_RETURN must be one where the ORACLE ITSELF answers differently when its
    # start state moves by one unit in the last place.
    off = np.nonzero(rel >= TOL_RETURN)[0]
    assert np.median(rel) < 1e-9 and len(off) <= 3, rel
    for k in off:
        c = rp.candidates[k // 3]
        moved = 0.0
        for dof in (2, 7, 10, 13):
            for side in (-np.inf, np.inf):
                st = state.copy()
                st[dof] = np.nextafter(st[dof], side)
                r2, _ = oracle.rollout_noisy(st, time, pl.candidate_times, pl.candidate_values[c], H, pl.interpolation_,
                                             0.5, 0.1, xnoise[k])
                moved = max(moved, abs(r2 - ref[k]) / max(abs(ref[k]), 1.0))
        assert moved >= TOL_RETURN, (k, rel[k], moved)
    # the wrench matters (the same candidate under three different noise draws gives three different returns)
    assert len(np.unique(rp.noisy_return[:3])) == 3 and abs(rp.noisy_return[0] - ret[rp.candidates[0]]) > 1e-9
    want = int(np.argmin(ref.reshape(4, 3).mean(axis=1)))
    gap = np.diff(np.sort(ref.reshape(4, 3).mean(axis=1)))[0]
    if (rel < 1e-9).all() and gap > GAP:
        assert rp.best_candidate == want
    assert np.allclose(rp.scores, rp.noisy_return.reshape(4, 3).mean(axis=1), rtol=1e-12)
    assert np.array_equal(pl.policy.plan.arrays()[1], pl.candidate_values[rp.candidates[rp.best_candidate]])
    # device-drawn OU noise: runs, differs between calls, finite
    a, fa = rp.batch.plan_rollouts_noisy(state, time, pl.candidate_times, np.repeat(pl.candidate_values[:4], 3, axis=0), H, 2, 0.5, 0.1, None, seed=1)
    b2, _ = rp.batch.plan_rollouts_noisy(state, time, pl.candidate_times, np.repeat(pl.candidate_values[:4], 3, axis=0), H, 2, 0.5, 0.1, None, seed=1)
    assert np.all(np.isfinite(a)) and not fa.any() and not np.array_equal(a, b2)
    rp.close()
