#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path: batched humanoid env-steps/s (BASELINE.json metric).

Workload at N GPUs (weak scaling, one process per GPU): BASELINE.json configs[2] ("C3") per GPU —
4096 lock-step humanoid environments with rf = 1 domain randomisation (cpu_env.py:189-280 ranges),
contact-rich start (standup pose + 300 zero-control settle steps), uniform random controls,
dt = 0.002, one mj_step + observation/reward epilogue per env-step.

  value   env-steps/s with inputs resident in HBM (controls streamed from a device ring larger than L2),
          K steps timed with CUDA events on the engine's stream, max over ranks;
  e2e     the same metric through the public API (BatchedHumanoidEnv.step): host actions in pinned
          memory -> H2D -> step -> D2H of obs/reward/flags every step, wall clock incl. host logic;
  roofline / fp64   dominant kernel's achieved algorithmic bytes/s (2530 B/env-step) and
          whole-step achieved fp64 flop/s (oracle counters) against measured peaks;
  cpu_baseline      the oracle (CPU restatement of the reference's MuJoCo path, kind "port": no MuJoCo
          binary exists on this box) on all host cores over a bounded sample of the same workload;
  plan    MJPC predictive-sampling iterations/s (configs[1] and the per-GPU slice of configs[3]).

`--impl reference` times the reference-path CPU implementation (oracle port) on the same workload.
"""
from __future__ import annotations

import argparse
import gc
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
for _p in (REPO, os.path.join(REPO, "tools")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "batched humanoid env-steps/sec"
UNIT = "env-steps/s"
BYTES_PER_ENV_STEP = 2530  # algorithmic HBM bytes per env-step with per-env DR params (C3)
SETTLE_STEPS = 300
ENVS_PER_GPU = 4096
CPU_SAMPLE_ENVS = 1024


def workload_name(envs, n_gpus):
    return (f"C3: {envs} lockstep humanoid envs/GPU x {n_gpus} GPU, DR rf=1 (cpu_env.py:189-280 ranges), starts through "
            f"reset's no-contact rejection, {SETTLE_STEPS} zero-control settle steps, then iid U(-1,1) controls at their "
            f"steady state (150 untimed lead-in steps), dt=0.002; second regime C3-settled = 0.02 x U(-1,1) in `workloads`")


def load_peaks():
    try:
        with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("timestamp,index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, indices):
        """`indices`: the GPUs of the job (one nvidia-smi process on rank 0 watches all of them); empty = no sampling."""
        self.index = ",".join(str(i) for i in indices)
        self.proc, self.path, self.t_region = None, None, None

    def wait_first(self, timeout=3.0):
        """Block until nvidia-smi has printed its first poll (its start-up takes longer than a short timed region)."""
        t_end = time.time() + timeout
        while self.proc is not None and time.time() < t_end:
            try:
                if os.path.getsize(self.path) > 0:
                    return
            except OSError:
                return
            time.sleep(0.02)

    def mark(self):
        """Start of the timed region: rows stamped before this moment are dropped (the sampler is started ahead of the
        warm-up so that nvidia-smi's start-up latency does not eat the region)."""
        self.t_region = time.time()

    def start(self):
        if not self.index:
            return
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.f = open(self.path, "w")
            import shutil
            pre = ["stdbuf", "-oL"] if shutil.which("stdbuf") else []   # line-buffered rows (stdio buffers fully into a file)
            self.proc = subprocess.Popen(pre + ["nvidia-smi", "-i", self.index, f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        try:
            self.proc.terminate()
            self.proc.wait(timeout=5)
        except Exception:
            pass
        try:
            self.f.close()
            rows = [[c.strip() for c in r.split(",")] for r in open(self.path).read().strip().splitlines() if r.count(",") >= 9]
            os.unlink(self.path)

            def stamp(r):
                import datetime
                try:
                    return datetime.datetime.strptime(r[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                except ValueError:
                    return 0.0

            inside = [r for r in rows if self.t_region is None or stamp(r) >= self.t_region - 0.05]
            note = None
            if not inside and rows:   # region shorter than one poll of all GPUs: the last poll stands in
                last = stamp(rows[-1])
                inside = [r for r in rows if stamp(r) >= last - 0.05]
                note = "no poll fell inside the timed region; nearest poll reported"
            rows = [r[1:] for r in inside]
            sm = [float(r[1]) for r in rows]
            reasons = set()
            for r in rows:
                for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7), ("sw_power_cap", 8)):
                    if "Active" in r[col] and "Not" not in r[col]:
                        reasons.add(name)
            if rows:
                out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(rows[0][2]), "reasons": sorted(reasons),
                       "samples": len(rows), "power_w_max": max(float(r[3]) for r in rows)}
                if note:
                    out["note"] = note
        except Exception:
            pass
        return out


# ---------------------------------------------------------------------------------------------- workload
# Two control regimes on the SAME environments (same DR draw, same accepted start poses, same settle):
#   C3          iid U(-1,1) position targets every 2 ms (BASELINE.json configs[2] as specified): the robots thrash, the
#               steady state has ~0.5 contacts per env (limits dominate the constraint rows);
#   C3-settled  0.02 x U(-1,1): the robots stay on the floor, ~3.5 contacts / ~36 constraint rows / ~4 Newton iterations
#               per env — the regime a policy that actually stands up lives in (zero controls: 4.4 / 48 / 2.9).
# `pre` untimed steps take each regime to its steady state, so that a 20-step and a 200-step timed region sample the
# same contact statistics (the driver runs --steps 20).
WORKLOADS = (("C3", 1.0, 150), ("C3-settled", 0.02, 50))


def make_workload(base_params, qpos0, n, seed, probe):
    """Per-env DR parameters and start poses of C3 (numpy; shared by both arms), through CPUEnv.reset's rejection loop
    (cpu_env.py:411-414): `probe(params, qpos) -> bool[n]` runs the first zero-control step on the arm's own engine and
    reports the envs that touch anything; those are re-drawn (model and pose, as the reference's recursive reset does).
    Contacts are bit-exact between the engines, so both arms accept the same draws."""
    from humanoid_mujoco_b200.env import randomize_params, standup_start_state
    rng = np.random.default_rng(seed)
    params = randomize_params(np.repeat(base_params[:1], n), rng, 1.0)
    qpos = standup_start_state(qpos0, rng, 1.0, n)
    attempts = 1
    for _ in range(100):
        bad = np.nonzero(probe(params, qpos))[0]
        if len(bad) == 0:
            break
        attempts += 1
        params[bad] = randomize_params(np.repeat(base_params[:1], len(bad)), rng, 1.0)
        qpos[bad] = standup_start_state(qpos0, rng, 1.0, len(bad))
    else:
        raise RuntimeError("make_workload: start poses still in contact after 100 attempts")
    return params, qpos, attempts


def control_ring(rank, entries, n):
    """[entries][n][12] iid U(-1,1): a prefix of the same numpy stream whatever `entries` is."""
    return np.random.default_rng(2 + rank).uniform(-1, 1, (entries, n, 12))


def contact_stats(ncon, nefc, niter, status=None):
    out = {"mean_ncon": float(np.mean(ncon)), "max_ncon": int(np.max(ncon)), "frac_no_contact": float(np.mean(np.asarray(ncon) == 0)),
           "mean_nefc": float(np.mean(nefc)), "max_nefc": int(np.max(nefc)), "mean_newton_iter": float(np.mean(niter)),
           "ncon_hist": np.bincount(np.minimum(ncon, 16), minlength=17).tolist(),
           "nefc_hist_by_10": np.bincount(np.minimum(np.asarray(nefc) // 10, 12), minlength=13).tolist(),
           "niter_hist": np.bincount(np.minimum(niter, 12), minlength=13).tolist()}
    if status is not None:
        out["envs_with_bad_status"] = int((np.asarray(status) != 0).sum())
    return out


# ---------------------------------------------------------------------------------------------- reference arm / CPU baseline
def _oracle_batch(n):
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    from oracle_py import OracleBatch  # bench.py's cpu_baseline / --impl reference legs may execute oracle/
    return OracleBatch(n)


def cpu_steps_per_s(params, qpos, qvel, ws, ctrl_ring, nsteps, warm, nthreads):
    """Oracle (CPU port of the reference path) on `nthreads` host threads; returns (env-steps/s, stats, batch)."""
    n = qpos.shape[0]
    ob = _oracle_batch(n)
    ob.params[:] = params
    ob.set_state(qpos, qvel, ws)
    d = ob.data
    if warm:
        ob.step(ctrl_ring, warm, nthreads)
    fl0 = (d["flops_smooth"] + d["flops_collision"] + d["flops_constraint"] + d["flops_solver"]).copy()
    start = contact_stats(d["ncon"], d["nefc"], d["solver_niter"])
    secs = ob.step(ctrl_ring, nsteps, nthreads)
    fl = d["flops_smooth"] + d["flops_collision"] + d["flops_constraint"] + d["flops_solver"] - fl0
    end = contact_stats(d["ncon"], d["nefc"], d["solver_niter"], d["status"])
    stats = {"mean_ncon": end["mean_ncon"], "mean_nefc": end["mean_nefc"], "mean_niter": end["mean_newton_iter"],
             "flops_per_step": float(fl.mean() / max(1, nsteps)), "bad_status": end["envs_with_bad_status"],
             "seconds": secs, "contacts_start": {k: start[k] for k in ("mean_ncon", "mean_nefc", "mean_newton_iter")}}
    return n * nsteps / secs, stats, ob


REF_SUBSTEPS = 16   # one reference-arm "step" = 16 mj_steps of the 1024-env sample (16384 env-steps; the GPU step is 4096)


def run_reference(args):
    """The reference path's CPU implementation (oracle port: MuJoCo itself is not available here) on all host threads, on
    the first 1024 of the SAME 4096 envs the GPU arm builds with seed 1 (same DR draws, same accepted starts, same
    settle, same controls stream, same steady-state lead-in), both control regimes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    nthreads = os.cpu_count() or 1
    n_all, m = args.envs, min(CPU_SAMPLE_ENVS, args.envs)
    ob_all = _oracle_batch(n_all)

    def probe(params, qpos):
        ob_all.params[:] = params
        ob_all.set_state(qpos, np.zeros((n_all, 18)), np.zeros((n_all, 18)), np.zeros(n_all))
        ob_all.data["ctrl"][:] = 0.0
        ob_all.step(None, 1, nthreads)
        return ob_all.data["ncon"] > 0

    t0 = time.perf_counter()
    params, qpos, attempts = make_workload(ob_all.params, ob_all.model["qpos0"][0], n_all, 1, probe)
    ob = _oracle_batch(m)
    ob.params[:] = params[:m]
    ob.set_state(qpos[:m], np.zeros((m, 18)), np.zeros((m, 18)), np.zeros(m))
    ob.data["ctrl"][:] = 0.0
    ob.step(None, SETTLE_STEPS, nthreads)
    settled = (ob.data["qpos"].copy(), ob.data["qvel"].copy(), ob.data["qacc_warmstart"].copy())
    ring = control_ring(0, 64, n_all)[:, :m].copy()
    res = {}
    for name, scale, pre in WORKLOADS:
        value, stats, _ = cpu_steps_per_s(params[:m], *settled, ring * scale, args.steps * REF_SUBSTEPS,
                                          pre + args.warmup * REF_SUBSTEPS, nthreads)
        res[name] = (value, stats)
    wall = time.perf_counter() - t0
    value, stats = res["C3"]
    sample = (f"first {m} of the {n_all} envs of C3 (seed 1: same DR draws / accepted starts / settle / control stream as the "
              f"GPU arm), {args.steps} x {REF_SUBSTEPS} mj_steps each after the steady-state lead-in")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * m * REF_SUBSTEPS / value, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args.envs, args.gpus),
                   "step": f"{REF_SUBSTEPS} mj_steps over the {m}-env sample ({m * REF_SUBSTEPS} env-steps)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": nthreads, "kind": "port", "sample": sample, **stats},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "workloads": {k: {"value": v, "unit": UNIT, **st} for k, (v, st) in res.items()},
        "reset_attempts": attempts,
        "note": ("MuJoCo 3.1.4 (the reference's physics) is not in /root/reference nor installed on this box: the timed CPU "
                 "implementation is the repo's fp64 C restatement (oracle/, -O3 -march=native -ffp-contract=off), all host threads"),
        "wall_s": wall,
    }
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------- planner (metric M2)
def plan_start_state(device=0, timestep=0.02, nsteps=50):
    """C2/C4 start: qpos0, qvel = 0, 50 zero-control steps at the agent timestep, produced on the GPU."""
    import humanoid_mujoco_b200 as hmj
    b = hmj.HumanoidBatch(1, device=device)
    p = b.default_params()
    p["timestep"] = timestep
    b.set_params(p)
    b.set_state(np.array(b.model["qpos0"][0])[None], np.zeros((1, 18)), np.zeros(1), np.zeros((1, 18)))
    b.step(np.zeros((1, 12)), nsteps)
    q, v, t, _ = b.get_state()
    b.close()
    return q[0].copy(), v[0].copy(), float(t[0])


def fd_bench(local, cpu):
    """Finite-differenced transition matrices for the 18 knots of the C2 horizon (iLQG's ModelDerivatives::Compute,
    mjpc/planners/model_derivatives.cc:51-96): ONE batched step of 18 x 97 envs (hmj_fd_transition, forward differences)
    against the oracle's mjd_transitionFD restatement, one knot after the other on one host thread."""
    import humanoid_mujoco_b200 as hmj
    q, v, t = plan_start_state(local)
    T = 18
    b = hmj.HumanoidBatch(97 * T, device=local)
    p = b.default_params()
    p["timestep"] = 0.02
    b.set_params(p)
    rng = np.random.default_rng(0)
    x = np.tile(np.concatenate([q, v]), (T, 1))
    x[:, 7:19] += rng.normal(0, 0.02, (T, 12))
    u = rng.normal(0, 0.05, (T, 12))
    tt = t + 0.02 * np.arange(T)
    for _ in range(3):
        b.fd_transition(x, u, tt)
    reps = 20
    t0 = time.perf_counter()
    for _ in range(reps):
        A, _B, _C, _D = b.fd_transition(x, u, tt)
    dt = (time.perf_counter() - t0) / reps
    out = {"knots": T, "ms_per_call": 1e3 * dt, "mj_steps_per_call": T * 49, "api": "hmj_fd_transition (host arrays in, A/B/C/D out)"}
    if cpu:
        sys.path.insert(0, os.path.join(REPO, "oracle"))
        from oracle_py import Oracle   # cpu_baseline leg
        o = Oracle()
        o.params["timestep"][0] = 0.02
        t0 = time.perf_counter()
        for k in range(T):
            Ar, _, _, _ = o.transition_fd(x[k], u[k], tt[k])
        cdt = time.perf_counter() - t0
        out["cpu_baseline"] = {"ms_per_call": 1e3 * cdt, "cores": 1, "kind": "port",
                               "max_rel_diff_A_last_knot": float(np.max(np.abs(A[T - 1] - Ar) / np.maximum(np.abs(Ar), 1.0)))}
    b.close()
    return out


def plan_bench(local, rank, world, iters, cpu):
    """MJPC predictive-sampling iterations/s: one iteration = SamplingPlanner.OptimizePolicy (resample nominal, noise,
    N rollouts of H-1 steps + 1 forward with fused policy/residual/cost, device argmin, winner exchange over the
    process group, policy update).  C2: 10 x H=18 on one GPU; C4: 8192/8 = 1024 samples per GPU x H=100."""
    import torch
    import humanoid_mujoco_b200 as hmj
    q, v, t = plan_start_state(local)
    out = {}
    for name, n_local, horizon, nit in (("C2", 10, 18, 5 * iters), ("C4_slice", 1024, 100, iters)):
        if name == "C2" and (world > 1 and rank > 0):
            continue
        pl = hmj.SamplingPlanner(num_trajectory=n_local, device=local, seed=100 + rank)
        if name == "C2":
            pl.rank, pl.world = 0, 1   # the reference's own 10-rollout case is single-GPU by definition
        pl.SetState(q, v, t)
        for _ in range(3):
            pl.OptimizePolicy(horizon)
        l0, h0 = pl.batch.kernel_launches, pl.batch.host_launches
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(nit):
            pl.OptimizePolicy(horizon)
        pl.batch.sync()
        dt = time.perf_counter() - t0
        n_glob = n_local * (pl.world if name != "C2" else 1)
        out[name] = {"iterations_per_s": nit / dt, "ms_per_iteration": 1e3 * dt / nit, "samples": n_glob, "horizon": horizon,
                     "sample_steps_per_s": n_glob * (horizon - 1) * nit / dt, "iterations": nit,
                     "kernel_launches_per_iteration": (pl.batch.host_launches - h0) / nit,
                     "kernel_nodes_per_iteration": (pl.batch.kernel_launches - l0) / nit,
                     "rollouts": "one CUDA-graph launch per iteration (candidates, rollouts, arg-min on the device)",
                     "winner": int(pl.winner), "winner_return": float(pl.winner_return),
                     "exchange": "none" if n_glob == n_local else "one all-gather of the winner records {cost, global index, nominal return, spline} over NCCL on the batch stream + device selection (hmj_plan_select)"}
        if cpu and rank == 0:
            sys.path.insert(0, os.path.join(REPO, "oracle"))
            from oracle_py import OracleBatch   # cpu_baseline leg: the oracle is the timed CPU implementation
            nthreads = os.cpu_count() or 1
            m = n_local if name == "C2" else 4 * nthreads
            ob = OracleBatch(1)
            ob.params["timestep"] = pl.timestep
            state = np.concatenate([q, v])
            vals = pl.candidate_values[:m] if pl.candidate_values.shape[0] >= m else np.resize(pl.candidate_values, (m,) + pl.candidate_values.shape[1:])
            ob.rollouts(state, t, pl.candidate_times, vals, horizon, pl.interpolation_, nthreads)
            reps = 20 if name == "C2" else 1
            secs = sum(ob.rollouts(state, t, pl.candidate_times, vals, horizon, pl.interpolation_, nthreads)[2] for _ in range(reps)) / reps
            cpu_it = (m / secs) / n_glob if name != "C2" else 1.0 / secs
            out[name]["cpu_baseline"] = {"iterations_per_s": cpu_it, "cores": nthreads, "kind": "port",
                                         "sample": f"{m} rollouts x H={horizon} on {nthreads} threads"
                                                   + ("" if name == "C2" else f", scaled to {n_glob} samples per iteration")}
        pl.batch.close()
    return out

# ---------------------------------------------------------------------------------------------- C5 / C1 (BASELINE.json configs[4], [0])
def c5_bench(local, rank, world, steps, barrier, max_over_ranks):
    """rl/train.py rollout collection (rl/train.py:134-151): 16384 envs in total (STRONG scaling: 16384 / world per GPU),
    rf = 0 (RANDOMIZATION_FACTOR_INIT), random actions, DummyVecEnv semantics with the auto-reset inside the timed region.
    Two numbers: `BatchedHumanoidEnv.step` with auto_reset (the C-ABI call: host actions in, obs/reward/flags out) and
    `B200VecEnv.step` (adds SB3's per-step copies and info dicts).  Plus the auto-reset cost itself: the same envs with a
    0.2 s time limit and staggered clocks, so that ~1 % of the envs end in EVERY step, against the run where none ends."""
    import humanoid_mujoco_b200 as hmj
    total = 16384
    n = total // world
    out = {"envs_total": total, "envs_per_gpu": n, "steps": steps, "scaling": "strong", "randomization_factor": 0.0}
    acts = hmj.pinned_empty((16, n, 12))
    acts[...] = np.random.default_rng(40 + rank).uniform(-1, 1, acts.shape)
    v = hmj.B200VecEnv(n, randomization_factor=0.0, device=local, seed=20 + rank)
    t0 = time.perf_counter()
    v.reset()
    out["reset_all_ms"] = 1e3 * max_over_ranks(time.perf_counter() - t0)
    env = v.env
    for s in range(10):
        env.step(acts[s % 16])
    barrier()
    ended = 0
    t0 = time.perf_counter()
    for s in range(steps):
        _o, _r, term, trunc, _i = env.step(acts[s % 16])
        if term is not env._all_false or trunc is not env._all_false:
            ended += int((term | trunc).sum())
    env.batch.sync()
    dt = max_over_ranks(time.perf_counter() - t0)
    out.update({"value": total * steps / dt, "unit": UNIT, "ms_per_step": 1e3 * dt / steps, "episodes_ended_rank0": ended,
                "api": "BatchedHumanoidEnv.step, auto_reset on (episode pool on the device)"})
    vsteps = max(50, steps // 5)
    barrier()
    t0 = time.perf_counter()
    for s in range(vsteps):
        v.step(acts[s % 16])
    dt = max_over_ranks(time.perf_counter() - t0)
    out["vecenv"] = {"value": total * vsteps / dt, "unit": UNIT, "ms_per_step": 1e3 * dt / vsteps, "steps": vsteps,
                     "api": "B200VecEnv.step (DummyVecEnv contract: copies of obs/reward, dones, infos)"}
    v.close()
    # ---- what an auto-reset costs: ~1 % of the envs end every step vs nobody ends
    res = {}
    for name, limit in (("none_ending", None), ("one_percent_ending_per_step", 0.2)):
        e = hmj.BatchedHumanoidEnv(n, randomization_factor=0.0, max_simulation_time_override=limit, device=local, seed=20 + rank)
        e.auto_reset = True
        e.reset()
        if limit is not None:   # stagger the clocks over one episode length (100 steps): n / 100 envs time out per step
            t = e.batch.get("time")
            t += (np.arange(n) % 100) * 0.002
            e.batch.set("time", t)
        for s in range(120):
            e.step(acts[s % 16])
        barrier()
        k, cnt = 200, 0
        t0 = time.perf_counter()
        for s in range(k):
            _o, _r, term, trunc, _i = e.step(acts[s % 16])
            if term is not e._all_false:
                cnt += int(term.sum())
        e.batch.sync()
        dt = max_over_ranks(time.perf_counter() - t0)
        res[name] = {"ms_per_step": 1e3 * dt / k, "endings_per_step_rank0": cnt / k}
        e.close()
    res["overhead_frac"] = res["one_percent_ending_per_step"]["ms_per_step"] / res["none_ending"]["ms_per_step"] - 1.0
    out["autoreset_cost"] = res
    return out


def c1_bench(local, cpu):
    """simulation/benchmark.py:28-55: ONE env (rf = 1) behind DummyVecEnv, 1000 random-action steps, seconds per step."""
    import humanoid_mujoco_b200 as hmj
    v = hmj.B200VecEnv(1, randomization_factor=1.0, device=local, seed=3)
    v.reset()
    rng = np.random.default_rng(0)
    acts = rng.uniform(-1, 1, (1000, 1, 12))
    for s in range(20):
        v.step(acts[s])
    t0 = time.perf_counter()
    for s in range(1000):
        v.step(acts[s])
    dt = time.perf_counter() - t0
    out = {"steps": 1000, "seconds_per_step": dt / 1000, "value": 1000 / dt, "unit": UNIT,
           "api": "B200VecEnv(1).step — one env is one warp: the GPU path is latency-bound here by construction"}
    if cpu:
        e = v.env
        q, vel, _t, ws = e.batch.get_state()
        value, st, _ = cpu_steps_per_s(e.batch.get_params()[:1], q[:1], vel[:1], ws[:1], acts[:64, :1], 1000, 20, 1)
        out["cpu_baseline"] = {"value": value, "unit": UNIT, "cores": 1, "kind": "port", "seconds_per_step": 1.0 / value,
                               "sample": "the same env (DR params and state after the warm-up), 1000 mj_steps on one thread; the "
                                         "reference adds CPUEnv's Python/JAX per-step work on top of mj_step"}
    v.close()
    return out


# ---------------------------------------------------------------------------------------------- GPU arm
def run_gpu(args):
    import torch
    import torch.distributed as dist

    import humanoid_mujoco_b200 as hmj

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # the contract is ONE JSON line on stdout: libraries that chat on fd 1 (NCCL prints its version there) go to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the engine has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n, K, W = args.envs, args.steps, args.warmup
    # ---- public API object: the CPUEnv contract over n envs, rf = 1
    env = hmj.BatchedHumanoidEnv(n, randomization_factor=1.0, device=local, seed=1 + rank)
    b = env.batch
    zeros18 = np.zeros((n, 18))

    def probe(params, qpos):   # first zero-control step on the engine: who touches anything?
        b.set_params(params)
        b.set_state(qpos, zeros18, np.zeros(n), zeros18)
        b.step(np.zeros((n, 12)), 1)
        return b.get("ncon") > 0

    params, qpos, attempts = make_workload(b.default_params(), b.model["qpos0"][0], n, 1 + rank, probe)
    b.set("status", np.zeros(n, dtype=np.int32))
    env.reset_to(params, qpos)          # delays / constant noise vectors / trackers of the public API
    b.step(np.zeros((n, 12)), SETTLE_STEPS)  # robots fall and lie on the bumpy floor
    q_s, v_s, _t, ws_s = b.get_state()
    settled = (q_s, v_s, np.zeros(n), ws_s)
    stream = torch.cuda.ExternalStream(b.stream, device=dev)

    # ---- device-resident inputs: control ring larger than L2 (R x n x 12 fp64), epilogue buffers
    ring_n = max(8, int(np.ceil(160e6 / (n * 12 * 8))))
    ring_host = control_ring(rank, ring_n, n)
    ring1 = torch.from_numpy(ring_host).to(dev)
    prev = torch.zeros((n, 12), dtype=torch.float64, device=dev)
    obs = torch.zeros((n, 30), dtype=torch.float64, device=dev)
    rew = torch.zeros(n, dtype=torch.float64, device=dev)
    flags = torch.zeros((3, n), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    lib, h = b.lib, b._h
    from humanoid_mujoco_b200._lib import check
    host_actions = hmj.pinned_empty((min(ring_n, 64), n, 12))   # the policy's output buffer: page-locked host memory
    hbm_peak, peak_kind = load_peaks()
    import ctypes as C
    tf = C.c_double(0)
    lib.hmj_fp64_peak(local, C.byref(tf))
    nthreads = os.cpu_count() or 1
    clocks = ClockSampler(range(world) if rank == 0 else [])   # ranks are local GPUs 0..world-1 of one node
    clocks.start()
    clocks.wait_first()

    def measure(name, scale, pre, sample_clocks):
        """One control regime: lead-in to the steady state, W warm-up steps, K timed steps (device resident), K profiled
        steps, then W + K steps end to end through BatchedHumanoidEnv.step."""
        ring = ring1 if scale == 1.0 else ring1 * scale

        def device_step(s):
            cur = ring[s % ring_n]
            b.step_dev(cur.data_ptr(), 1)
            check(lib.hmj_env_obs_reward_dev(h, prev.data_ptr(), cur.data_ptr(), None, 10.0, obs.data_ptr(), rew.data_ptr(),
                                             flags[0].data_ptr(), flags[1].data_ptr(), flags[2].data_ptr()))

        def stats():
            return contact_stats(b.get("ncon"), b.get("nefc"), b.get("niter"), b.get("status"))

        b.set_state(*settled)
        for s in range(pre + W):
            device_step(s)
        b.sync()
        st_start = stats()
        start_state = b.get_state()
        barrier()
        torch.cuda.synchronize()
        if sample_clocks:
            clocks.mark()
        l0 = b.kernel_launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for s in range(K):
            device_step(pre + W + s)
        e1.record(stream)
        b.sync()
        torch.cuda.synchronize()
        barrier()
        ms_total = max_over_ranks(e0.elapsed_time(e1))
        launches = b.kernel_launches - l0
        st_end = stats()
        # per-kernel device time (CUDA events around every launch on the engine stream), the next K steps
        b.profile(True)
        for s in range(K):
            device_step(pre + W + K + s)
        prof = b.profile_read()
        b.profile(False)
        kern = {k: prof[k + "_ms"] / prof["steps"] for k in ("smooth", "collide", "solve")}
        # end to end through the public API: host actions in, obs/reward/flags out, every step.  BatchedHumanoidEnv.step
        # DMAs the actions from the caller's pinned array, runs prologue, physics step and epilogue on the device and
        # reads the results from the pinned IO block (one D2H copy).
        host_actions[...] = np.random.default_rng(5 + rank).uniform(-scale, scale, host_actions.shape)
        for s in range(W):
            env.step(host_actions[s % host_actions.shape[0]])
        # the interpreter's cyclic collector can fire inside a 6 ms timed region and cost a millisecond on this heap
        # (torch + numpy loaded): collect now and park the survivors, as a serving loop would at start-up
        gc.collect()
        gc.freeze()
        barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for s in range(K):
            env.step(host_actions[(W + s) % host_actions.shape[0]])
        b.sync()
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        out = {"value": world * n * K / (ms_total * 1e-3), "unit": UNIT, "ms_per_step": ms_total / K,
               "controls": f"{scale} x U(-1,1)", "lead_in_steps": pre,
               "e2e": {"value": world * n * K / e2e_s, "unit": UNIT, "ms_per_step": 1e3 * e2e_s / K},
               "kernel_ms": kern, "contacts_start": st_start, "contacts_end": st_end}
        out["_start_state"] = start_state
        return out, launches

    results = {}
    launches = 0
    for name, scale, pre in WORKLOADS:
        results[name], ln = measure(name, scale, pre, sample_clocks=(name == "C3"))
        if name == "C3":
            launches = ln
            clk = clocks.stop()
    # ---- CPU baseline legs (rank 0, N = 1 only): the oracle on the first 1024 envs, from the GPU arm's own states
    for name, scale, pre in WORKLOADS:
        out = results[name]
        start_state = out.pop("_start_state")
        if rank == 0 and world == 1 and not args.no_cpu:
            m = min(CPU_SAMPLE_ENVS, n)
            q, v, _tt, ws = start_state
            cpu_value, cst, _ = cpu_steps_per_s(params[:m], q[:m], v[:m], ws[:m], ring_host[:64, :m] * scale, args.cpu_steps, 0,
                                                nthreads)
            out["cpu_baseline"] = {"value": cpu_value, "unit": UNIT, "cores": nthreads, "kind": "port",
                                   "sample": (f"first {m} of the {n} envs from the GPU arm's own state at the start of the "
                                              f"timed region, same DR params and control stream, {args.cpu_steps} steps each"),
                                   **cst}
            fl = cst["flops_per_step"]
            out["fp64"] = {"flops_per_env_step_oracle_counters": fl, "achieved_tflops": out["value"] * fl / 1e12,
                           "peak_tflops": tf.value, "frac": (out["value"] * fl / 1e12) / tf.value if tf.value else None}
    main = results["C3"]
    kern = main["kernel_ms"]
    dom = max(kern, key=kern.get)
    achieved = BYTES_PER_ENV_STEP * n / (kern[dom] * 1e-3) / 1e9
    traffic = None
    try:   # DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture (same 4096-env launch)
        with open(os.path.join(REPO, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get("k_" + dom) if n == ENVS_PER_GPU else None
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "k_" + dom, "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak, "traffic": traffic, "peak_kind": peak_kind + " (MEASURED_PEAKS.json hbm_gbs)",
                "kernel_ms": kern, "algorithmic_bytes_per_env_step": BYTES_PER_ENV_STEP,
                "note": "path is fp64-issue/latency bound, not HBM bound (SURVEY 8d): see roofline_fp64"}
    h2d = n * 12 * 8
    d2h = n * ((30 + 1 + 1 + 4) * 8 + 3 * 4)

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": main["value"], "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": workload_name(n, world), "envs_per_gpu": n, "timestep": 0.002,
                       "reset_attempts": attempts,
                       "l2": (f"per-step controls stream from a {ring_n}x{n}x12 fp64 device ring "
                              f"({ring_n * n * 96 / 1e6:.0f} MB > 126 MB L2); the state is each step's own previous output"),
                       "step": "one mj_step of all envs + obs/reward epilogue"},
            "roofline": roofline,
            "roofline_fp64": {"bound": "fp64", "peak": tf.value, "unit": "TFLOP/s",
                              "peak_kind": "builder-measured hmj_fp64_peak (DFMA issue peak, 8 chains/thread)",
                              **({k: main["fp64"][k] for k in ("achieved_tflops", "frac", "flops_per_env_step_oracle_counters")}
                                 if "fp64" in main else {})},
            "e2e": {"value": main["e2e"]["value"], "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": main["e2e"]["ms_per_step"],
                    "api": "BatchedHumanoidEnv.step (hmj_env_step_fused_from C-ABI: actions DMA'd from the caller's pinned array, results into the pinned IO block, device-resident env bookkeeping)"},
            "gpu_launches": int(launches), "clocks": clk,
            "contacts": {"start": {k: v for k, v in main["contacts_start"].items() if not k.endswith("hist") and "hist" not in k},
                         "end": {k: v for k, v in main["contacts_end"].items() if "hist" not in k}},
            "workloads": results,
        }
        if "cpu_baseline" in main:
            line["cpu_baseline"] = main["cpu_baseline"]
            line["fp64"] = main["fp64"]
    env.close()
    if not args.no_c5:
        c5 = c5_bench(local, rank, world, args.c5_steps, barrier, max_over_ranks)
        if rank == 0:
            line["c5"] = c5
            line["c1"] = c1_bench(local, cpu=(world == 1 and not args.no_cpu))
    if not args.no_plan:
        plan = plan_bench(local, rank, world, args.plan_iters, cpu=(world == 1 and not args.no_cpu))
        if rank == 0:
            line["plan"] = plan
            line["fd"] = fd_bench(local, cpu=(world == 1 and not args.no_cpu))
    if world > 1:
        dist.destroy_process_group()
    sys.stdout.flush()
    if rank == 0:
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    os.close(json_fd)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--envs", type=int, default=ENVS_PER_GPU, help="environments per GPU")
    ap.add_argument("--cpu-steps", type=int, default=300, help="steps per env of the bounded CPU-baseline sample")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-plan", action="store_true", help="skip the planner (M2) measurement")
    ap.add_argument("--no-c5", action="store_true", help="skip the C5 (16384-env rollout collection, strong scaling) and C1 legs")
    ap.add_argument("--c5-steps", type=int, default=1000)
    ap.add_argument("--plan-iters", type=int, default=20, help="timed planner iterations of the C4 slice (C2 runs 5x as many)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
