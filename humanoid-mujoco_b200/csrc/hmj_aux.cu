// hmj_aux.cu — small per-env kernels around the physics step.
//
//  * planner bookkeeping of Trajectory::NoisyRollout (mjpc/trajectory.cc:100-210): broadcast of the
//    shared initial state, optional trajectory recording, total_return = mean of knot costs
//    (UpdateReturn, trajectory.cc:312-326), local arg-min with lowest-index tie-break;
//  * the CPUEnv.step prologue/epilogue (simulation/cpu_env.py:656-674, 465-616) and standupReward
//    (simulation/reward_functions.py:247-374), one thread per environment.
// These are HBM/L2-bound streaming kernels (a few hundred bytes per env), thread-per-env.
#include "hmj_dev.cuh"
#include "hmj_envstep.cuh"

// ------------------------------------------------------------------ planner
// `state` = qpos | qvel | time (HMJ_NSTATE + 1 doubles on the device: no by-value argument changes between planning
// iterations, so the whole rollout replays as one CUDA graph)
__global__ void k_plan_init(hmj_buffers_t B, const double* __restrict__ state) {
  const int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= B.n) return;
  const double time = state[HMJ_NSTATE];
  for (int k = 0; k < HMJ_NQ; k++) B.qpos[(size_t)env * HMJ_NQ + k] = state[k];
  for (int k = 0; k < HMJ_NV; k++) {
    B.qvel[(size_t)env * HMJ_NV + k] = state[HMJ_NQ + k];
    B.warmstart[(size_t)env * HMJ_NV + k] = 0.0;  // deterministic stand-in for per-thread mjData reuse (planner.cc:374)
  }
  for (int k = 0; k < 6; k++) B.xfrc[(size_t)env * 6 + k] = 0.0;
  // fresh mjData: ctrl = 0 (only visible to a horizon-1 rollout, whose single mj_forward runs before any policy call;
  // the reference leaves the worker's stale ctrl there, trajectory.cc:186-198)
  for (int k = 0; k < HMJ_NU; k++) B.ctrl[(size_t)env * HMJ_NU + k] = 0.0;
  B.xfrc_body[env] = 0;
  // deterministic stand-in for the worker's stale xfrc_applied (NoisyRollout never clears it, trajectory.cc:100-155)
  for (int k = 0; k < HMJ_NBODY * 6; k++) B.xfrc_all[(size_t)env * HMJ_NBODY * 6 + k] = 0.0;
  B.time[env] = time;
  B.cost_sum[env] = 0.0;
  B.status[env] = 0;
}

// Ornstein-Uhlenbeck force perturbation of Trajectory::NoisyRollout (mjpc/trajectory.cc:147-155), step t of the rollout:
// xfrc_applied[i] = rate * xfrc_applied[i] + N(0, scale) for all 6 * nbody entries, applied after the policy and before the
// step.  `noise` [N][nsteps][90] holds unit normals injected by the caller (parity tests); NULL draws them from the
// counter-based stream (seed, call counter, global sample, t * 90 + k).  Thread per (sample, entry).
__global__ void k_plan_ou(hmj_buffers_t B, int t, int nsteps, double rate, double scale, const double* __restrict__ noise,
                          const unsigned long long* __restrict__ seed_iter, int first_global) {
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  const int per = HMJ_NBODY * 6;
  if (id >= B.n * per) return;
  const int env = id / per, k = id - env * per;
  double z;
  if (noise) {
    z = noise[((size_t)env * nsteps + t) * per + k];
  } else {
    const uint64_t key = hmj_splitmix64(seed_iter[0] ^ hmj_splitmix64(seed_iter[1] + 0x0B0B0B0B0B0B0B0Bull));
    const uint32_t c = 2u * (uint32_t)(t * per + k);
    const double u1 = 1.0 - hmj_rand_u01(key, (uint32_t)(first_global + env), c);
    const double u2 = hmj_rand_u01(key, (uint32_t)(first_global + env), c + 1u);
    z = sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
  }
  double* x = B.xfrc_all + (size_t)env * per + k;
  *x = rate * (*x) + z * scale;
}

void hmj_launch_plan_ou(hmj_buffers_t buf, int t, int nsteps, double rate, double scale, const double* noise_dev,
                        const unsigned long long* seed_iter_dev, int first_global, cudaStream_t stream) {
  const int tot = buf.n * HMJ_NBODY * 6;
  k_plan_ou<<<(tot + 255) / 256, 256, 0, stream>>>(buf, t, nsteps, rate, scale, noise_dev, seed_iter_dev, first_global);
}

__global__ void k_plan_record(hmj_buffers_t B, hmj_record_t R, int t) {
  const int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= B.n) return;
  const size_t row = (size_t)t * B.n + env;
  for (int k = 0; k < HMJ_NQ; k++) R.states[row * HMJ_NSTATE + k] = B.qpos[(size_t)env * HMJ_NQ + k];
  for (int k = 0; k < HMJ_NV; k++) R.states[row * HMJ_NSTATE + HMJ_NQ + k] = B.qvel[(size_t)env * HMJ_NV + k];
  for (int k = 0; k < HMJ_NU; k++) R.actions[row * HMJ_NU + k] = B.ctrl[(size_t)env * HMJ_NU + k];
  for (int k = 0; k < HMJ_NRESIDUAL; k++) R.residual[row * HMJ_NRESIDUAL + k] = B.residual[(size_t)env * HMJ_NRESIDUAL + k];
  for (int k = 0; k < 3; k++) R.trace[row * 3 + k] = B.trace[(size_t)env * 3 + k];
  R.times[row] = B.time[env];
  R.costs[row] = B.cost_last[env];
}

__global__ void k_plan_finish(hmj_buffers_t B, int horizon, double* __restrict__ returns, int32_t* __restrict__ failure) {
  const int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= B.n) return;
  const int fail = B.status[env] != 0;
  returns[env] = fail ? 1.0e6 : B.cost_sum[env] / (double)(horizon > 1 ? horizon : 1);  // kMaxReturnValue
  if (failure) failure[env] = fail;
}

// single-CTA arg-min; ties -> lowest index (std::partial_sort leaves ties unspecified)
__global__ void k_argmin(const double* __restrict__ returns, int n, double* __restrict__ out) {
  __shared__ double sv[1024];
  __shared__ int si[1024];
  double best = INFINITY;
  int bi = 0x7fffffff;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const double v = returns[i];
    if (v < best || (v == best && i < bi)) { best = v; bi = i; }
  }
  sv[threadIdx.x] = best; si[threadIdx.x] = bi;
  __syncthreads();
  for (int s = blockDim.x / 2; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      const double ov = sv[threadIdx.x + s];
      const int oi = si[threadIdx.x + s];
      if (ov < sv[threadIdx.x] || (ov == sv[threadIdx.x] && oi < si[threadIdx.x])) { sv[threadIdx.x] = ov; si[threadIdx.x] = oi; }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) { out[0] = sv[0]; out[1] = (double)si[0]; }
}

void hmj_launch_plan_init(hmj_buffers_t buf, const double* state_dev, cudaStream_t stream) {
  k_plan_init<<<(buf.n + 127) / 128, 128, 0, stream>>>(buf, state_dev);
}
void hmj_launch_plan_record(hmj_buffers_t buf, hmj_record_t rec, int t, cudaStream_t stream) {
  k_plan_record<<<(buf.n + 127) / 128, 128, 0, stream>>>(buf, rec, t);
}
void hmj_launch_plan_finish(hmj_buffers_t buf, int horizon, double* returns_dev, int32_t* failure_dev, cudaStream_t stream) {
  k_plan_finish<<<(buf.n + 127) / 128, 128, 0, stream>>>(buf, horizon, returns_dev, failure_dev);
}
void hmj_launch_argmin(const double* returns_dev, int n, double* out, cudaStream_t stream) {
  k_argmin<<<1, 1024, 0, stream>>>(returns_dev, n, out);
}

// ------------------------------------------------------------------ one planning iteration on the device
// SamplingPlanner::AddNoiseToPolicy (mjpc/planners/sampling/planner.cc:313-339): candidate s = clamp(nominal + noise_s);
// the sample with GLOBAL index 0 keeps the nominal (planner.cc:361).  `noise` != NULL: caller-provided values (parity
// tests inject them); NULL: drawn here from the counter-based stream (seed, iteration, global sample, element) with
// Box-Muller — the reference draws absl::Gaussian from an entropy-seeded BitGen (unpinned by construction).
// std_s = std1 with probability 0.2 when std1 > 0 (planner.cc:324-329), else std0.  Thread per (sample, node, actuator).
// Cross-entropy flavour (mjpc/planners/cross_entropy/planner.cc:346-378): `std_vec` [nnode * 12] is the ABSOLUTE standard
// deviation of every spline parameter (max(sqrt(variance), std_min), no control-range scaling) and replaces std0 / std1.
__global__ void k_plan_candidates(const hmj_tables_t* __restrict__ T, int n, int nnode, const double* __restrict__ nominal,
                                  const double* __restrict__ noise, double std0, double std1, unsigned long long seed,
                                  unsigned long long iteration, int first_global, double* __restrict__ values,
                                  const double* __restrict__ std_vec) {
  const int per = nnode * HMJ_NU;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)n * per) return;
  const int s = (int)(t / per), e = (int)(t - (long long)s * per), a = e % HMJ_NU;
  const double lo = T->m.act_ctrlrange[a][0], hi = T->m.act_ctrlrange[a][1];
  double z = 0.0;
  if (first_global + s != 0) {
    if (noise) {
      z = noise[t];
    } else {
      const uint64_t key = hmj_splitmix64(seed ^ hmj_splitmix64(iteration + 0x632BE59BD9B4E019ull));
      const uint32_t gs = (uint32_t)(first_global + s);
      double sd = std0;
      if (std1 > 0 && hmj_rand_u01(key, gs, 0xFFFFFFFFu) < 0.2) sd = std1;
      const double u1 = 1.0 - hmj_rand_u01(key, gs, 2u * (uint32_t)e);        // (0, 1]
      const double u2 = hmj_rand_u01(key, gs, 2u * (uint32_t)e + 1u);
      const double g = sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
      z = std_vec ? g * std_vec[e] : g * (0.5 * (hi - lo)) * sd;
    }
  }
  const double v = nominal[e] + z;
  values[t] = fmin(fmax(v, lo), hi);
}

// Local arg-min of the returns (lowest index wins ties) and the winner record this rank contributes to the exchange:
// rec = { cost, GLOBAL index, return of local sample 0 (the nominal on the rank that owns global sample 0), winner spline
// values [nnode][12] }.  Single CTA.
__global__ void k_plan_winner(const double* __restrict__ returns, int n, const double* __restrict__ values, int nnode,
                              int first_global, double* __restrict__ argmin_out, double* __restrict__ rec) {
  __shared__ double sv[1024];
  __shared__ int si[1024];
  double best = INFINITY;
  int bi = 0x7fffffff;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const double v = returns[i];
    if (v < best || (v == best && i < bi)) { best = v; bi = i; }
  }
  sv[threadIdx.x] = best; si[threadIdx.x] = bi;
  __syncthreads();
  for (int s = blockDim.x / 2; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      const double ov = sv[threadIdx.x + s];
      const int oi = si[threadIdx.x + s];
      if (ov < sv[threadIdx.x] || (ov == sv[threadIdx.x] && oi < si[threadIdx.x])) { sv[threadIdx.x] = ov; si[threadIdx.x] = oi; }
    }
    __syncthreads();
  }
  const int w = si[0] == 0x7fffffff ? 0 : si[0];   // every return NaN: sample 0 stands
  if (threadIdx.x == 0) {
    argmin_out[0] = sv[0]; argmin_out[1] = (double)w;
    rec[0] = returns[w]; rec[1] = (double)(first_global + w); rec[2] = returns[0];
  }
  for (int e = threadIdx.x; e < nnode * HMJ_NU; e += blockDim.x) rec[3 + e] = values[(size_t)w * nnode * HMJ_NU + e];
}

// The exchange's reduction: `gathered` holds one winner record per rank (stride doubles each).  The global winner is the
// lexicographic minimum of (cost, global index) — NaN costs never win unless all are NaN — and the nominal return is rank
// 0's (it owns global sample 0).  One warp; equals dist.select_winner on the host (tests/test_dist_gloo.py).
__global__ void k_plan_select(const double* __restrict__ gathered, int nranks, int stride, double* __restrict__ out) {
  const int lane = threadIdx.x;
  double best = INFINITY, bidx = INFINITY;
  int br = 0x7fffffff;
  for (int r = lane; r < nranks; r += 32) {
    double c = gathered[(size_t)r * stride];
    if (!(c == c)) c = INFINITY;
    const double ix = gathered[(size_t)r * stride + 1];
    if (c < best || (c == best && ix < bidx) || br == 0x7fffffff) { best = c; bidx = ix; br = r; }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const double ob = __shfl_xor_sync(0xffffffffu, best, o), oi = __shfl_xor_sync(0xffffffffu, bidx, o);
    const int orr = __shfl_xor_sync(0xffffffffu, br, o);
    if (orr != 0x7fffffff && (br == 0x7fffffff || ob < best || (ob == best && oi < bidx))) { best = ob; bidx = oi; br = orr; }
  }
  const double* src = gathered + (size_t)br * stride;
  for (int e = lane; e < stride; e += 32) out[e] = (e == 2) ? gathered[2] : src[e];
}

void hmj_launch_plan_candidates(const hmj_tables_t* tab, int n, int nnode, const double* nominal_dev, const double* noise_dev,
                                double std0, double std1, unsigned long long seed, unsigned long long iteration, int first_global,
                                double* values_dev, cudaStream_t stream, const double* std_vec_dev) {
  const long long tot = (long long)n * nnode * HMJ_NU;
  k_plan_candidates<<<(int)((tot + 255) / 256), 256, 0, stream>>>(tab, n, nnode, nominal_dev, noise_dev, std0, std1, seed, iteration,
                                                                  first_global, values_dev, std_vec_dev);
}
void hmj_launch_plan_winner(const double* returns_dev, int n, const double* values_dev, int nnode, int first_global,
                            double* argmin_dev, double* rec_dev, cudaStream_t stream) {
  k_plan_winner<<<1, 1024, 0, stream>>>(returns_dev, n, values_dev, nnode, first_global, argmin_dev, rec_dev);
}
void hmj_launch_plan_select(const double* gathered_dev, int nranks, int stride, double* out_dev, cudaStream_t stream) {
  k_plan_select<<<1, 32, 0, stream>>>(gathered_dev, nranks, stride, out_dev);
}

// ------------------------------------------------------------------ CPUEnv epilogue (per-warp bodies: hmj_envstep.cuh)
// One warp per environment of rows [first, last).  `post`: also the delay FIFOs / torso_quat of k_env_post (no mask).
#define OBS_WARPS 4
__global__ void __launch_bounds__(32 * OBS_WARPS)
k_env_obs_reward(const hmj_tables_t* __restrict__ T, hmj_buffers_t B, const double* __restrict__ prev_action,
                 const double* __restrict__ latest_action, const double* __restrict__ noise, double max_sim_time,
                 hmj_results_t R, int first, int last, hmj_envstate_t E, int post) {
  __shared__ double st[OBS_WARPS][5 * HMJ_NU];   // tq, dc, reg, u per actuator; sym terms
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int env = first + blockIdx.x * OBS_WARPS + w;
  if (env >= last) return;
  if (B.active && !B.active[env]) return;
  env_obs_reward_warp(T, B, env, env, lane, st[w], prev_action, latest_action, noise, max_sim_time, R, post, E);
}

void hmj_launch_env_obs_reward(const hmj_tables_t* tab, hmj_buffers_t buf, const double* prev_action, const double* latest_action,
                               const double* noise, double max_sim_time, double* obs, double* reward, int32_t* terminated,
                               int32_t* truncated, int32_t* self_collision, cudaStream_t stream) {
  hmj_results_t R{};
  R.obs = obs; R.reward = reward; R.terminated = terminated; R.truncated = truncated; R.self_collision = self_collision;
  k_env_obs_reward<<<(buf.n + OBS_WARPS - 1) / OBS_WARPS, 32 * OBS_WARPS, 0, stream>>>(
      tab, buf, prev_action, latest_action, noise, max_sim_time, R, 0, buf.n, hmj_envstate_t{}, 0);
}

// Epilogue of the control step for envs [first, first + count) of the batch as a kernel of its own: observation + reward
// + the delay FIFOs and torso_quat of k_env_post (no mask).
void hmj_launch_env_epilogue(const hmj_tables_t* tab, hmj_buffers_t buf, hmj_envstate_t env, double max_sim_time, hmj_results_t R,
                             int first, int count, cudaStream_t stream) {
  k_env_obs_reward<<<(count + OBS_WARPS - 1) / OBS_WARPS, 32 * OBS_WARPS, 0, stream>>>(
      tab, buf, env.prev_action, env.latest_action, env.obs_noise, max_sim_time, R, first, first + count, env, 1);
}

// ------------------------------------------------------------------ CPUEnv prologue / delay FIFOs on the device
// CPUEnv._apply_action (cpu_env.py:656-674) + _apply_external_forces (:618-654).  Warp per environment: lane a < 12 owns
// actuator a (coalesced rows of the FIFO), lane 0 runs the push scheduler.
#define ENV_WARPS 4
__global__ void __launch_bounds__(32 * ENV_WARPS)
k_env_pre(hmj_buffers_t B, hmj_envstate_t E, const double* __restrict__ action, const int32_t* __restrict__ mask,
          double rf, unsigned long long seed, int sid0, int first, int last) {
  const int lane = threadIdx.x & 31;
  const int env = first + blockIdx.x * ENV_WARPS + (threadIdx.x >> 5);
  if (env >= last || (mask && !mask[env])) return;
  // ---- action: + constant noise, delay FIFO, NO clipping (cpu_env.py:670-671); action == NULL re-sends data.ctrl
  const int cnt = E.act_count[env];
  const int d = E.act_delay[env];
  double* hist = E.act_hist + (size_t)env * HMJ_ENV_HIST * HMJ_NU;
  double* ctrl = B.ctrl + (size_t)env * HMJ_NU;
  const int slot = cnt % HMJ_ENV_HIST;
  const int src = ((cnt - d) % HMJ_ENV_HIST + HMJ_ENV_HIST) % HMJ_ENV_HIST;
  if (lane < HMJ_NU) {
    const int a = lane;
    const double v = action ? action[(size_t)env * HMJ_NU + a] + E.act_noise[(size_t)env * HMJ_NU + a] : ctrl[a];
    hist[slot * HMJ_NU + a] = v;
    const double take = (cnt >= d) ? (src == slot ? v : hist[src * HMJ_NU + a]) : 0.0;
    ctrl[a] = take;
    E.prev_action[(size_t)env * HMJ_NU + a] = E.latest_action[(size_t)env * HMJ_NU + a];
    E.latest_action[(size_t)env * HMJ_NU + a] = take;
  }
  __syncwarp();
  if (lane != 0) return;
  E.act_count[env] = cnt + 1;
  // ---- random pushes: every U(1,3) s, for U(.05,.15) s, U(5,15) rf N along a random horizontal direction on a random body
  double* fs = E.force_sched + (size_t)env * 5;
  double* xf = B.xfrc + (size_t)env * 6;
  const double t = B.time[env];
  if (t >= fs[0] + fs[1]) {
    const uint32_t c = (uint32_t)E.force_draws[env] * 8u;
    // contraction-free (__dmul_rn / __dadd_rn): the numpy restatement in env.py must see the same bits
    fs[0] = __dadd_rn(t, __dadd_rn(1.0, __dmul_rn(2.0, hmj_rand_u01(seed, sid0 + env, c + 0))));
    fs[1] = __dadd_rn(0.05, __dmul_rn(0.10, hmj_rand_u01(seed, sid0 + env, c + 1)));
    fs[2] = __dadd_rn(5.0 * rf, __dmul_rn(15.0 * rf - 5.0 * rf, hmj_rand_u01(seed, sid0 + env, c + 2)));
    double dx = __dadd_rn(-1.0, __dmul_rn(2.0, hmj_rand_u01(seed, sid0 + env, c + 3)));
    double dy = __dadd_rn(-1.0, __dmul_rn(2.0, hmj_rand_u01(seed, sid0 + env, c + 4)));
    const double nrm = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    if (nrm == 0) { dx = 1.0; dy = 0.0; } else { dx /= nrm; dy /= nrm; }
    fs[3] = dx; fs[4] = dy;
    xf[0] = 0.0; xf[1] = 0.0;
    int body = 1 + (int)(14.0 * hmj_rand_u01(seed, sid0 + env, c + 5));
    if (body > 14) body = 14;
    E.force_body[env] = body;
    E.force_draws[env] += 1;
  }
  if (t > fs[0] && t < fs[0] + fs[1]) { xf[0] = fs[3] * fs[2]; xf[1] = fs[4] * fs[2]; }
  B.xfrc_body[env] = E.force_body[env];
}

// Observation delay FIFOs of CPUEnv._get_obs (cpu_env.py:528-547): `obs` holds the undelayed observation on entry,
// the delayed one on exit; also CPUEnv.torso_quat (:516).
// Warp per environment, lane k < 30 owns observation element k.
__global__ void __launch_bounds__(32 * ENV_WARPS)
k_env_post(hmj_buffers_t B, hmj_envstate_t E, const int32_t* __restrict__ mask, double* __restrict__ obs,
           double* __restrict__ time_out, double* __restrict__ quat_out) {
  const int lane = threadIdx.x & 31;
  const int env = blockIdx.x * ENV_WARPS + (threadIdx.x >> 5);
  if (env >= B.n) return;
  hmj_results_t R{};
  R.obs = obs; R.time = time_out; R.quat = quat_out;
  env_post_warp(B, E, mask, R, env, env, lane);
}

void hmj_launch_env_pre(hmj_buffers_t buf, hmj_envstate_t env, const double* action, const int32_t* mask, double rf,
                        unsigned long long seed, cudaStream_t stream, int stream_id_offset) {
  k_env_pre<<<(buf.n + ENV_WARPS - 1) / ENV_WARPS, 32 * ENV_WARPS, 0, stream>>>(buf, env, action, mask, rf, seed, stream_id_offset, 0,
                                                                                buf.n);
}
// address of the prologue kernel (hmj_api.cu looks its nodes up in the captured control-step graph)
const void* hmj_env_pre_kernel() { return (const void*)k_env_pre; }
// the same for envs [first, first + count) of the batch (`buf` is the whole batch)
void hmj_launch_env_pre_rows(hmj_buffers_t buf, hmj_envstate_t env, const double* action, double rf, unsigned long long seed,
                             int first, int count, cudaStream_t stream) {
  k_env_pre<<<(count + ENV_WARPS - 1) / ENV_WARPS, 32 * ENV_WARPS, 0, stream>>>(buf, env, action, nullptr, rf, seed, 0, first,
                                                                                first + count);
}
void hmj_launch_env_post(hmj_buffers_t buf, hmj_envstate_t env, const int32_t* mask, double* obs, double* time_out, double* quat_out,
                         cudaStream_t stream) {
  k_env_post<<<(buf.n + ENV_WARPS - 1) / ENV_WARPS, 32 * ENV_WARPS, 0, stream>>>(buf, env, mask, obs, time_out, quat_out);
}

// ------------------------------------------------------------------ FP64 FMA peak (roofline denominator)
// MEASURED_PEAKS.json carries no FP64 figure; this micro-benchmark measures the DFMA issue peak of the device:
// 8 independent FMA chains per thread, 4096 iterations, 148 x 16 CTAs of 256 threads.
__global__ void k_fp64_peak(double* out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; i++) {
    x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
    x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
  }
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

extern "C" int hmj_fp64_peak(int device, double* tflops) {
  if (cudaSetDevice(device) != cudaSuccess) return -1;
  const int grid = 148 * 16, block = 256, iters = 4096;
  double* out = nullptr;
  if (cudaMalloc(&out, sizeof(double) * grid * block) != cudaSuccess) return -1;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  double best = 0.0;
  for (int rep = 0; rep < 6; rep++) {
    cudaEventRecord(e0);
    k_fp64_peak<<<grid, block>>>(out, iters, 0.999999, 1e-9);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(out); return -1; }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fl = 2.0 * 8.0 * iters * (double)grid * block;
    const double tf = fl / (ms * 1e-3) * 1e-12;
    if (rep > 0 && tf > best) best = tf;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(out);
  if (tflops) *tflops = best;
  return 0;
}
