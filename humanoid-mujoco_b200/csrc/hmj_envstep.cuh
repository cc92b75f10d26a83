// hmj_envstep.cuh — per-warp epilogue of CPUEnv.step: _get_obs (cpu_env.py:465-571), _get_reward + standupReward
// (cpu_env.py:586-616, reward_functions.py:247-374) and the observation delay FIFOs (cpu_env.py:528-547).
// The bodies of k_env_obs_reward / k_env_post (hmj_aux.cu), one warp per env.
#ifndef HMJ_ENVSTEP_CUH_
#define HMJ_ENVSTEP_CUH_
#include "hmj_dev.cuh"

__device__ __forceinline__ double scaled_exp(double x) { return exp(-x / 0.5); }

// inverseRotateVectors(q, (0,0,-1)) with scipy/JAX Rotation semantics (normalises q): -(third row of R(q))
__device__ __forceinline__ void local_gravity(const double* quat_in, double* g) {
  double q[4] = {quat_in[0], quat_in[1], quat_in[2], quat_in[3]};
  const double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  const double s = 1.0 / n;
  q[0] *= s; q[1] *= s; q[2] *= s; q[3] *= s;
  double R[9];
  d_quat2mat(R, q);
  g[0] = -R[6]; g[1] = -R[7]; g[2] = -R[8];
}

// Delay FIFOs + torso_quat of one env.  `env` indexes B (possibly a view of a pipe), `genv` the env-state and result arrays.
// `R.obs` holds the undelayed observation on entry and the delayed one on exit.  Lane k < 30 owns observation element k.
__device__ __forceinline__ void env_post_warp(const hmj_buffers_t& B, const hmj_envstate_t& E, const int32_t* __restrict__ mask,
                                              const hmj_results_t& R, int env, int genv, int lane) {
  if (lane == 0 && R.time) R.time[genv] = B.time[env];
  if (mask && !mask[genv]) {
    if (R.quat && lane < 4) R.quat[(size_t)genv * 4 + lane] = E.torso_quat[(size_t)genv * 4 + lane];
    return;
  }
  if (lane < 4) {
    const double v = B.sens[(size_t)env * SENS_NUM + SENS_ROOT_XQUAT + lane] + E.obs_noise[(size_t)genv * 31 + 27 + lane];
    E.torso_quat[(size_t)genv * 4 + lane] = v;
    if (R.quat) R.quat[(size_t)genv * 4 + lane] = v;
  }
  const int cnt = E.obs_count[genv];
  if (lane < 30) {
    const int k = lane;
    double* hist = E.obs_hist + (size_t)genv * HMJ_ENV_HIST * 30;
    double* o = R.obs + (size_t)genv * 30;
    const int slot = cnt % HMJ_ENV_HIST;
    const double v = o[k];
    hist[slot * 30 + k] = v;
    const int g = k < 24 ? 0 : (k < 27 ? 1 : 2);   // joint pos+vel / angular velocity / gravity groups
    const int d = E.obs_delay[(size_t)genv * 3 + g];
    const int src = ((cnt - d) % HMJ_ENV_HIST + HMJ_ENV_HIST) % HMJ_ENV_HIST;
    o[k] = (cnt >= d) ? (src == slot ? v : hist[src * 30 + k]) : ((k == 29) ? -1.0 : 0.0);
  }
  __syncwarp();
  if (lane == 0) E.obs_count[genv] = cnt + 1;
}

// Observation + reward + flags of one env by one warp: lanes 0..11 own one actuator each (its two observation entries and
// its three exp() reward terms), lanes 0..5 then the six symmetry terms, lane 15 the noisy-quaternion gravity of the
// observation, lane 0 the remaining scalar terms and the final sum.  Every sum is formed by lane 0 in the reference's
// order (a = 0..11), so the result is the same bits as the one-thread-per-env form; the 45 exp() calls are what made that
// form ~20 us of latency.  `st`: 60 doubles of shared memory owned by the warp.  `post`: run env_post_warp on the fresh
// observation (the reward terms read the state, never `obs`).  All 32 lanes must call.
__device__ __forceinline__ void env_obs_reward_warp(const hmj_tables_t* __restrict__ T, const hmj_buffers_t& B, int env, int genv,
                                                    int lane, double* st, const double* prev_action, const double* latest_action,
                                                    const double* noise, double max_sim_time, const hmj_results_t& R,
                                                    int post, const hmj_envstate_t& E) {
  const hmj_model_t& M = T->m;
  const double* qpos = B.qpos + (size_t)env * HMJ_NQ;
  const double* qvel = B.qvel + (size_t)env * HMJ_NV;
  const double* sens = B.sens + (size_t)env * SENS_NUM;
  const double* nz = noise ? noise + (size_t)genv * 31 : nullptr;
  const double* tq = sens + SENS_ROOT_XQUAT;  // data.xquat[torso_idx]: one step stale
  double* o = R.obs + (size_t)genv * 30;
  // ---- observation (cpu_env.py:465-571), undelayed; JOINT_NAMES follow the actuator order
  if (lane < HMJ_NU) {
    const int a = lane;
    o[a] = qpos[M.act_qposadr[a]] + (nz ? nz[a] : 0.0);
    o[12 + a] = qvel[M.act_dofadr[a]] + (nz ? nz[12 + a] : 0.0);
    // ---- per-actuator terms of standupReward (reward_functions.py:247-374)
    const int d = M.act_dofadr[a];
    const double tau = B.qfrc_constraint[(size_t)env * HMJ_NV + d] + B.qfrc_smooth[(size_t)env * HMJ_NV + d];
    const double ex = fmax(fabs(tau) - 1.0, 0.0);
    const double lat = latest_action[(size_t)genv * HMJ_NU + a] / (HMJ_PI / 2);
    const double prv = prev_action[(size_t)genv * HMJ_NU + a] / (HMJ_PI / 2);
    const double dc = (lat - prv) * 500.0;
    st[0 * HMJ_NU + a] = scaled_exp(ex * ex);
    st[1 * HMJ_NU + a] = scaled_exp(dc * dc);
    st[2 * HMJ_NU + a] = scaled_exp(lat * lat);
    st[3 * HMJ_NU + a] = lat;
  } else if (lane < HMJ_NU + 3) {
    const int k = lane - HMJ_NU;
    o[24 + k] = qvel[3 + k] + (nz ? nz[24 + k] : 0.0);
  } else if (lane == HMJ_NU + 3) {
    double nq[4] = {tq[0] + (nz ? nz[27] : 0.0), tq[1] + (nz ? nz[28] : 0.0), tq[2] + (nz ? nz[29] : 0.0), tq[3] + (nz ? nz[30] : 0.0)};
    double g[3];
    local_gravity(nq, g);
    o[27] = g[0]; o[28] = g[1]; o[29] = g[2];
  }
  __syncwarp();
  if (post) env_post_warp(B, E, nullptr, R, env, genv, lane);
  if (lane < 6) {
    const double* u = st + 3 * HMJ_NU;
    // elbows equal; hip roll / hip pitch / knee / shoulder pitch / shoulder roll opposite
    const int i1 = (0x438765 >> (4 * lane)) & 15;   // {5, 6, 7, 8, 3, 4}  (nibble tables: no local-memory array)
    const int i2 = (0x10ba92 >> (4 * lane)) & 15;   // {2, 9, 10, 11, 0, 1}
    const double v = (lane == 0) ? (u[i1] - u[i2]) : (u[i1] + u[i2]);
    st[4 * HMJ_NU + lane] = scaled_exp(v * v);
  }
  __syncwarp();
  if (lane != 0) return;
  double gl[3];
  local_gravity(tq, gl);
  const double z = sens[SENS_ROOT_Z];
  double r = 0.0;
  r += 5.0 * scaled_exp(qvel[0] * qvel[0] + qvel[1] * qvel[1]);                 // horizontal velocity (target 0)
  {
    const double dx = gl[0], dy = gl[1], dz = gl[2] + 1.0;
    r += 10.0 * scaled_exp(dx * dx + dy * dy + dz * dz);                        // upright (yaw weight 0)
  }
  {
    const double z0 = -0.6, z1 = -0.375;                                         // np.interp(z, [MIN_Z, TARGET_Z], [0, 15])
    r += (z <= z0) ? 0.0 : (z >= z1 ? 15.0 : (z - z0) / (z1 - z0) * 15.0);
  }
  double tq_sum = 0.0, dc_sum = 0.0, reg_sum = 0.0;
  for (int a = 0; a < HMJ_NU; a++) { tq_sum += st[a]; dc_sum += st[HMJ_NU + a]; reg_sum += st[2 * HMJ_NU + a]; }
  r += 2.5 * tq_sum / 12.0;
  r += 2.0 * dc_sum / 12.0;
  r += 0.5 * reg_sum / 12.0;
  {
    double sym = 0.0;
    for (int k = 0; k < 6; k++) sym += st[4 * HMJ_NU + k];
    r += 1.0 * sym / 6.0;
  }
  int selfc = 0;
  const int nc = B.ncon[env];
  for (int c = 0; c < nc; c++)
    if (B.contact[(size_t)env * HMJ_MAXCON + c].g1 != 0) selfc = 1;
  if (selfc) r += -20.0;
  const int term = B.time[env] >= max_sim_time;
  if (term) r = -100.0;
  const int upright = fmax(fabs(gl[0]), fabs(gl[1])) < 0.7;
  const int trunc = (z >= -0.375) && upright;
  R.reward[genv] = r;
  R.terminated[genv] = term;
  R.truncated[genv] = trunc;
  if (R.self_collision) R.self_collision[genv] = selfc;
}

#endif  // HMJ_ENVSTEP_CUH_
