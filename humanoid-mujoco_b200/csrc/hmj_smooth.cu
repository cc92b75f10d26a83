// hmj_smooth.cu — everything in mj_forward that does not depend on contacts, as TWO kernels:
//
//   k_kin  (critical path, feeds the collision stage)
//          mj_checkPos/Vel -> [planner: spline policy -> ctrl] -> mj_kinematics -> inertial / geom / site frames
//          -> mj_comPos -> cdof
//   k_dyn  (side stream, beside the collision kernels; feeds k_factor and k_solve)
//          mj_crb -> qM, mj_comVel, mj_passive, mj_rne, mj_fwdActuation -> qfrc_smooth (+ xfrc_applied)
//          -> [planner: sensors -> humanoid_cap Stand residual -> norms -> running cost]
//
// Restates (B200-first) the stages declared at simulation/mujoco/include/mujoco/mujoco.h:301-304,
// 310-349,253 and mjpc/planners/sampling/policy.cc:52-59, mjpc/spline/spline.cc:103-156,
// mjpc/tasks/humanoid_cap/stand/stand.cc:40-99, mjpc/task.cc:71-110, mjpc/norm.cc:50-210.
//
// Mapping: one warp per environment; lanes own bodies (15), dofs (18), geoms (10) or sparse-M
// entries (117); the kinematic tree is walked level by level (bodies of equal depth in
// parallel); subtree reductions use warp shuffles.  The split exists because only the geom frames gate the
// collision stage: k_kin keeps 4 KB of shared memory per env (every env of a 4096-env batch is resident at once) and is
// ~1/3 of the old single kernel's latency; the inertia / bias half runs concurrently with broad- and narrowphase.  The
// two halves exchange xipos / ximat (180 doubles per env, L2) and the site positions (in the `sens` row).
// Roofline: FP64 issue-latency bound; HBM traffic is the state row (qpos,qvel,ctrl,params: ~1.7 KB in) and the scratch
// rows (geom frames, cdof, qM, xipos/ximat: ~4.5 KB out, L2 resident).
#include "hmj_dev.cuh"

struct KinSmem {
  double qpos[HMJ_NQ + 1];
  double xpos[HMJ_NBODY][3];
  double xquat[HMJ_NBODY][4];
  double xmat[HMJ_NBODY][9];
  double xipos[HMJ_NBODY][3];
  double xanchor[HMJ_NJNT][3];
  double xaxis[HMJ_NJNT][3];
  double com[4];
};

struct DynSmem {
  double qpos[HMJ_NQ + 1];
  double qvel[HMJ_NV];
  double ctrl[HMJ_NU];
  double xipos[HMJ_NBODY][3];
  double cinert[HMJ_NBODY][10];
  double crb[HMJ_NBODY][10];   // later reused for accumulated cfrc
  double cdof[HMJ_NV][6];
  double cdof_dot[HMJ_NV][6];
  double buf[HMJ_NV][6];
  double cvel[HMJ_NBODY][6];
  double cacc[HMJ_NBODY][6];
  double cfrc[HMJ_NBODY][6];
  double qfrc_passive[HMJ_NV];
  double qfrc_bias[HMJ_NV];
  double qfrc_actuator[HMJ_NV];
  double com[4];
  double residual[HMJ_NRESIDUAL];
};

__device__ __forceinline__ void inert_com(double* res, const double* inert, const double* mat, const double* dif, double mass) {
  double t0 = mat[0] * inert[0], t1 = mat[3] * inert[0], t2 = mat[6] * inert[0];
  double t3 = mat[1] * inert[1], t4 = mat[4] * inert[1], t5 = mat[7] * inert[1];
  double t6 = mat[2] * inert[2], t7 = mat[5] * inert[2], t8 = mat[8] * inert[2];
  res[0] = mat[0] * t0 + mat[1] * t3 + mat[2] * t6 + mass * (dif[1] * dif[1] + dif[2] * dif[2]);
  res[1] = mat[3] * t1 + mat[4] * t4 + mat[5] * t7 + mass * (dif[0] * dif[0] + dif[2] * dif[2]);
  res[2] = mat[6] * t2 + mat[7] * t5 + mat[8] * t8 + mass * (dif[0] * dif[0] + dif[1] * dif[1]);
  res[3] = mat[0] * t1 + mat[1] * t4 + mat[2] * t7 - mass * dif[0] * dif[1];
  res[4] = mat[0] * t2 + mat[1] * t5 + mat[2] * t8 - mass * dif[0] * dif[2];
  res[5] = mat[3] * t2 + mat[4] * t5 + mat[5] * t8 - mass * dif[1] * dif[2];
  res[6] = mass * dif[0]; res[7] = mass * dif[1]; res[8] = mass * dif[2];
  res[9] = mass;
}

// TimeSpline::Sample for one value dimension (spline.cc:103-156, Slope :259-277, CubicCoefficients :240-257)
__device__ double spline_slope(const double* times, const double* vals, int stride, int n, int node) {
  if (node == 0) return (vals[stride] - vals[0]) / (times[1] - times[0]);
  const double* v = vals + (size_t)node * stride;
  if (node == n - 1) return (v[0] - v[-stride]) / (times[node] - times[node - 1]);
  return 0.5 * (v[stride] - v[0]) / (times[node + 1] - times[node]) + 0.5 * (v[0] - v[-stride]) / (times[node] - times[node - 1]);
}

__device__ double spline_sample1(const double* times, const double* vals, int stride, int n, int interpolation, double time) {
  if (n == 0) return 0.0;
  int upper = 0;
  while (upper < n && !(time < times[upper])) upper++;
  if (upper == n) return vals[(size_t)(n - 1) * stride];
  if (upper == 0) return vals[0];
  int lower = upper - 1;
  double dt = times[upper] - times[lower];
  double t = (time - times[lower]) / dt;
  double lo = vals[(size_t)lower * stride], up = vals[(size_t)upper * stride];
  if (interpolation == 0) return lo;
  if (interpolation == 1) return lo * (1 - t) + up * t;
  double c0 = 2.0 * t * t * t - 3.0 * t * t + 1.0;
  double c1 = (t * t * t - 2.0 * t * t + t) * dt;
  double c2 = -2.0 * t * t * t + 3 * t * t;
  double c3 = (t * t * t - t * t) * dt;
  double m0 = spline_slope(times, vals, stride, n, lower);
  double m1 = spline_slope(times, vals, stride, n, upper);
  return c0 * lo + c1 * m0 + c2 * up + c3 * m1;
}

// mjpc Norm, value only (norm.cc:50-210)
__device__ double norm_value(const double* x, const double* prm, int n, int type) {
  double y = 0, p = prm[0], q = prm[1];
  switch (type) {
    case -1: y = x[0]; break;
    case 0:
      for (int i = 0; i < n; i++) y += x[i] * x[i];
      y *= 0.5;
      break;
    case 1: {
      double c = 0;
      for (int i = 0; i < n; i++) c += x[i] * x[i];
      y = pow(pow(c, q / 2) + pow(p, q), 1 / q) - p;
      break;
    }
    case 2: {
      double c = 0;
      for (int i = 0; i < n; i++) c += x[i] * x[i];
      y = sqrt(c + p * p) - p;
      break;
    }
    case 3:
      for (int i = 0; i < n; i++) y += p * p * (cosh(x[i] / p) - 1.0);
      break;
    case 5:
      for (int i = 0; i < n; i++) y += pow(fabs(x[i]), p);
      break;
    case 6:
      for (int i = 0; i < n; i++) y += sqrt(x[i] * x[i] + p * p) - p;
      break;
    case 7:
      for (int i = 0; i < n; i++) y += pow(pow(fabs(x[i]), q) + pow(p, q), 1 / q) - p;
      break;
    case 8:
      for (int i = 0; i < n; i++) y += (p > 0) ? p * log(1 + exp(x[i] / p)) : (x[i] > 0 ? x[i] : 0);
      break;
    default: y = nan("");
  }
  return y;
}

// ------------------------------------------------------------------ k_kin
__global__ void __launch_bounds__(32 * HMJ_WARPS_PER_CTA, 7)   // 7 CTAs x 4 warps x 148 SMs >= 4096 envs: one wave
k_kin(const hmj_tables_t* __restrict__ T, hmj_buffers_t B, int plan_mode, const double* __restrict__ spline_times,
      const double* __restrict__ spline_values, int nnode, int interpolation, int values_per_env) {
  __shared__ KinSmem smem[HMJ_WARPS_PER_CTA];
  const int lane = threadIdx.x & 31;
  const int wib = threadIdx.x >> 5;
  const int env = blockIdx.x * HMJ_WARPS_PER_CTA + wib;
  // Zero the per-step work counters of the stages that follow on this stream (narrowphase task counts and item counter,
  // contact-slot bitmask, solver class counts): three memset nodes less in every step's dependency chain.  The previous
  // step's consumers of these arrays finished before this kernel started (stream order).
  if (blockIdx.x == 0 && threadIdx.x < 64) B.wq_count[threadIdx.x] = 0;
  if (blockIdx.x == 0 && threadIdx.x < 4) B.solve_count[threadIdx.x] = 0;
  if (env >= B.n) return;  // whole warp exits together
  if (B.active && !B.active[env]) return;
  if (lane < 4) B.slot_mask[(size_t)env * 4 + lane] = 0u;
  KinSmem& S = smem[wib];
  const hmj_model_t& M = T->m;
  const hmj_params_t* P = B.params + env;
  const unsigned FULL = 0xffffffffu;

  // ---------------- load state, mj_checkPos / mj_checkVel
  double* g_qpos = B.qpos + (size_t)env * HMJ_NQ;
  double* g_qvel = B.qvel + (size_t)env * HMJ_NV;
  double* g_ctrl = B.ctrl + (size_t)env * HMJ_NU;
  double qp = lane < HMJ_NQ ? g_qpos[lane] : 0.0;
  double qv = lane < HMJ_NV ? g_qvel[lane] : 0.0;
  const double pf_mass = (lane >= 1 && lane < HMJ_NBODY) ? P->body_mass[lane] : 0.0;   // requested now: an L2 round trip
  // model constants of the body this lane owns in the tree walk (independent loads, in flight together with the state)
  const int kc_nlevel = T->nlevel;
  int kc_depth = -1, kc_pid = 0, kc_jadr = 0, kc_jnum = 0, kc_jtype = 0, kc_qadr = 0;
  double kc_bpos[3] = {0, 0, 0}, kc_bquat[4] = {1, 0, 0, 0}, kc_jaxis[3] = {0, 0, 0}, kc_jpos[3] = {0, 0, 0}, kc_q0 = 0.0;
  if (lane >= 1 && lane < HMJ_NBODY) {
    const int b = lane;
    kc_depth = M.body_depth[b];
    kc_pid = M.body_parentid[b];
    kc_jadr = M.body_jntadr[b];
    kc_jnum = M.body_jntnum[b];
    for (int k = 0; k < 3; k++) kc_bpos[k] = M.body_pos[b][k];
    for (int k = 0; k < 4; k++) kc_bquat[k] = M.body_quat[b][k];
    if (kc_jnum == 1) {
      kc_jtype = M.jnt_type[kc_jadr];
      kc_qadr = M.jnt_qposadr[kc_jadr];
      for (int k = 0; k < 3; k++) { kc_jaxis[k] = M.jnt_axis[kc_jadr][k]; kc_jpos[k] = M.jnt_pos[kc_jadr][k]; }
      kc_q0 = M.qpos0[kc_qadr];
    }
  }
  int status = 0;
  bool badq = __any_sync(FULL, d_is_bad(qp));
  bool badv = __any_sync(FULL, d_is_bad(qv));
  if (badq || badv) {  // mj_resetData semantics for this path (k_dyn starts after this kernel and sees the reset state)
    status |= (badq ? HMJ_STATUS_BADQPOS : HMJ_STATUS_BADQVEL) | HMJ_STATUS_RESET_PENDING;   // ctrl and xfrc: zeroed by k_dyn
    qp = lane < HMJ_NQ ? M.qpos0[lane] : 0.0;
    qv = 0.0;
    if (lane < HMJ_NQ) g_qpos[lane] = qp;
    if (lane < HMJ_NV) { g_qvel[lane] = 0.0; B.warmstart[(size_t)env * HMJ_NV + lane] = 0.0; }
    if (lane == 0) B.time[env] = 0.0;
  }
  if (lane < HMJ_NQ) S.qpos[lane] = qp;

  // ---------------- control: planner spline policy (SamplingPolicy::Action) or caller-provided ctrl
  double u = 0.0;
  if (lane < HMJ_NU) {
    if ((plan_mode & 2) && !(badq || badv)) {
      const double* vals = spline_values + (values_per_env ? (size_t)env * nnode * HMJ_NU : 0) + lane;
      u = spline_sample1(spline_times, vals, HMJ_NU, nnode, interpolation, B.time[env]);
      u = fmin(fmax(u, M.act_ctrlrange[lane][0]), M.act_ctrlrange[lane][1]);  // Clamp (utilities.cc:94-98)
      g_ctrl[lane] = u;
    }
  }
  // mjWARN_BADCTRL (controls zeroed): checked here for the controls this kernel makes itself (the spline policy);
  // caller-provided controls are checked by their consumer, k_dyn — this kernel does not read them, so whatever writes
  // them (k_env_pre behind the action upload) can run beside it
  if ((plan_mode & 2) && __any_sync(FULL, lane < HMJ_NU && d_is_bad(u))) {
    status |= HMJ_STATUS_BADCTRL;
    if (lane < HMJ_NU) g_ctrl[lane] = 0.0;
  }
  if (lane == 0) {
    S.xpos[0][0] = S.xpos[0][1] = S.xpos[0][2] = 0;
    S.xquat[0][0] = 1; S.xquat[0][1] = S.xquat[0][2] = S.xquat[0][3] = 0;
    for (int k = 0; k < 9; k++) S.xmat[0][k] = (k % 4 == 0) ? 1.0 : 0.0;
    S.xipos[0][0] = S.xipos[0][1] = S.xipos[0][2] = 0;
  }
  __syncwarp();

  // ---------------- mj_kinematics, level by level.  Lane b owns body b and holds that body's model constants in
  // registers (requested with the state at kernel entry, see kc_*): the loop itself touches no table, where it used to
  // pay three dependent L1-cold lookups per level (depth -> body -> joint -> axis / offsets), 7 levels deep.
  for (int L = 1; L < kc_nlevel; L++) {
    if (kc_depth == L) {
      const int b = lane;
      const int pid = kc_pid;
      const int jadr = kc_jadr;
      double xpos[3], xquat[4];
      if (kc_jnum == 1 && kc_jtype == HMJ_JNT_FREE) {
        const int qadr = kc_qadr;
        xpos[0] = S.qpos[qadr]; xpos[1] = S.qpos[qadr + 1]; xpos[2] = S.qpos[qadr + 2];
        xquat[0] = S.qpos[qadr + 3]; xquat[1] = S.qpos[qadr + 4]; xquat[2] = S.qpos[qadr + 5]; xquat[3] = S.qpos[qadr + 6];
        d_normalize4(xquat);
        for (int k = 0; k < 3; k++) { S.xanchor[jadr][k] = xpos[k]; S.xaxis[jadr][k] = kc_jaxis[k]; }
      } else {
        if (pid) {
          d_rot_vec_mat(xpos, kc_bpos, S.xmat[pid]);
          xpos[0] += S.xpos[pid][0]; xpos[1] += S.xpos[pid][1]; xpos[2] += S.xpos[pid][2];
          d_mul_quat(xquat, S.xquat[pid], kc_bquat);
        } else {
          for (int k = 0; k < 3; k++) xpos[k] = kc_bpos[k];
          for (int k = 0; k < 4; k++) xquat[k] = kc_bquat[k];
        }
        if (kc_jnum == 1) {  // one hinge
          const int qadr = kc_qadr;
          double axis[3], anchor[3], qloc[4], vec[3];
          d_rot_vec_quat(axis, kc_jaxis, xquat);
          d_rot_vec_quat(anchor, kc_jpos, xquat);
          anchor[0] += xpos[0]; anchor[1] += xpos[1]; anchor[2] += xpos[2];
          d_axis_angle2quat(qloc, kc_jaxis, S.qpos[qadr] - kc_q0);
          d_mul_quat(xquat, xquat, qloc);
          d_rot_vec_quat(vec, kc_jpos, xquat);
          for (int k = 0; k < 3; k++) {
            xpos[k] = anchor[k] - vec[k];
            S.xanchor[jadr][k] = anchor[k];
            S.xaxis[jadr][k] = axis[k];
          }
        }
      }
      d_normalize4(xquat);
      for (int k = 0; k < 3; k++) S.xpos[b][k] = xpos[k];
      for (int k = 0; k < 4; k++) S.xquat[b][k] = xquat[k];
      d_quat2mat(S.xmat[b], xquat);
    }
    __syncwarp();
  }

  // ---------------- inertial frames (lanes 1..14) -> global for k_dyn, geom frames (lanes 16..25), sites (lanes 26,27)
  if (lane >= 1 && lane < HMJ_NBODY) {
    const int b = lane;
    double* gi = B.xipos + ((size_t)env * HMJ_NBODY + b) * 3;
    double* gm = B.ximat + ((size_t)env * HMJ_NBODY + b) * 9;
    if (M.body_sameframe[b]) {
      for (int k = 0; k < 3; k++) { S.xipos[b][k] = S.xpos[b][k]; gi[k] = S.xpos[b][k]; }
      for (int k = 0; k < 9; k++) gm[k] = S.xmat[b][k];
    } else {
      double t[4], p[3], mat[9];
      d_rot_vec_mat(p, M.body_ipos[b], S.xmat[b]);
      for (int k = 0; k < 3; k++) { S.xipos[b][k] = p[k] + S.xpos[b][k]; gi[k] = S.xipos[b][k]; }
      d_mul_quat(t, S.xquat[b], M.body_iquat[b]);
      d_quat2mat(mat, t);
      for (int k = 0; k < 9; k++) gm[k] = mat[k];
    }
  } else if (lane >= 16 && lane < 16 + HMJ_NGEOM) {
    const int g = lane - 16, b = M.geom_bodyid[g];
    double t[4], p[3], mat[9];
    d_rot_vec_mat(p, M.geom_pos[g], S.xmat[b]);
    d_mul_quat(t, S.xquat[b], M.geom_quat[g]);
    d_quat2mat(mat, t);
    double* gp = B.geom_xpos + ((size_t)env * HMJ_NGEOM + g) * 3;
    double* gm = B.geom_xmat + ((size_t)env * HMJ_NGEOM + g) * 9;
    for (int k = 0; k < 3; k++) gp[k] = p[k] + S.xpos[b][k];
    for (int k = 0; k < 9; k++) gm[k] = mat[k];
  } else if (lane >= 26 && lane < 26 + HMJ_NSITE) {
    const int s = lane - 26, b = M.site_bodyid[s];
    double p[3];
    d_rot_vec_mat(p, M.site_pos[s], S.xmat[b]);
    for (int k = 0; k < 3; k++) B.sens[(size_t)env * SENS_NUM + SENS_SITE_XPOS + 3 * s + k] = p[k] + S.xpos[b][k];
  }
  __syncwarp();

  // ---------------- mj_comPos: subtree CoM of the root (all moving bodies share root body 1)
  const double mass = pf_mass;
  {
    double cx = 0, cy = 0, cz = 0;
    if (lane >= 1 && lane < HMJ_NBODY) { cx = mass * S.xipos[lane][0]; cy = mass * S.xipos[lane][1]; cz = mass * S.xipos[lane][2]; }
    cx = warp_sum(cx); cy = warp_sum(cy); cz = warp_sum(cz);
    const double stm = M.body_subtreemass[1];  // compile-time constant (no mj_setConst after DR)
    if (stm < HMJ_MINVAL) { cx = S.xipos[1][0]; cy = S.xipos[1][1]; cz = S.xipos[1][2]; }
    else { double s = 1.0 / fmax(HMJ_MINVAL, stm); cx *= s; cy *= s; cz *= s; }
    if (lane == 0) {
      S.com[0] = cx; S.com[1] = cy; S.com[2] = cz;
      double* gc = B.com + (size_t)env * 3;
      gc[0] = cx; gc[1] = cy; gc[2] = cz;
    }
  }
  __syncwarp();
  if (lane < HMJ_NV) {  // cdof, one dof per lane
    const int d = lane, j = M.dof_jntid[d], bi = M.jnt_bodyid[j];
    double off[3] = {S.com[0] - S.xanchor[j][0], S.com[1] - S.xanchor[j][1], S.com[2] - S.xanchor[j][2]};
    double c[6];
    if (M.jnt_type[j] == HMJ_JNT_FREE) {
      const int k = d - M.jnt_dofadr[j];
      if (k < 3) {
        for (int q = 0; q < 6; q++) c[q] = 0;
        c[3 + k] = 1;
      } else {
        double ax[3] = {S.xmat[bi][k - 3], S.xmat[bi][k - 3 + 3], S.xmat[bi][k - 3 + 6]};
        c[0] = ax[0]; c[1] = ax[1]; c[2] = ax[2];
        d_cross(c + 3, ax, off);
      }
    } else {
      c[0] = S.xaxis[j][0]; c[1] = S.xaxis[j][1]; c[2] = S.xaxis[j][2];
      d_cross(c + 3, S.xaxis[j], off);
    }
    double* gc = B.cdof + ((size_t)env * HMJ_NV + d) * 6;
    for (int q = 0; q < 6; q++) gc[q] = c[q];
  }
  // ---------------- values the env epilogue reads after the step (one-step-stale by contract)
  if (lane < 3) B.sens[(size_t)env * SENS_NUM + SENS_TORSO_XPOS + lane] = S.xpos[2][lane];
  if (lane < 4) B.sens[(size_t)env * SENS_NUM + SENS_ROOT_XQUAT + lane] = S.xquat[1][lane];
  if (lane == 0) B.sens[(size_t)env * SENS_NUM + SENS_ROOT_Z] = S.xpos[1][2];
  if (status && lane == 0) atomicOr(&B.status[env], status);
}

// ------------------------------------------------------------------ k_dyn
__global__ void __launch_bounds__(32 * HMJ_WARPS_PER_CTA)
k_dyn(const hmj_tables_t* __restrict__ T, hmj_buffers_t B, const hmj_task_t* __restrict__ task, int plan_mode) {
  extern __shared__ double smem_raw[];
  const int lane = threadIdx.x & 31;
  const int wib = threadIdx.x >> 5;
  const int env = blockIdx.x * HMJ_WARPS_PER_CTA + wib;
  if (env >= B.n) return;  // whole warp exits together
  if (B.active && !B.active[env]) return;
  DynSmem& S = reinterpret_cast<DynSmem*>(smem_raw)[wib];
  const hmj_model_t& M = T->m;
  const hmj_params_t* P = B.params + env;
  const unsigned FULL = 0xffffffffu;

  // ---------------- state and the kinematic quantities k_kin left in L2; per-env parameters requested up front
  if (lane < HMJ_NQ) S.qpos[lane] = B.qpos[(size_t)env * HMJ_NQ + lane];
  if (lane < HMJ_NV) S.qvel[lane] = B.qvel[(size_t)env * HMJ_NV + lane];
  {
    // the tail of k_kin's mj_resetData: a bad state this step zeroes the controls and the applied wrench
    const bool reset = (B.status[env] & HMJ_STATUS_RESET_PENDING) != 0;
    if (reset) {
      if (lane < HMJ_NU) B.ctrl[(size_t)env * HMJ_NU + lane] = 0.0;
      if (lane < 6) B.xfrc[(size_t)env * 6 + lane] = 0.0;
      __syncwarp();
      if (lane == 0) atomicAnd(&B.status[env], ~HMJ_STATUS_RESET_PENDING);
    }
    const double u = (lane < HMJ_NU && !reset) ? B.ctrl[(size_t)env * HMJ_NU + lane] : 0.0;
    const bool badu = !(plan_mode & 2) && __any_sync(FULL, lane < HMJ_NU && d_is_bad(u));   // mjWARN_BADCTRL: controls zeroed
    if (lane < HMJ_NU) {
      S.ctrl[lane] = badu ? 0.0 : u;
      if (badu) B.ctrl[(size_t)env * HMJ_NU + lane] = 0.0;
    }
    if (badu && lane == 0) atomicOr(&B.status[env], HMJ_STATUS_BADCTRL);
  }
  const double pf_mass = (lane >= 1 && lane < HMJ_NBODY) ? P->body_mass[lane] : 0.0;
  const double pf_arm = lane < HMJ_NV ? P->dof_armature[lane] : 0.0;
  const double pf_stiff = lane < HMJ_NV ? P->jnt_stiffness[M.dof_jntid[lane]] : 0.0;
  const double pf_gain = lane < HMJ_NU ? P->act_gainprm0[lane] : 0.0;
  const double pf_fr0 = lane < HMJ_NU ? P->act_forcerange[lane][0] : 0.0;
  const double pf_fr1 = lane < HMJ_NU ? P->act_forcerange[lane][1] : 0.0;
  const int pf_xb = B.xfrc_body[env];
  for (int t = lane; t < HMJ_NBODY * 3; t += 32) (&S.xipos[0][0])[t] = t < 3 ? 0.0 : B.xipos[(size_t)env * HMJ_NBODY * 3 + t];
  for (int t = lane; t < HMJ_NV * 6; t += 32) (&S.cdof[0][0])[t] = B.cdof[(size_t)env * HMJ_NV * 6 + t];
  if (lane < 3) S.com[lane] = B.com[(size_t)env * 3 + lane];
  if (lane == 0) {
    for (int k = 0; k < 6; k++) { S.cvel[0][k] = 0; S.cfrc[0][k] = 0; }
    S.cacc[0][0] = S.cacc[0][1] = S.cacc[0][2] = 0;
    S.cacc[0][3] = -M.gravity[0]; S.cacc[0][4] = -M.gravity[1]; S.cacc[0][5] = -M.gravity[2];
    for (int k = 0; k < 10; k++) S.cinert[0][k] = 0;
  }
  __syncwarp();
  const double mass = pf_mass;
  if (lane >= 1 && lane < HMJ_NBODY) {
    const double* gm = B.ximat + ((size_t)env * HMJ_NBODY + lane) * 9;
    double ximat[9];
    for (int k = 0; k < 9; k++) ximat[k] = gm[k];
    double off[3] = {S.xipos[lane][0] - S.com[0], S.xipos[lane][1] - S.com[1], S.xipos[lane][2] - S.com[2]};
    inert_com(S.cinert[lane], M.body_inertia[lane], ximat, off, mass);
  }
  __syncwarp();

  // ---------------- mj_crb: composite inertia = sum of cinert over the body's subtree; sparse M
  if (lane >= 1 && lane < HMJ_NBODY) {
    double acc[10];
    for (int k = 0; k < 10; k++) acc[k] = S.cinert[lane][k];
    for (int c = lane + 1; c < T->subtree_end[lane]; c++)
      for (int k = 0; k < 10; k++) acc[k] += S.cinert[c][k];
    for (int k = 0; k < 10; k++) S.crb[lane][k] = acc[k];
  }
  __syncwarp();
  if (lane < HMJ_NV) d_mul_inert_vec(S.buf[lane], S.crb[M.dof_bodyid[lane]], S.cdof[lane]);
  __syncwarp();
#pragma unroll
  for (int kk = 0; kk < (HMJ_NM + 31) / 32; kk++) {   // uniform trip count: the armature comes from its owner lane
    const int e = lane + 32 * kk;
    const bool ok = e < HMJ_NM;
    const int i = ok ? T->M_row[e] : 0, j = ok ? T->M_col[e] : 0;
    const double arm = __shfl_sync(FULL, pf_arm, i);
    if (ok) {
      double v = d_dot6(S.cdof[j], S.buf[i]);
      if (i == j) v += arm;
      B.qM[(size_t)env * HMJ_NM + e] = v;
    }
  }

  // ---------------- mj_comVel (cvel, cdof_dot) and the acceleration pass of mj_rne, by level
  for (int L = 1; L < T->nlevel; L++) {
    if (lane < T->level_n[L]) {
      const int b = T->level_body[L][lane], pid = M.body_parentid[b];
      const int bda = M.body_dofadr[b], dn = M.body_dofnum[b];
      double cvel[6], cacc[6];
      for (int k = 0; k < 6; k++) { cvel[k] = S.cvel[pid][k]; cacc[k] = S.cacc[pid][k]; }
      if (dn == 6) {
        for (int j = 0; j < 3; j++) {
          for (int k = 0; k < 6; k++) { S.cdof_dot[bda + j][k] = 0; cvel[k] += S.cdof[bda + j][k] * S.qvel[bda + j]; }
        }
        for (int j = 3; j < 6; j++) d_cross_motion(S.cdof_dot[bda + j], cvel, S.cdof[bda + j]);
        for (int j = 3; j < 6; j++)
          for (int k = 0; k < 6; k++) cvel[k] += S.cdof[bda + j][k] * S.qvel[bda + j];
      } else {
        for (int j = 0; j < dn; j++) {
          d_cross_motion(S.cdof_dot[bda + j], cvel, S.cdof[bda + j]);
          for (int k = 0; k < 6; k++) cvel[k] += S.cdof[bda + j][k] * S.qvel[bda + j];
        }
      }
      for (int j = 0; j < dn; j++)
        for (int k = 0; k < 6; k++) cacc[k] += S.cdof_dot[bda + j][k] * S.qvel[bda + j];
      for (int k = 0; k < 6; k++) { S.cvel[b][k] = cvel[k]; S.cacc[b][k] = cacc[k]; }
    }
    __syncwarp();
  }

  // ---------------- mj_rne: body forces, subtree accumulation, projection on dofs
  if (lane >= 1 && lane < HMJ_NBODY) {
    double f[6], t[6], t1[6];
    d_mul_inert_vec(f, S.cinert[lane], S.cacc[lane]);
    d_mul_inert_vec(t, S.cinert[lane], S.cvel[lane]);
    d_cross_force(t1, S.cvel[lane], t);
    for (int k = 0; k < 6; k++) S.cfrc[lane][k] = f[k] + t1[k];
  }
  __syncwarp();
  if (lane >= 1 && lane < HMJ_NBODY) {
    double acc[6];
    for (int k = 0; k < 6; k++) acc[k] = S.cfrc[lane][k];
    for (int c = lane + 1; c < T->subtree_end[lane]; c++)
      for (int k = 0; k < 6; k++) acc[k] += S.cfrc[c][k];
    for (int k = 0; k < 6; k++) S.crb[lane][k] = acc[k];  // reuse crb rows as accumulated cfrc
  }
  __syncwarp();

  // ---------------- mj_passive, mj_fwdActuation
  if (lane < HMJ_NV) {
    const int d = lane, j = M.dof_jntid[d];
    const double stiff = pf_stiff;
    double spring = 0.0;
    if (stiff != 0) {
      const int padr = M.jnt_qposadr[j], dadr = M.jnt_dofadr[j];
      if (M.jnt_type[j] == HMJ_JNT_FREE) {
        const int k = d - dadr;
        if (k < 3) {
          spring = -stiff * (S.qpos[padr + k] - M.qpos_spring[padr + k]);
        } else {  // mju_subQuat(qpos quat, spring quat) as a 3D velocity
          double quat[4] = {S.qpos[padr + 3], S.qpos[padr + 4], S.qpos[padr + 5], S.qpos[padr + 6]};
          double qneg[4] = {M.qpos_spring[padr + 3], -M.qpos_spring[padr + 4], -M.qpos_spring[padr + 5], -M.qpos_spring[padr + 6]};
          double qdif[4], axis[3];
          d_normalize4(quat);
          d_mul_quat(qdif, qneg, quat);
          axis[0] = qdif[1]; axis[1] = qdif[2]; axis[2] = qdif[3];
          double sin_a_2 = d_normalize3(axis);
          double speed = 2 * atan2(sin_a_2, qdif[0]);
          if (speed > HMJ_PI) speed -= 2 * HMJ_PI;
          spring = -stiff * (axis[k - 3] * speed);
        }
      } else {
        spring = -stiff * (S.qpos[padr] - M.qpos_spring[padr]);
      }
    }
    S.qfrc_passive[d] = spring + (-M.dof_damping[d] * S.qvel[d]);
    S.qfrc_bias[d] = d_dot6(S.cdof[d], S.crb[M.dof_bodyid[d]]);
    S.qfrc_actuator[d] = 0.0;
  }
  __syncwarp();
  if (lane < HMJ_NU) {
    double c = fmin(fmax(S.ctrl[lane], M.act_ctrlrange[lane][0]), M.act_ctrlrange[lane][1]);
    double force = pf_gain * c;
    force = fmin(fmax(force, pf_fr0), pf_fr1);
    S.qfrc_actuator[M.act_dofadr[lane]] = M.act_gear[lane] * force;
  }
  __syncwarp();

  // ---------------- qfrc_smooth = passive - bias + actuator + J' xfrc_applied
  if (lane < HMJ_NV) {
    const int d = lane;
    double f = S.qfrc_passive[d] - S.qfrc_bias[d];
    f += S.qfrc_actuator[d];
    double ext = S.qfrc_actuator[d];
    const int xb = pf_xb;
    if (xb > 0 && xb < HMJ_NBODY) {
      const double* w = B.xfrc + (size_t)env * 6;
      if ((w[0] != 0 || w[1] != 0 || w[2] != 0 || w[3] != 0 || w[4] != 0 || w[5] != 0) && ((T->body_dofmask[xb] >> d) & 1u)) {
        double off[3] = {S.xipos[xb][0] - S.com[0], S.xipos[xb][1] - S.com[1], S.xipos[xb][2] - S.com[2]};
        double tmp[3];
        d_cross(tmp, S.cdof[d], off);
        const double fa = (S.cdof[d][3] + tmp[0]) * w[0] + (S.cdof[d][4] + tmp[1]) * w[1] + (S.cdof[d][5] + tmp[2]) * w[2];
        const double fb = S.cdof[d][0] * w[3] + S.cdof[d][1] * w[4] + S.cdof[d][2] * w[5];
        f += fa;
        f += fb;
        ext += fa;
        ext += fb;
      }
    }
    if (plan_mode & HMJ_PLAN_XFRC) {   // xfrc_applied on every body (mj_xfrcAccumulate): force and torque at the body's CoM
      const double* wa = B.xfrc_all + (size_t)env * HMJ_NBODY * 6;
      for (int bb = 1; bb < HMJ_NBODY; bb++) {
        if (!((T->body_dofmask[bb] >> d) & 1u)) continue;
        const double* w = wa + bb * 6;
        double off[3] = {S.xipos[bb][0] - S.com[0], S.xipos[bb][1] - S.com[1], S.xipos[bb][2] - S.com[2]};
        double tmp[3];
        d_cross(tmp, S.cdof[d], off);
        const double fa = (S.cdof[d][3] + tmp[0]) * w[0] + (S.cdof[d][4] + tmp[1]) * w[1] + (S.cdof[d][5] + tmp[2]) * w[2];
        const double fb = S.cdof[d][0] * w[3] + S.cdof[d][1] * w[4] + S.cdof[d][2] * w[5];
        f += fa; f += fb;
        ext += fa; ext += fb;
      }
    }
    B.qfrc_smooth[(size_t)env * HMJ_NV + d] = f;
    B.qfrc_ext[(size_t)env * HMJ_NV + d] = ext;
  }

  // ---------------- planner: sensors -> Stand residual -> cost accumulation
  if (plan_mode & 1) {
    const double* sens = B.sens + (size_t)env * SENS_NUM;
    // mj_subtreeVel for the root: whole-robot CoM velocity (torso_subtreelinvel is attached to body 1)
    double lx = 0, ly = 0, lz = 0;
    if (lane >= 1 && lane < HMJ_NBODY) {
      double dx[3] = {S.xipos[lane][0] - S.com[0], S.xipos[lane][1] - S.com[1], S.xipos[lane][2] - S.com[2]};
      double dv[3];
      d_cross(dv, S.cvel[lane], dx);
      lx = mass * (S.cvel[lane][3] + dv[0]); ly = mass * (S.cvel[lane][4] + dv[1]); lz = mass * (S.cvel[lane][5] + dv[2]);
    }
    lx = warp_sum(lx); ly = warp_sum(ly); lz = warp_sum(lz);
    const double sinv = 1.0 / fmax(HMJ_MINVAL, M.body_subtreemass[1]);
    lx *= sinv; ly *= sinv; (void)lz;
    if (lane == 0) {
      // torso_subtreecom is attached to body torso (no children): m2 * xipos2 / subtreemass2
      double s2 = M.body_subtreemass[2] < HMJ_MINVAL ? 1.0 : P->body_mass[2] / fmax(HMJ_MINVAL, M.body_subtreemass[2]);
      double comx = S.xipos[2][0] * s2, comy = S.xipos[2][1] * s2;
      if (M.body_subtreemass[2] < HMJ_MINVAL) { comx = S.xipos[2][0]; comy = S.xipos[2][1]; }
      const double* f1 = sens + SENS_SITE_XPOS;
      const double* f2 = sens + SENS_SITE_XPOS + 3;
      S.residual[0] = (sens[SENS_TORSO_XPOS + 2] - 0.5 * (f1[2] + f2[2])) - task->height_goal;
      double cpx = comx + lx * 0.2, cpy = comy + ly * 0.2;
      double fx = (f1[0] + f2[0]) * 0.5 - cpx, fy = (f1[1] + f2[1]) * 0.5 - cpy;
      S.residual[1] = sqrt(fx * fx + fy * fy);
      S.residual[2] = lx;
      S.residual[3] = ly;
    }
    if (lane < 12) { S.residual[4 + lane] = S.qvel[6 + lane]; S.residual[16 + lane] = S.ctrl[lane]; }
    __syncwarp();
    if (lane < HMJ_NRESIDUAL) B.residual[(size_t)env * HMJ_NRESIDUAL + lane] = S.residual[lane];
    if (lane < 3) B.trace[(size_t)env * 3 + lane] = sens[SENS_TORSO_XPOS + lane];
    // CostTerms: one lane per term, then sum in term order (task.cc:71-110)
    double term = 0.0;
    if (lane < task->nterm) {
      int shift = 0;
      for (int k = 0; k < lane; k++) shift += task->dim[k];
      term = task->weight[lane] * norm_value(S.residual + shift, task->normprm[lane], task->dim[lane], task->norm[lane]);
    }
    double cost = 0.0;
    for (int k = 0; k < task->nterm; k++) cost += __shfl_sync(FULL, term, k);
    if (fabs(task->risk) >= 1e-6) cost = (exp(task->risk * cost) - 1.0) / task->risk;
    if (lane == 0) { B.cost_sum[env] += cost; B.cost_last[env] = cost; }
  }
}

// k_kin on `stream`; k_dyn behind it on ctx->side[0] (where hmj_launch_factor queues k_factor next), or on `stream` itself
// when there is no side stream.  hmj_launch_solve joins the side stream.
void hmj_launch_smooth(const hmj_tables_t* tab, hmj_buffers_t buf, const hmj_task_t* task, int plan_mode,
                       const double* spline_times, const double* spline_values, int nnode, int interpolation,
                       int values_per_env, cudaStream_t stream, const hmj_solve_ctx_t* ctx) {
  static bool configured[HMJ_MAX_DEVICES] = {false};   // function attributes are per device
  const size_t smem = sizeof(DynSmem) * HMJ_WARPS_PER_CTA;
  const int dev = hmj_current_device();
  if (!configured[dev]) {
    cudaFuncSetAttribute(k_dyn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured[dev] = true;
  }
  const int grid = (buf.n + HMJ_WARPS_PER_CTA - 1) / HMJ_WARPS_PER_CTA;
  k_kin<<<grid, 32 * HMJ_WARPS_PER_CTA, 0, stream>>>(tab, buf, plan_mode, spline_times, spline_values, nnode, interpolation,
                                                      values_per_env);
  cudaStream_t s2 = stream;
  if (ctx) {
    cudaEventRecord(ctx->fac_fork, stream);
    cudaStreamWaitEvent(ctx->side[0], ctx->fac_fork, 0);
    s2 = ctx->side[0];
  }
  k_dyn<<<grid, 32 * HMJ_WARPS_PER_CTA, smem, s2>>>(tab, buf, task, plan_mode);
}
