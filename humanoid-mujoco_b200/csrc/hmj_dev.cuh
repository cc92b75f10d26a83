// hmj_dev.cuh — device-side tables, per-batch buffers and small math shared by the sm_100a kernels.
//
// Thread mapping everywhere: ONE WARP PER ENVIRONMENT (or per collision query), 4 warps per CTA.
// Per-env arrays are separate device arrays ("structure of arrays" across fields), each laid out
// env-major so a warp's read of its environment's row is one contiguous, coalesced segment.
// All intermediates that cross kernels (geom frames, cdof, qM, qfrc_smooth, contact slots) stay
// L2-resident: ~13 KB/env, 53 MB at 4096 envs against 126 MB of L2.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/hmj_model.h"
#include "../../include/hmj_sincos.h"

#define HMJ_WARPS_PER_CTA 4
#define HMJ_NSLOT (9 * HMJ_MAXCONPAIR + 28) /* contact slots per env: 9 hfield pairs x cap + 28 mesh pairs */
#define HMJ_SLOT_DOUBLES 8                  /* dist, pos[3], normal[3], valid */
#define HMJ_FAC_DOUBLES 192                 /* per-env Euler factor: 171 lower-triangle entries + 18 reciprocal pivots, padded */
#define HMJ_MESH_PAD 4
#define HMJ_NGROUP (HMJ_NMESH + 28)           /* narrowphase task groups: 7 hulls (hfield prisms) + 28 mesh-mesh pairs */
#define HMJ_NVERT_SMEM (HMJ_MAXVERT + HMJ_NMESH * HMJ_MESH_PAD)

#define HMJ_MAX_DEVICES 64
static inline int hmj_current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return (dev >= 0 && dev < HMJ_MAX_DEVICES) ? dev : 0;
}

#define HMJ_MINVAL 1e-15
#define HMJ_MAXVAL 1e10
#define HMJ_PI 3.14159265358979323846

// Host-derived lookup tables (built once in hmj_batch_create, read through L1/L2).
#define HMJ_SUP_G 8                                  /* cells per cube-face edge */
#define HMJ_SUP_CELLS (6 * HMJ_SUP_G * HMJ_SUP_G)
#define HMJ_SUP_CAP 16384                            /* candidate entries over all meshes (this model: ~7100) */
#define HMJ_SUP_DELTA 1e-3                           /* cells overlap by this much in cube-map coordinates: covers the fp32 cell look-up */
struct hmj_tables_t {
  hmj_model_t m;
  // kinematic levels (bodies of equal depth are independent)
  int32_t nlevel;
  int32_t level_n[8];
  int32_t level_body[8][8];
  int32_t subtree_end[HMJ_NBODY];  // bodies [b, subtree_end[b]) form b's subtree (DFS numbering)
  // sparse inertia index maps: entry e of qM is M(M_row[e], M_col[e]), col an ancestor-or-self dof of row
  int32_t M_row[HMJ_NM];
  int32_t M_col[HMJ_NM];
  uint32_t body_dofmask[HMJ_NBODY];  // bit d set: dof d moves body b
  // per-dof joint info
  int32_t dof_hinge_jnt[HMJ_NV];  // joint id for hinge dofs, -1 for free-joint dofs
  // collision
  int32_t pair_slot[HMJ_NPAIR + 1];  // first contact slot of each candidate pair
  int32_t slot_pair[HMJ_NSLOT];      // inverse map
  int32_t pair_group[HMJ_NPAIR + 1]; // narrowphase task group of each pair (hfield pairs: mesh id of geom2)
  int32_t group_base[64];            // first task of group g in wq_task, in units of N; group_base[NGROUP] = total
  int32_t broad_order[HMJ_NPAIR + 1]; // k_broad serves the pairs in this order: height-field pairs by hull size (descending), then mesh-mesh
  int32_t group_order[64];           // k_narrow hands out items group by group in this order: most expensive hull(s) first (LPT)
  int32_t group_shiftcap[64];        // log2 of the largest item (queries per warp) of group g: 2 for big hulls ... 5 for small ones
  // convex hull vertices, SoA float32, each mesh padded to a multiple of 4 with copies of its vertex 0;
  // mesh_adr[HMJ_NMESH] = total padded count (k_narrow stages them in shared memory as fp64)
  int32_t mesh_adr[HMJ_NMESH + 1];
  int32_t mesh_npad[HMJ_NMESH + 1];
  float vx[HMJ_MAXVERT + HMJ_NMESH * HMJ_MESH_PAD];
  float vy[HMJ_MAXVERT + HMJ_NMESH * HMJ_MESH_PAD];
  float vz[HMJ_MAXVERT + HMJ_NMESH * HMJ_MESH_PAD];
  // the same vertices as fp64 (what the kernels stage in shared memory: a plain 16-byte copy, no conversion)
  alignas(16) double dvx[HMJ_MAXVERT + HMJ_NMESH * HMJ_MESH_PAD];
  alignas(16) double dvy[HMJ_MAXVERT + HMJ_NMESH * HMJ_MESH_PAD];
  alignas(16) double dvz[HMJ_MAXVERT + HMJ_NMESH * HMJ_MESH_PAD];
  double mesh_lo[HMJ_NMESH + 1][3];  // local bounding box of each hull (exact min / max of its vertices)
  double mesh_hi[HMJ_NMESH + 1][3];
  // support candidates by direction cell (hmj_api.cu: build_support_cells): the directions of space are cut into
  // 6 x G x G cells (cube map); sup_vert[sup_start[mesh][c] .. sup_start[mesh][c + 1]) lists, in ascending order, every
  // vertex of the hull that can be the support point for SOME direction of cell c (a strict superset, built with a margin
  // far above rounding), so the arg-max over the list is the arg-max over the hull, bit for bit.
  int32_t tree_chol;                                     // 1: dofs are a free root + four 3-hinge chains (hmj_solve.cu: tree_factor_solve)
  int32_t sup_ok;                                        // 0: tables not built (capacity): kernels scan the whole hull
  uint16_t sup_start[HMJ_NMESH][HMJ_SUP_CELLS + 2];      // offsets into sup_vert
  uint16_t sup_vert[HMJ_SUP_CAP];                        // vertex indices within the mesh
  // lower-triangle index maps for the 18x18 dense routines
  int32_t tri_i[HMJ_NV * (HMJ_NV + 1) / 2];
  int32_t tri_j[HMJ_NV * (HMJ_NV + 1) / 2];
};

// planner task description pushed at plan time (mjpc Task -> numeric tables)
struct hmj_task_t {
  int32_t nterm;
  int32_t dim[HMJ_NTERM + 1];
  int32_t norm[HMJ_NTERM + 1];
  double weight[HMJ_NTERM + 1];
  double normprm[HMJ_NTERM + 1][HMJ_MAXNORMPRM];
  double risk;
  double height_goal;
};

// Device buffers of one batch (all pointers are device pointers; N = n_envs).
struct hmj_buffers_t {
  int32_t n;
  const int32_t* active;  // optional [N]: rows with 0 are skipped by every stage (episode-pool rows that are not being generated)
  // ---- state
  double* qpos;       // [N][19]
  double* qvel;       // [N][18]
  double* time;       // [N]
  double* warmstart;  // [N][18]
  double* ctrl;       // [N][12]
  double* xfrc;       // [N][6]   Cartesian wrench on one body (cpu_env.py:618-654)
  int32_t* xfrc_body; // [N]
  double* xfrc_all;   // [N][15][6] wrench on EVERY body (planner robustness rollouts: OU noise, mjpc/trajectory.cc:147-155); used with HMJ_PLAN_XFRC
  hmj_params_t* params;  // [N]
  // ---- cross-kernel scratch (L2 resident)
  double* geom_xpos;  // [N][10][3]
  double* geom_xmat;  // [N][10][9]
  double* cdof;       // [N][18][6]
  double* com;        // [N][3]     subtree_com of the root body
  double* xipos;      // [N][15][3] inertial frames, k_kin -> k_dyn
  double* ximat;      // [N][15][9]
  double* qM;         // [N][117]
  double* qfrc_smooth;// [N][18]
  double* euler_factor; // [N][192]   Cholesky factor of M + h diag(damping): lower triangle, reciprocal diagonal
  double* sens;       // [N][16]    torso xpos(3), xquat body1 (4), xpos body1 z, site z..., see SENS_*
  double* slots;      // [N][NSLOT][8]  dist, pos[3], normal[3], (unused)
  uint32_t* slot_mask;// [N][4]         bit s set: slot s holds a contact of this step (zeroed before k_broad)
  int32_t* wq_count;  // [64]       narrowphase tasks per group (zeroed every step)
  int2* wq_task;      // [group_base[NGROUP] * N]  (env, packed pair / slot ordinal / prism position)
  int32_t* solve_count; // [4]     envs per solver class (zeroed every step)
  int32_t* solve_list;  // [4][N]  env ids of each solver class
  // ---- outputs
  double* qacc;            // [N][18]
  double* qacc_smooth;     // [N][18]
  double* qfrc_constraint; // [N][18]
  double* qfrc_ext;        // [N][18]  qfrc_actuator + J' xfrc_applied (k_dyn), for mj_inverse
  double* qfrc_inverse;    // [N][18]  output of hmj_batch_inverse
  hmj_contact_t* contact;  // [N][MAXCON]
  int32_t* ncon;           // [N]
  int32_t* nefc;           // [N]
  int32_t* niter;          // [N]
  int32_t* status;         // [N]
  // ---- planner accumulators
  double* residual;   // [N][28] residual of the latest forward pass
  double* cost_sum;   // [N]     running sum of CostValue(residual_t)
  double* cost_last;  // [N]     CostValue of the latest forward pass
  double* trace;      // [N][3]
};

// Device-resident CPUEnv bookkeeping (simulation/cpu_env.py:135-187 delays, :618-674 pushes / action path):
// the per-step host logic of CPUEnv.step lives here so a control step is H2D(actions) -> kernels -> D2H(results).
#define HMJ_ENV_HIST 26  /* delay FIFO depth: MAX_DELAY / TIMESTEP + 1 */
struct hmj_envstate_t {
  double* act_noise;     // [N][12]  per-episode constant action noise (the JAX key is never split, cpu_env.py:664)
  double* obs_noise;     // [N][31]  q12 qd12 gyro3 quat4
  int32_t* act_delay;    // [N]      action delay in steps
  int32_t* obs_delay;    // [N][3]   joint (angles+velocities) / gyro / gravity delays in steps
  int32_t* act_count;    // [N]
  int32_t* obs_count;    // [N]
  double* act_hist;      // [N][HIST][12]
  double* obs_hist;      // [N][HIST][30]
  double* prev_action;   // [N][12]
  double* latest_action; // [N][12]
  double* force_sched;   // [N][5]   start, duration, magnitude, dir x, dir y
  int32_t* force_body;   // [N]
  int32_t* force_draws;  // [N]      renewals so far (counter of the per-env random stream)
  double* torso_quat;    // [N][4]   noisy root quaternion (CPUEnv.torso_quat, read by controllers/mpc_in_env.py:25)
};

// Result block of a control step (CPUEnv.step's return values), rows indexed by the env id of the whole batch.
struct hmj_results_t {
  double* obs;           // [N][30]
  double* reward;        // [N]
  double* time;          // [N]
  double* quat;          // [N][4]  CPUEnv.torso_quat
  int32_t* terminated;   // [N]
  int32_t* truncated;    // [N]
  int32_t* self_collision; // [N]
};
// Episode pool (hmj_reset.cu): S spare rows behind the N envs of a batch hold pre-validated fresh episodes.
#define HMJ_POOL_EMPTY 0     /* row free: the next generation round draws a new episode into it */
#define HMJ_POOL_PROBING 1   /* being generated: the round's kernels run on it (active mask) */
#define HMJ_POOL_PENDING 2   /* accepted by the rejection rule; becomes READY when the main stream joins the round */
#define HMJ_POOL_READY 3     /* can be taken by an env whose episode ended */
#define HMJ_POOL_TAKEN 4     /* assigned this step, being copied */
#define HMJ_POOL_FREED 5     /* consumed; becomes EMPTY at the step that launches the next round (deterministic membership) */
#define HMJ_POOL_REJECTED 6  /* failed the rejection rule; becomes EMPTY at the step that launches the next round */
#define HMJ_POOL_MAXTERM 64  /* terminal observations carried in the per-step host block (more: fetched on demand) */
struct hmj_pool_t {
  int32_t S;
  int32_t* state;        // [S]
  int32_t* active;       // [S]  the spare view's `active` mask
  int32_t* gen;          // [S]  episodes drawn so far in this row (counter block of the random stream)
  double* obs;           // [S][30] reset's return value (second _get_obs) of the pooled episode
  double* reward;        // [S]  scratch of the probe step's epilogue
  int32_t* term;         // [S]
  int32_t* trunc;        // [S]
  int32_t* selfc;        // [S]
  int32_t* pending;      // [N]  env waits for a fresh episode
  int32_t* assign;       // [N]  spare row taken this step, -1 none
  int32_t* served_list;  // [N]  envs served this step, in order
  int32_t* episode_src;  // [N]  row | generation << 16 of the env's current episode
  int32_t* ended_ids_full; // [N]  envs that needed an episode this step, in env order
  double* term_obs_full;   // [N][30] their terminal observations
  // host block extension (device mirror of the pinned block): counts[8], ended_ids[64], term_obs[64][30]
  int32_t* counts;
  int32_t* ended_ids;
  double* term_obs;
};

// Counter-based uniform in [0, 1): splitmix64 of (seed, env, counter).  Restated in numpy by env.py for the
// host mirror; replaces the reference's unseeded Python `random` stream (unpinned).
__host__ __device__ __forceinline__ uint64_t hmj_splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  uint64_t z = x;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ double hmj_rand_u01(uint64_t seed, uint32_t env, uint32_t counter) {
  const uint64_t h = hmj_splitmix64(seed ^ hmj_splitmix64(((uint64_t)env << 32) | counter));
  return (double)(h >> 11) * (1.0 / 9007199254740992.0);
}

#define SENS_TORSO_XPOS 0   /* xpos[2] (3) */
#define SENS_ROOT_XQUAT 3   /* xquat[1] (4) */
#define SENS_ROOT_Z 7       /* xpos[1][2] */
#define SENS_SITE_XPOS 8    /* site_xpos[0..1] (6): the feet, for the planner residual */
#define SENS_NUM 16

// ------------------------------------------------------------------ small math (device)
__device__ __forceinline__ void d_mul_quat(double* r, const double* a, const double* b) {
  double t0 = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double t1 = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double t2 = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double t3 = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = t0; r[1] = t1; r[2] = t2; r[3] = t3;
}

__device__ __forceinline__ void d_quat2mat(double* r, const double* q) {
  double q00 = q[0] * q[0], q01 = q[0] * q[1], q02 = q[0] * q[2], q03 = q[0] * q[3];
  double q11 = q[1] * q[1], q12 = q[1] * q[2], q13 = q[1] * q[3];
  double q22 = q[2] * q[2], q23 = q[2] * q[3], q33 = q[3] * q[3];
  r[0] = q00 + q11 - q22 - q33;
  r[4] = q00 - q11 + q22 - q33;
  r[8] = q00 - q11 - q22 + q33;
  r[1] = 2 * (q12 - q03);
  r[2] = 2 * (q13 + q02);
  r[3] = 2 * (q12 + q03);
  r[5] = 2 * (q23 - q01);
  r[6] = 2 * (q13 - q02);
  r[7] = 2 * (q23 + q01);
}

__device__ __forceinline__ void d_rot_vec_quat(double* res, const double* v, const double* q) {
  double t0 = q[0] * v[0] + q[2] * v[2] - q[3] * v[1];
  double t1 = q[0] * v[1] + q[3] * v[0] - q[1] * v[2];
  double t2 = q[0] * v[2] + q[1] * v[1] - q[2] * v[0];
  double r0 = v[0] + 2 * (q[2] * t2 - q[3] * t1);
  double r1 = v[1] + 2 * (q[3] * t0 - q[1] * t2);
  double r2 = v[2] + 2 * (q[1] * t1 - q[2] * t0);
  res[0] = r0; res[1] = r1; res[2] = r2;
}

__device__ __forceinline__ void d_rot_vec_mat(double* res, const double* v, const double* m) {
  double r0 = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
  double r1 = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
  double r2 = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  res[0] = r0; res[1] = r1; res[2] = r2;
}

__device__ __forceinline__ void d_cross(double* r, const double* a, const double* b) {
  double r0 = a[1] * b[2] - a[2] * b[1];
  double r1 = a[2] * b[0] - a[0] * b[2];
  double r2 = a[0] * b[1] - a[1] * b[0];
  r[0] = r0; r[1] = r1; r[2] = r2;
}

__device__ __forceinline__ double d_normalize3(double* v) {
  double n = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  if (n < HMJ_MINVAL) {
    v[0] = 1; v[1] = 0; v[2] = 0;
  } else {
    double s = 1.0 / n;
    v[0] *= s; v[1] *= s; v[2] *= s;
  }
  return n;
}

__device__ __forceinline__ double d_normalize4(double* q) {
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < HMJ_MINVAL) {
    q[0] = 1; q[1] = 0; q[2] = 0; q[3] = 0;
  } else if (fabs(n - 1) > HMJ_MINVAL) {
    double s = 1.0 / n;
    q[0] *= s; q[1] *= s; q[2] *= s; q[3] *= s;
  }
  return n;
}

__device__ __forceinline__ void d_axis_angle2quat(double* q, const double* axis, double angle) {
  if (angle == 0) {
    q[0] = 1; q[1] = 0; q[2] = 0; q[3] = 0;
  } else {
    double s, c;
    hmj_sincos(angle * 0.5, &s, &c);  // shared deterministic sin/cos: same bits as the oracle
    q[0] = c; q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
  }
}

__device__ __forceinline__ void d_mul_inert_vec(double* res, const double* i, const double* v) {
  res[0] = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
  res[1] = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
  res[2] = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
  res[3] = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
  res[4] = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
  res[5] = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
}

__device__ __forceinline__ void d_cross_motion(double* res, const double* vel, const double* v) {
  res[0] = -vel[2] * v[1] + vel[1] * v[2];
  res[1] = vel[2] * v[0] - vel[0] * v[2];
  res[2] = -vel[1] * v[0] + vel[0] * v[1];
  res[3] = -vel[2] * v[4] + vel[1] * v[5] - vel[5] * v[1] + vel[4] * v[2];
  res[4] = vel[2] * v[3] - vel[0] * v[5] + vel[5] * v[0] - vel[3] * v[2];
  res[5] = -vel[1] * v[3] + vel[0] * v[4] - vel[4] * v[0] + vel[3] * v[1];
}

__device__ __forceinline__ void d_cross_force(double* res, const double* vel, const double* f) {
  res[0] = -vel[2] * f[1] + vel[1] * f[2] - vel[5] * f[4] + vel[4] * f[5];
  res[1] = vel[2] * f[0] - vel[0] * f[2] + vel[5] * f[3] - vel[3] * f[5];
  res[2] = -vel[1] * f[0] + vel[0] * f[1] - vel[4] * f[3] + vel[3] * f[4];
  res[3] = -vel[2] * f[4] + vel[1] * f[5];
  res[4] = vel[2] * f[3] - vel[0] * f[5];
  res[5] = -vel[1] * f[3] + vel[0] * f[4];
}

__device__ __forceinline__ double d_dot6(const double* a, const double* b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3] + a[4] * b[4] + a[5] * b[5];
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ bool d_is_bad(double x) { return isnan(x) || x > HMJ_MAXVAL || x < -HMJ_MAXVAL; }

// Side streams of hmj_launch_solve: the three solver classes are independent, so the (few, long) many-contact solves
// and the (many, short) unconstrained ones run beside the main class instead of after it.  Null -> one stream.
struct hmj_solve_ctx_t {
  cudaStream_t side[2];
  cudaEvent_t fork, join[2];
  cudaEvent_t fac_fork, fac_done;   // k_factor on side[0], forked after k_smooth, joined by hmj_launch_solve
};
// kernel launchers (defined in the .cu files, called from hmj_api.cu)
void hmj_launch_smooth(const hmj_tables_t* tab, hmj_buffers_t buf, const hmj_task_t* task, int plan_mode,
                       const double* spline_times, const double* spline_values, int nnode, int interpolation,
                       int values_per_env, cudaStream_t stream, const hmj_solve_ctx_t* ctx = nullptr);
void hmj_launch_collide(const hmj_tables_t* tab, hmj_buffers_t buf, cudaStream_t stream, bool zero = true);
// Contact-independent factorisations (qacc_smooth, Euler factor): call after hmj_launch_smooth, before hmj_launch_solve.
void hmj_launch_factor(const hmj_tables_t* tab, hmj_buffers_t buf, cudaStream_t stream, const hmj_solve_ctx_t* ctx = nullptr);
void hmj_launch_solve(const hmj_tables_t* tab, hmj_buffers_t buf, int integrate, cudaStream_t stream,
                      const hmj_solve_ctx_t* ctx = nullptr, bool zero = true);

// plan_mode bits for hmj_launch_smooth
// internal, transient status bit: k_kin found a bad state this step (mj_resetData) and k_dyn still has to zero the controls
// and the applied wrench of the env (k_kin leaves both alone so that their writers may run beside it); k_dyn clears it
#define HMJ_STATUS_RESET_PENDING (1 << 30)
#define HMJ_PLAN_COST 1    /* sensors -> residual -> cost accumulation */
#define HMJ_PLAN_POLICY 2  /* ctrl from the spline policy at data time */
#define HMJ_PLAN_XFRC 4    /* k_dyn also applies the per-body wrenches of xfrc_all */

// trajectory recording buffers (optional), layout [H][N][dim]
struct hmj_record_t {
  int32_t horizon;
  double* states;   // [H][N][37]
  double* actions;  // [H][N][12]
  double* times;    // [H][N]
  double* residual; // [H][N][28]
  double* costs;    // [H][N]
  double* trace;    // [H][N][3]
};

void hmj_launch_plan_init(hmj_buffers_t buf, const double* state_time_dev, cudaStream_t stream);
void hmj_launch_plan_ou(hmj_buffers_t buf, int t, int nsteps, double rate, double scale, const double* noise_dev,
                        const unsigned long long* seed_iter_dev, int first_global, cudaStream_t stream);
void hmj_launch_plan_candidates(const hmj_tables_t* tab, int n, int nnode, const double* nominal_dev, const double* noise_dev,
                                double std0, double std1, unsigned long long seed, unsigned long long iteration, int first_global,
                                double* values_dev, cudaStream_t stream, const double* std_vec_dev = nullptr);
void hmj_launch_plan_winner(const double* returns_dev, int n, const double* values_dev, int nnode, int first_global,
                            double* argmin_dev, double* rec_dev, cudaStream_t stream);
void hmj_launch_plan_select(const double* gathered_dev, int nranks, int stride, double* out_dev, cudaStream_t stream);
void hmj_launch_plan_record(hmj_buffers_t buf, hmj_record_t rec, int t, cudaStream_t stream);
void hmj_launch_plan_finish(hmj_buffers_t buf, int horizon, double* returns_dev, int32_t* failure_dev, cudaStream_t stream);
void hmj_launch_argmin(const double* returns_dev, int n, double* out_cost_idx_dev, cudaStream_t stream);
void hmj_launch_env_obs_reward(const hmj_tables_t* tab, hmj_buffers_t buf, const double* prev_action, const double* latest_action,
                               const double* noise, double max_sim_time, double* obs, double* reward, int32_t* terminated,
                               int32_t* truncated, int32_t* self_collision, cudaStream_t stream);
void hmj_launch_env_pre(hmj_buffers_t buf, hmj_envstate_t env, const double* action, const int32_t* mask, double rf,
                        unsigned long long seed, cudaStream_t stream, int stream_id_offset = 0);
void hmj_launch_env_reset(const hmj_tables_t* tab, hmj_buffers_t spare, hmj_envstate_t spare_env, hmj_pool_t pool, double rf,
                          double timestep, unsigned long long seed, int row0_global, cudaStream_t stream);
void hmj_launch_pool_validate(hmj_buffers_t spare, hmj_pool_t pool, cudaStream_t stream);
void hmj_launch_autoreset(hmj_buffers_t buf, hmj_envstate_t env, hmj_buffers_t spare, hmj_envstate_t spare_env, hmj_pool_t pool,
                          const int32_t* term, const int32_t* trunc, double* obs, double* time_out, double* quat_out, int flags /* 1 promote, 2 drain, 4 a round is launched right behind this step */,
                          cudaStream_t stream);
void hmj_launch_fd_setup(const hmj_tables_t* tab, hmj_buffers_t view, int nknot, const double* x_dev, const double* u_dev,
                         const double* time_dev, const double* ws_dev, double eps, int centered, int32_t* valid_dev,
                         cudaStream_t stream);
void hmj_launch_fd_diff(const hmj_tables_t* tab, hmj_buffers_t view, int nknot, const int32_t* valid_dev, double eps, double* A,
                        double* B, double* C, double* D, cudaStream_t stream);
void hmj_launch_pool_mark_pending(hmj_pool_t pool, const int32_t* ids_dev, int count, int n, cudaStream_t stream);
void hmj_launch_env_epilogue(const hmj_tables_t* tab, hmj_buffers_t buf, hmj_envstate_t env, double max_sim_time, hmj_results_t R,
                             int first, int count, cudaStream_t stream);
const void* hmj_env_pre_kernel();
void hmj_launch_env_pre_rows(hmj_buffers_t buf, hmj_envstate_t env, const double* action, double rf, unsigned long long seed,
                             int first, int count, cudaStream_t stream);
void hmj_launch_env_post(hmj_buffers_t buf, hmj_envstate_t env, const int32_t* mask, double* obs, double* time_out, double* quat_out,
                         cudaStream_t stream);
