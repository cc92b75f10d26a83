/* hmj_oracle.h — CPU oracle (TEST INFRASTRUCTURE ONLY — never linked into the product).
 *
 * A plain-C, fp64, single-environment restatement of the reference's hot path:
 *   - MuJoCo 3.1.4 mj_step / mj_forward for the humanoid model (third-party, pinned by
 *     mujoco_mpc/CMakeLists.txt:58-61; source NOT in /root/reference — restated from the
 *     published algorithm, API declared in simulation/mujoco/include/mujoco/mujoco.h:120,129,
 *     247-268,310-367) including libccd's MPR (bundled with MuJoCo) for mesh-mesh and
 *     mesh-heightfield narrowphase;
 *   - mjpc spline policy / residual / norm / cost / rollout
 *     (mjpc/spline/spline.cc:103-277, mjpc/planners/sampling/policy.cc:52-59,
 *      mjpc/tasks/humanoid_cap/stand/stand.cc:40-99, mjpc/norm.cc:50-210,
 *      mjpc/task.cc:71-110, mjpc/trajectory.cc:100-210,312-326).
 *
 * PARITY STATUS: the mjpc pieces (spline, norm) are pinned against the reference's own
 * sources compiled into oracle/_ref and against the reference tests' known answers;
 * the mj_step restatement is "parity unpinned" w.r.t. a MuJoCo binary (none exists in
 * this image) — it is the definition of expected output for the CUDA path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.
 */
#ifndef HMJ_ORACLE_H_
#define HMJ_ORACLE_H_

#include "hmj_model.h"

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAXCON HMJ_MAXCON
#define ORC_MAXEFC HMJ_MAXEFC

/* efc_type */
#define ORC_CNSTR_LIMIT_JOINT 3
#define ORC_CNSTR_CONTACT_PYRAMIDAL 6
/* efc_state */
#define ORC_STATE_SATISFIED 0
#define ORC_STATE_QUADRATIC 1

typedef struct orc_data_ {
  /* ---- state ---- */
  double qpos[HMJ_NQ];
  double qvel[HMJ_NV];
  double ctrl[HMJ_NU];
  double qacc_warmstart[HMJ_NV];
  double xfrc_applied[HMJ_NBODY][6];
  double time;
  /* ---- position stage ---- */
  double xpos[HMJ_NBODY][3];
  double xquat[HMJ_NBODY][4];
  double xmat[HMJ_NBODY][9];
  double xipos[HMJ_NBODY][3];
  double ximat[HMJ_NBODY][9];
  double xanchor[HMJ_NJNT][3];
  double xaxis[HMJ_NJNT][3];
  double geom_xpos[HMJ_NGEOM][3];
  double geom_xmat[HMJ_NGEOM][9];
  double site_xpos[HMJ_NSITE][3];
  double subtree_com[HMJ_NBODY][3];
  double cinert[HMJ_NBODY][10];
  double crb[HMJ_NBODY][10];
  double cdof[HMJ_NV][6];
  double qM[HMJ_NM];
  double qLD[HMJ_NM];
  double qLDiagInv[HMJ_NV];
  /* ---- velocity stage ---- */
  double cvel[HMJ_NBODY][6];
  double cdof_dot[HMJ_NV][6];
  double qfrc_spring[HMJ_NV];
  double qfrc_damper[HMJ_NV];
  double qfrc_passive[HMJ_NV];
  double qfrc_bias[HMJ_NV];
  double subtree_linvel[HMJ_NBODY][3];
  /* ---- actuation / acceleration ---- */
  double actuator_force[HMJ_NU];
  double qfrc_actuator[HMJ_NV];
  double qfrc_smooth[HMJ_NV];
  double qacc_smooth[HMJ_NV];
  double qacc[HMJ_NV];
  double qfrc_constraint[HMJ_NV];
  /* ---- constraints ---- */
  double efc_J[ORC_MAXEFC][HMJ_NV];
  double efc_pos[ORC_MAXEFC];
  double efc_margin[ORC_MAXEFC];
  double efc_diagApprox[ORC_MAXEFC];
  double efc_R[ORC_MAXEFC];
  double efc_D[ORC_MAXEFC];
  double efc_KBIP[ORC_MAXEFC][4];
  double efc_vel[ORC_MAXEFC];
  double efc_aref[ORC_MAXEFC];
  double efc_b[ORC_MAXEFC];
  double efc_force[ORC_MAXEFC];
  /* ---- task ---- */
  double residual[HMJ_NRESIDUAL];
  /* ---- counters (algorithmic work, FMA = 2 flops) ---- */
  double flops_smooth;
  double flops_collision;
  double flops_constraint;
  double flops_solver;
  double n_support;  /* support-function evaluations */
  double n_mpr;      /* MPR queries run */
  /* ---- contacts ---- */
  hmj_contact_t contact[ORC_MAXCON];
  int32_t efc_type[ORC_MAXEFC];
  int32_t efc_id[ORC_MAXEFC];
  int32_t efc_state[ORC_MAXEFC];
  int32_t contact_efc_address[ORC_MAXCON];
  int32_t ncon;
  int32_t nefc;
  int32_t solver_niter;
  int32_t solver_nls; /* total line-search evaluations */
  int32_t status;     /* HMJ_STATUS_* bits (MuJoCo warnings) */
  int32_t pad0;
} orc_data_t;

/* trajectory record written by orc_rollout (mjpc/trajectory.h:74-86) */

void orc_default_params(const hmj_model_t* m, hmj_params_t* p);
void orc_reset_data(const hmj_model_t* m, orc_data_t* d);

/* stage functions (MuJoCo names: mujoco.h:310-367) */
void orc_kinematics(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);
void orc_com_pos(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);
void orc_crb(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);
void orc_factor_m(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);
void orc_collision(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);
void orc_make_constraint(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);
void orc_fwd_velocity(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);
void orc_fwd_actuation(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);
void orc_fwd_acceleration(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);
void orc_fwd_constraint(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);
void orc_subtree_vel(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);
void orc_euler(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);

void orc_forward(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);
void orc_step(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d);
/* n lock-step steps with a fixed control; returns accumulated status */
int orc_step_n(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d, int n);

/* MPR on two hulls given explicit world frames (for unit tests of the narrowphase) */
int orc_mpr_mesh_mesh(const hmj_model_t* m, int mesh1, const double pos1[3], const double mat1[9],
                      int mesh2, const double pos2[3], const double mat2[9],
                      double* depth, double dir[3], double pos[3]);

/* ---- mjpc pieces ---- */
void orc_spline_sample(const double* times, const double* values, int nnode, int dim,
                       int interpolation, double time, double* out);
double orc_norm(const double* x, const double* params, int n, int type);
void orc_cost_terms(const hmj_model_t* m, const double* residual, double* terms);
double orc_cost_value(const hmj_model_t* m, const double* residual);
void orc_residual(const hmj_model_t* m, const orc_data_t* d, double* residual);

/* Trajectory::NoisyRollout with xfrc_std = 0 (mjpc/trajectory.cc:100-210) driven by a
 * SamplingPolicy spline (policy.cc:52-59). Arrays are row-major [horizon][dim].
 * Returns failure flag. */
int orc_rollout(const hmj_model_t* m, const hmj_params_t* p, const double* state, double time,
                const double* spline_times, const double* spline_values, int nnode, int interpolation,
                int horizon, double* states, double* actions, double* times, double* residual,
                double* costs, double* trace, double* total_return);

/* multi-threaded drivers (hmj_oracle_batch.c): CPU-baseline timing harness, return wall seconds */
double orc_step_batch(const hmj_model_t* m, const hmj_params_t* params, orc_data_t* data, int n_envs,
                      const double* ctrl, int ctrl_steps, int nsteps, int nthreads);
double orc_rollout_batch(const hmj_model_t* m, const hmj_params_t* p, const double* state, double time,
                         const double* spline_times, const double* spline_values, int n, int nnode,
                         int interpolation, int horizon, double* total_return, int32_t* failure, int nthreads);

/* mjd_transitionFD for one knot (hmj_oracle_fd.c): A [36][36], B [36][12], C [28][36], D [28][12] */
int orc_transition_fd(const hmj_model_t* m, const hmj_params_t* p, const double* x, const double* u, double time,
                      const double* warmstart, double eps, int centered, double* A, double* B, double* C, double* D);
int orc_rollout_noisy(const hmj_model_t* m, const hmj_params_t* p, const double* state, double time,
                      const double* spline_times, const double* spline_values, int nnode, int interpolation,
                      int horizon, double xfrc_std, double xfrc_rate, const double* noise, double* total_return);
/* mj_inverse for a given acceleration (hmj_oracle_fd.c) */
void orc_inverse(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d, const double* qacc, double* qfrc_inverse);
int orc_sizeof_data(void);
int orc_sizeof_model(void);
int orc_sizeof_params(void);

#ifdef __cplusplus
}
#endif
#endif /* HMJ_ORACLE_H_ */
