/* hmj_oracle_fd.c — CPU restatement of MuJoCo's finite-differenced transition matrices (mjd_transitionFD / mjd_stepFD,
 * declared at simulation/mujoco/include/mujoco/mujoco.h "Derivatives" section) as mjpc's ModelDerivatives::Compute calls it
 * for every knot of a trajectory (mjpc/planners/model_derivatives.cc:51-96).  TEST INFRASTRUCTURE: checker for
 * hmj_fd_transition (csrc/hmj_fd.cu).  Parity unpinned like the step itself (MuJoCo 3.1.4 is not available here); restated
 * from the published algorithm:
 *   y = step(x, u): next state in tangent coordinates (mj_differentiatePos for positions), s = sensordata computed by
 *   that step's forward pass (i.e. at the perturbed PRE-step state; here the 28 user sensors = the task residual);
 *   every perturbed step restarts from the same state, warm start and time; controls are nudged forward where the
 *   control range allows, backward otherwise (or both when centred); positions are perturbed with mj_integratePos.
 *   A [36][36] = dy/dx, B [36][12] = dy/du, C [28][36] = ds/dx, D [28][12] = ds/du, row-major. */
#include <math.h>
#include <string.h>

#include "hmj_oracle.h"
#include "hmj_oracle_internal.h"

#define NDX (2 * HMJ_NV)
#define NS HMJ_NRESIDUAL

static void mul_quat(double* r, const double* a, const double* b) {
  double t0 = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double t1 = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double t2 = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double t3 = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = t0; r[1] = t1; r[2] = t2; r[3] = t3;
}

/* mj_integratePos(m, qpos, dq, 1) */
static void integrate_pos(const hmj_model_t* m, double* qpos, const double* dq) {
  for (int j = 0; j < HMJ_NJNT; j++) {
    int padr = m->jnt_qposadr[j], vadr = m->jnt_dofadr[j];
    if (m->jnt_type[j] == HMJ_JNT_FREE) {
      for (int i = 0; i < 3; i++) qpos[padr + i] += dq[vadr + i];
      double w[3] = {dq[vadr + 3], dq[vadr + 4], dq[vadr + 5]}, qrot[4];
      double n = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
      if (n < 1e-15) { w[0] = 1; w[1] = 0; w[2] = 0; } else { w[0] /= n; w[1] /= n; w[2] /= n; }
      if (n == 0) { qrot[0] = 1; qrot[1] = qrot[2] = qrot[3] = 0; }
      else { double s = sin(n * 0.5), c = cos(n * 0.5); qrot[0] = c; qrot[1] = w[0] * s; qrot[2] = w[1] * s; qrot[3] = w[2] * s; }
      double* q = qpos + padr + 3;
      double qn = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
      if (qn > 1e-15 && fabs(qn - 1) > 1e-15) for (int i = 0; i < 4; i++) q[i] /= qn;
      mul_quat(q, q, qrot);
    } else {
      qpos[padr] += dq[vadr];
    }
  }
}

/* mj_differentiatePos(m, dq, h, q1, q2): dq = (q2 (-) q1) / h */
static void differentiate_pos(const hmj_model_t* m, double* dq, double h, const double* q1, const double* q2) {
  for (int j = 0; j < HMJ_NJNT; j++) {
    int padr = m->jnt_qposadr[j], vadr = m->jnt_dofadr[j];
    if (m->jnt_type[j] == HMJ_JNT_FREE) {
      for (int i = 0; i < 3; i++) dq[vadr + i] = (q2[padr + i] - q1[padr + i]) / h;
      const double* qa = q2 + padr + 3;
      const double* qb = q1 + padr + 3;
      double qneg[4] = {qb[0], -qb[1], -qb[2], -qb[3]}, qdif[4];
      mul_quat(qdif, qneg, qa);
      double axis[3] = {qdif[1], qdif[2], qdif[3]};
      double s = sqrt(axis[0] * axis[0] + axis[1] * axis[1] + axis[2] * axis[2]);
      if (s < 1e-15) { axis[0] = 1; axis[1] = 0; axis[2] = 0; } else { axis[0] /= s; axis[1] /= s; axis[2] /= s; }
      double speed = 2 * atan2(s, qdif[0]);
      if (speed > 3.14159265358979323846) speed -= 2 * 3.14159265358979323846;
      speed /= h;
      for (int i = 0; i < 3; i++) dq[vadr + 3 + i] = axis[i] * speed;
    } else {
      dq[vadr] = (q2[padr] - q1[padr]) / h;
    }
  }
}

typedef struct { double qpos[HMJ_NQ], qvel[HMJ_NV], sens[NS]; } fd_out_t;

static void run_step(const hmj_model_t* m, const hmj_params_t* p, const double* qpos, const double* qvel, const double* ctrl,
                     const double* ws, double time, fd_out_t* out) {
  orc_data_t d;
  orc_reset_data(m, &d);
  memcpy(d.qpos, qpos, sizeof(d.qpos));
  memcpy(d.qvel, qvel, sizeof(d.qvel));
  memcpy(d.ctrl, ctrl, sizeof(d.ctrl));
  memcpy(d.qacc_warmstart, ws, sizeof(d.qacc_warmstart));
  d.time = time;
  orc_step(m, p, &d);
  memcpy(out->qpos, d.qpos, sizeof(out->qpos));
  memcpy(out->qvel, d.qvel, sizeof(out->qvel));
  memcpy(out->sens, d.residual, sizeof(out->sens));
}

/* clampedStateDiff / clampedDiff: centred when both sides exist, one-sided otherwise, zero when neither */
static void column(const hmj_model_t* m, const fd_out_t* y0, const fd_out_t* yp, const fd_out_t* ym, double eps, double* dy,
                   double* ds) {
  const fd_out_t* a = yp ? yp : y0;
  const fd_out_t* b = ym ? ym : y0;
  const double h = (yp && ym) ? 2 * eps : eps;
  if (!yp && !ym) { memset(dy, 0, NDX * sizeof(double)); memset(ds, 0, NS * sizeof(double)); return; }
  differentiate_pos(m, dy, h, b->qpos, a->qpos);
  for (int i = 0; i < HMJ_NV; i++) dy[HMJ_NV + i] = (a->qvel[i] - b->qvel[i]) / h;
  for (int i = 0; i < NS; i++) ds[i] = (a->sens[i] - b->sens[i]) / h;
}

int orc_transition_fd(const hmj_model_t* m, const hmj_params_t* p, const double* x, const double* u, double time,
                      const double* warmstart, double eps, int centered, double* A, double* B, double* C, double* D) {
  fd_out_t y0, yp, ym;
  double ws[HMJ_NV] = {0};
  if (warmstart) memcpy(ws, warmstart, sizeof(ws));
  const double* qpos = x;
  const double* qvel = x + HMJ_NQ;
  run_step(m, p, qpos, qvel, u, ws, time, &y0);
  double dy[NDX], ds[NS];
  /* controls */
  for (int i = 0; i < HMJ_NU; i++) {
    double c[HMJ_NU];
    const double lo = m->act_ctrlrange[i][0], hi = m->act_ctrlrange[i][1];
    const int fwd = (u[i] + eps <= hi) && (u[i] + eps >= lo);
    if (fwd) { memcpy(c, u, sizeof(c)); c[i] += eps; run_step(m, p, qpos, qvel, c, ws, time, &yp); }
    const int back = (centered || !fwd) && (u[i] - eps >= lo) && (u[i] - eps <= hi);
    if (back) { memcpy(c, u, sizeof(c)); c[i] -= eps; run_step(m, p, qpos, qvel, c, ws, time, &ym); }
    column(m, &y0, fwd ? &yp : 0, back ? &ym : 0, eps, dy, ds);
    for (int r = 0; r < NDX; r++) B[r * HMJ_NU + i] = dy[r];
    for (int r = 0; r < NS; r++) D[r * HMJ_NU + i] = ds[r];
  }
  /* velocities */
  for (int i = 0; i < HMJ_NV; i++) {
    double v[HMJ_NV];
    memcpy(v, qvel, sizeof(v)); v[i] += eps;
    run_step(m, p, qpos, v, u, ws, time, &yp);
    if (centered) { memcpy(v, qvel, sizeof(v)); v[i] -= eps; run_step(m, p, qpos, v, u, ws, time, &ym); }
    column(m, &y0, &yp, centered ? &ym : 0, eps, dy, ds);
    for (int r = 0; r < NDX; r++) A[r * NDX + HMJ_NV + i] = dy[r];
    for (int r = 0; r < NS; r++) C[r * NDX + HMJ_NV + i] = ds[r];
  }
  /* positions (tangent space) */
  for (int i = 0; i < HMJ_NV; i++) {
    double q[HMJ_NQ], dq[HMJ_NV] = {0};
    memcpy(q, qpos, sizeof(q)); dq[i] = eps; integrate_pos(m, q, dq);
    run_step(m, p, q, qvel, u, ws, time, &yp);
    if (centered) { memcpy(q, qpos, sizeof(q)); dq[i] = -eps; integrate_pos(m, q, dq); run_step(m, p, q, qvel, u, ws, time, &ym); }
    column(m, &y0, &yp, centered ? &ym : 0, eps, dy, ds);
    for (int r = 0; r < NDX; r++) A[r * NDX + i] = dy[r];
    for (int r = 0; r < NS; r++) C[r * NDX + i] = ds[r];
  }
  return 0;
}

/* ------------------------------------------------------------------ mj_inverse (mujoco.h: mj_inverse; used by
 * controllers/mpc_utils.py:36-55): position / velocity stages, the constraint forces of the GIVEN acceleration
 * (mj_invConstraint: Jaref = J qacc - aref, f = -D Jaref on the active rows, no solver), and
 * qfrc_inverse = M qacc + qfrc_bias - qfrc_passive - qfrc_constraint.  Checker for hmj_batch_inverse. */
void orc_inverse(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d, const double* qacc, double* qfrc_inverse) {
  orc_kinematics(m, p, d);
  orc_com_pos(m, p, d);
  orc_crb(m, p, d);
  orc_factor_m(m, p, d);
  orc_collision(m, p, d);
  orc_make_constraint(m, p, d);
  orc_fwd_velocity(m, p, d);
  double force[ORC_MAXEFC];
  for (int r = 0; r < d->nefc; r++) {
    double jar = -d->efc_aref[r];
    for (int i = 0; i < HMJ_NV; i++) jar += d->efc_J[r][i] * qacc[i];
    force[r] = jar < 0 ? -d->efc_D[r] * jar : 0.0;
  }
  double Ma[HMJ_NV];
  orc_mul_m(m, d->qM, Ma, qacc);
  for (int i = 0; i < HMJ_NV; i++) {
    double fc = 0.0;
    for (int r = 0; r < d->nefc; r++) fc += d->efc_J[r][i] * force[r];
    d->qfrc_constraint[i] = fc;
    qfrc_inverse[i] = Ma[i] + d->qfrc_bias[i] - d->qfrc_passive[i] - fc;
  }
}
