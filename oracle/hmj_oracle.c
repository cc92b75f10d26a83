/* hmj_oracle.c — oracle: mj_forward / mj_step drivers and the mjpc rollout pieces.
 * TEST INFRASTRUCTURE ONLY (see hmj_oracle.h).
 */
#include <math.h>
#include <string.h>

#include "hmj_oracle_internal.h"

static int is_bad(double x) { return isnan(x) || x > MJ_MAXVAL || x < -MJ_MAXVAL; }

/* mj_resetData as seen by this path: state to qpos0, everything else cleared */
static void reset_state(const hmj_model_t* m, orc_data_t* d) {
  memcpy(d->qpos, m->qpos0, sizeof(d->qpos));
  memset(d->qvel, 0, sizeof(d->qvel));
  memset(d->ctrl, 0, sizeof(d->ctrl));
  memset(d->qacc_warmstart, 0, sizeof(d->qacc_warmstart));
  memset(d->xfrc_applied, 0, sizeof(d->xfrc_applied));
  memset(d->qacc, 0, sizeof(d->qacc));
  d->time = 0;
}

/* mj_forward (mujoco.h:129): position -> velocity -> actuation -> acceleration -> constraint */
void orc_forward(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  /* mj_checkCtrl-equivalent (mj_fwdActuation): bad controls are zeroed with a warning */
  for (int i = 0; i < HMJ_NU; i++)
    if (is_bad(d->ctrl[i])) {
      d->status |= HMJ_STATUS_BADCTRL;
      memset(d->ctrl, 0, sizeof(d->ctrl));
      break;
    }
  orc_kinematics(m, p, d);
  orc_com_pos(m, p, d);
  orc_crb(m, p, d);
  orc_factor_m(m, p, d);
  orc_collision(m, p, d);
  orc_make_constraint(m, p, d);
  orc_fwd_velocity(m, p, d);
  orc_subtree_vel(m, p, d); /* sensorVel: subtreelinvel (planner task model) */
  orc_fwd_actuation(m, p, d);
  orc_fwd_acceleration(m, p, d);
  orc_fwd_constraint(m, p, d);
  orc_residual(m, d, d->residual); /* sensorAcc: user-sensor callback (mjcb_sensor) */
}

/* mj_step (mujoco.h:120) with the Euler integrator */
void orc_step(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  for (int i = 0; i < HMJ_NQ; i++)
    if (is_bad(d->qpos[i])) { reset_state(m, d); d->status |= HMJ_STATUS_BADQPOS; break; }
  for (int i = 0; i < HMJ_NV; i++)
    if (is_bad(d->qvel[i])) { reset_state(m, d); d->status |= HMJ_STATUS_BADQVEL; break; }
  orc_forward(m, p, d);
  for (int i = 0; i < HMJ_NV; i++)
    if (is_bad(d->qacc[i])) {
      reset_state(m, d);
      d->status |= HMJ_STATUS_BADQACC;
      orc_forward(m, p, d);
      break;
    }
  orc_euler(m, p, d);
}

int orc_step_n(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d, int n) {
  for (int i = 0; i < n; i++) orc_step(m, p, d);
  return d->status;
}

/* ------------------------------------------------------------------ TimeSpline::Sample (spline.cc:103-156) */
static double spline_slope(const double* times, const double* values, int n, int dim, int node, int vi) {
  const double* v = values + (size_t)node * dim;
  if (node == 0) {
    const double* nx = values + (size_t)(node + 1) * dim;
    return (nx[vi] - v[vi]) / (times[node + 1] - times[node]);
  }
  const double* pv = values + (size_t)(node - 1) * dim;
  if (node == n - 1) return (v[vi] - pv[vi]) / (times[node] - times[node - 1]);
  const double* nx = values + (size_t)(node + 1) * dim;
  return 0.5 * (nx[vi] - v[vi]) / (times[node + 1] - times[node]) +
         0.5 * (v[vi] - pv[vi]) / (times[node] - times[node - 1]);
}

void orc_spline_sample(const double* times, const double* values, int n, int dim, int interpolation,
                       double time, double* out) {
  if (n == 0) { memset(out, 0, dim * sizeof(double)); return; }
  int upper = 0; /* std::upper_bound: first node with times[upper] > time */
  while (upper < n && !(time < times[upper])) upper++;
  if (upper == n) { memcpy(out, values + (size_t)(n - 1) * dim, dim * sizeof(double)); return; }
  if (upper == 0) { memcpy(out, values, dim * sizeof(double)); return; }
  int lower = upper - 1;
  double t = (time - times[lower]) / (times[upper] - times[lower]);
  const double* lo = values + (size_t)lower * dim;
  const double* up = values + (size_t)upper * dim;
  if (interpolation == 0) {
    memcpy(out, lo, dim * sizeof(double));
  } else if (interpolation == 1) {
    for (int i = 0; i < dim; i++) out[i] = lo[i] * (1 - t) + up[i] * t;
  } else {
    double dt = times[upper] - times[lower];
    double c0 = 2.0 * t * t * t - 3.0 * t * t + 1.0;
    double c1 = (t * t * t - 2.0 * t * t + t) * dt;
    double c2 = -2.0 * t * t * t + 3 * t * t;
    double c3 = (t * t * t - t * t) * dt;
    for (int i = 0; i < dim; i++) {
      double m0 = spline_slope(times, values, n, dim, lower, i);
      double m1 = spline_slope(times, values, n, dim, upper, i);
      out[i] = c0 * lo[i] + c1 * m0 + c2 * up[i] + c3 * m1;
    }
  }
}

/* ------------------------------------------------------------------ Norm (norm.cc:50-210), value only */
double orc_norm(const double* x, const double* params, int n, int type) {
  double y = 0, p = params ? params[0] : 0, q = params ? params[1] : 0;
  switch (type) {
    case -1: y = x[0]; break;
    case 0:
      for (int i = 0; i < n; i++) y += x[i] * x[i];
      y *= 0.5;
      break;
    case 1: {
      double c = 0;
      for (int i = 0; i < n; i++) c += x[i] * x[i];
      double a = pow(c, q / 2) + pow(p, q);
      y = pow(a, 1 / q) - p;
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
      for (int i = 0; i < n; i++) {
        if (p > 0) y += p * log(1 + exp(x[i] / p));
        else y += x[i] > 0 ? x[i] : 0;
      }
      break;
    default: y = NAN;
  }
  return y;
}

/* BaseResidualFn::CostTerms / CostValue (task.cc:71-110) */
void orc_cost_terms(const hmj_model_t* m, const double* residual, double* terms) {
  int shift = 0;
  for (int k = 0; k < m->task_nterm; k++) {
    terms[k] = m->task_weight[k] * orc_norm(residual + shift, m->task_normprm[k], m->task_dim[k], m->task_norm[k]);
    shift += m->task_dim[k];
  }
}

double orc_cost_value(const hmj_model_t* m, const double* residual) {
  double terms[HMJ_NTERM + 1], cost = 0;
  orc_cost_terms(m, residual, terms);
  for (int k = 0; k < m->task_nterm; k++) cost += terms[k];
  if (fabs(m->task_risk) < 1e-6) return cost; /* kRiskNeutralTolerance */
  return (exp(m->task_risk * cost) - 1.0) / m->task_risk;
}

/* humanoid_cap::Stand::ResidualFn::Residual (stand.cc:40-99) */
void orc_residual(const hmj_model_t* m, const orc_data_t* d, double* residual) {
  int counter = 0;
  const double* f1 = d->site_xpos[0];  /* right_bottom_foot */
  const double* f2 = d->site_xpos[1];  /* left_bottom_foot */
  const double* head = d->xpos[2];     /* torso_position: framepos of body torso */
  residual[counter++] = (head[2] - 0.5 * (f1[2] + f2[2])) - m->task_height_goal;
  const double* com = d->subtree_com[2];      /* torso_subtreecom: body torso (no children) */
  const double* comvel = d->subtree_linvel[1];/* torso_subtreelinvel: body humanoid_world_link */
  const double kFallTime = 0.2;
  double cp[2] = {com[0] + comvel[0] * kFallTime, com[1] + comvel[1] * kFallTime};
  double fxy[2] = {0.0, 0.0};
  fxy[0] += f1[0]; fxy[1] += f1[1];
  fxy[0] += f2[0]; fxy[1] += f2[1];
  fxy[0] *= 0.5; fxy[1] *= 0.5;
  fxy[0] -= cp[0]; fxy[1] -= cp[1];
  residual[counter++] = sqrt(fxy[0] * fxy[0] + fxy[1] * fxy[1]);
  residual[counter++] = comvel[0];
  residual[counter++] = comvel[1];
  for (int i = 6; i < HMJ_NV; i++) residual[counter++] = d->qvel[i];
  for (int i = 0; i < HMJ_NU; i++) residual[counter++] = d->ctrl[i];
}

/* ------------------------------------------------------------------ Trajectory::NoisyRollout, xfrc_std = 0 */
int orc_rollout(const hmj_model_t* m, const hmj_params_t* p, const double* state, double time,
                const double* spline_times, const double* spline_values, int nnode, int interpolation,
                int horizon, double* states, double* actions, double* times, double* residual,
                double* costs, double* trace, double* total_return) {
  orc_data_t dd;
  orc_data_t* d = &dd;
  const int ns = HMJ_NSTATE, nu = HMJ_NU, nr = HMJ_NRESIDUAL;
  orc_reset_data(m, d); /* deterministic stand-in for the per-thread mjData reuse (planner.cc:374) */
  memcpy(states, state, ns * sizeof(double));
  memcpy(d->qpos, state, HMJ_NQ * sizeof(double));
  memcpy(d->qvel, state + HMJ_NQ, HMJ_NV * sizeof(double));
  times[0] = time;
  d->time = time;
  for (int t = 0; t < horizon - 1; t++) {
    double* a = actions + (size_t)t * nu;
    orc_spline_sample(spline_times, spline_values, nnode, nu, interpolation, d->time, a);
    for (int i = 0; i < nu; i++) {
      if (a[i] < m->act_ctrlrange[i][0]) a[i] = m->act_ctrlrange[i][0];
      else if (a[i] > m->act_ctrlrange[i][1]) a[i] = m->act_ctrlrange[i][1];
    }
    memcpy(d->ctrl, a, nu * sizeof(double));
    orc_step(m, p, d);
    memcpy(residual + (size_t)t * nr, d->residual, nr * sizeof(double));
    memcpy(trace + (size_t)t * 3, d->xpos[2], 3 * sizeof(double)); /* trace0 = torso framepos */
    if (d->status) { *total_return = 1.0e6; return 1; }               /* kMaxReturnValue */
    memcpy(states + (size_t)(t + 1) * ns, d->qpos, HMJ_NQ * sizeof(double));
    memcpy(states + (size_t)(t + 1) * ns + HMJ_NQ, d->qvel, HMJ_NV * sizeof(double));
    times[t + 1] = d->time;
  }
  if (horizon > 1) memcpy(actions + (size_t)(horizon - 1) * nu, actions + (size_t)(horizon - 2) * nu, nu * sizeof(double));
  else memset(actions, 0, nu * sizeof(double));
  orc_forward(m, p, d);
  memcpy(residual + (size_t)(horizon - 1) * nr, d->residual, nr * sizeof(double));
  memcpy(trace + (size_t)(horizon - 1) * 3, d->xpos[2], 3 * sizeof(double));
  /* UpdateReturn (trajectory.cc:312-326) */
  double tot = 0;
  for (int t = 0; t < horizon; t++) {
    costs[t] = orc_cost_value(m, residual + (size_t)t * nr);
    tot += costs[t];
  }
  *total_return = tot / (horizon > 1 ? horizon : 1);
  return 0;
}

/* Trajectory::NoisyRollout with xfrc_std > 0 (mjpc/trajectory.cc:100-210, perturbation :147-155): orc_rollout with an
 * Ornstein-Uhlenbeck wrench on every body; `noise` [horizon-1][6*nbody] are the unit normals (the reference draws them from an
 * entropy-seeded absl::BitGen).  Returns the failure flag; total_return as UpdateReturn (:312-326). */
int orc_rollout_noisy(const hmj_model_t* m, const hmj_params_t* p, const double* state, double time,
                      const double* spline_times, const double* spline_values, int nnode, int interpolation,
                      int horizon, double xfrc_std, double xfrc_rate, const double* noise, double* total_return) {
  orc_data_t dd;
  orc_data_t* d = &dd;
  const int nu = HMJ_NU;
  orc_reset_data(m, d);
  memcpy(d->qpos, state, HMJ_NQ * sizeof(double));
  memcpy(d->qvel, state + HMJ_NQ, HMJ_NV * sizeof(double));
  d->time = time;
  const double rate = exp(-p->timestep / xfrc_rate);
  const double scale = xfrc_std * sqrt(1 - rate * rate);
  double tot = 0;
  for (int t = 0; t < horizon - 1; t++) {
    double a[HMJ_NU];
    orc_spline_sample(spline_times, spline_values, nnode, nu, interpolation, d->time, a);
    for (int i = 0; i < nu; i++) {
      if (a[i] < m->act_ctrlrange[i][0]) a[i] = m->act_ctrlrange[i][0];
      else if (a[i] > m->act_ctrlrange[i][1]) a[i] = m->act_ctrlrange[i][1];
    }
    memcpy(d->ctrl, a, nu * sizeof(double));
    if (xfrc_std > 0) {
      double* x = &d->xfrc_applied[0][0];
      for (int i = 0; i < 6 * HMJ_NBODY; i++) x[i] = rate * x[i] + noise[(size_t)t * 6 * HMJ_NBODY + i] * scale;
    }
    orc_step(m, p, d);
    tot += orc_cost_value(m, d->residual);
    if (d->status) { *total_return = 1.0e6; return 1; }
  }
  orc_forward(m, p, d);
  tot += orc_cost_value(m, d->residual);
  *total_return = tot / (horizon > 1 ? horizon : 1);
  return 0;
}
