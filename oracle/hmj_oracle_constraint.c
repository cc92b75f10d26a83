/* hmj_oracle_constraint.c — oracle: constraint assembly + Newton solver.
 * TEST INFRASTRUCTURE ONLY (see hmj_oracle.h).
 *
 * Restates MuJoCo 3.1.4 mj_makeConstraint (mujoco.h:358), mj_referenceConstraint (:367),
 * mj_fwdConstraint (:259) for this model: joint-limit rows then pyramidal contact rows
 * (condim 6 -> 10 rows per contact), impedance / reference acceleration, dense Newton
 * (primal) solver with Cholesky factor, incremental rank-one updates and exact 1-D line
 * search.  Options: mjOption defaults, mjmodel.h:403-444 (tolerance 1e-8, 100 iterations,
 * ls_tolerance 0.01, 50 line-search iterations, impratio 1, refsafe on, warmstart on).
 */
#include <math.h>
#include <stdio.h>
#include <string.h>

#include "hmj_oracle_internal.h"

#define NV HMJ_NV

/* ------------------------------------------------------------------ rows */
static int add_row(orc_data_t* d, const double* jac, double pos, double margin, int type, int id) {
  if (d->nefc >= ORC_MAXEFC) { d->status |= HMJ_STATUS_CONTACTFULL; return 0; }
  int r = d->nefc++;
  memcpy(d->efc_J[r], jac, NV * sizeof(double));
  d->efc_pos[r] = pos;
  d->efc_margin[r] = margin;
  d->efc_type[r] = type;
  d->efc_id[r] = id;
  return 1;
}

static void instantiate_limits(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  for (int i = 0; i < HMJ_NJNT; i++) {
    if (!m->jnt_limited[i] || m->jnt_type[i] != HMJ_JNT_HINGE) continue;
    double margin = p->jnt_margin[i];
    double value = d->qpos[m->jnt_qposadr[i]];
    for (int side = -1; side <= 1; side += 2) {
      double dist = side * (p->jnt_range[i][(side + 1) / 2] - value);
      if (dist < margin) {
        double jac[NV];
        memset(jac, 0, sizeof(jac));
        jac[m->jnt_dofadr[i]] = -(double)side;
        add_row(d, jac, dist, margin, ORC_CNSTR_LIMIT_JOINT, i);
      }
    }
  }
}

static void instantiate_contacts(const hmj_model_t* m, orc_data_t* d) {
  for (int i = 0; i < d->ncon; i++) {
    const hmj_contact_t* c = &d->contact[i];
    int b1 = m->geom_bodyid[c->g1], b2 = m->geom_bodyid[c->g2];
    double j1p[3 * NV], j1r[3 * NV], j2p[3 * NV], j2r[3 * NV], difp[3 * NV], difr[3 * NV], jac[6 * NV];
    d->contact_efc_address[i] = d->nefc;
    orc_jac(m, d, j1p, j1r, c->pos, b1);
    orc_jac(m, d, j2p, j2r, c->pos, b2);
    for (int k = 0; k < 3 * NV; k++) { difp[k] = j2p[k] - j1p[k]; difr[k] = j2r[k] - j1r[k]; }
    /* rotate into the contact frame: rows = normal, tangent1, tangent2 (translation), then rotation */
    for (int r = 0; r < 3; r++)
      for (int k = 0; k < NV; k++) {
        jac[r * NV + k] = c->frame[3 * r] * difp[k] + c->frame[3 * r + 1] * difp[NV + k] + c->frame[3 * r + 2] * difp[2 * NV + k];
        jac[(3 + r) * NV + k] = c->frame[3 * r] * difr[k] + c->frame[3 * r + 1] * difr[NV + k] + c->frame[3 * r + 2] * difr[2 * NV + k];
      }
    /* pyramidal cone: one pair of opposing edges per friction dimension */
    for (int k = 1; k < 6; k++) {
      double row[NV];
      for (int q = 0; q < NV; q++) row[q] = jac[q] + jac[k * NV + q] * c->friction[k - 1];
      if (!add_row(d, row, c->dist, 0.0, ORC_CNSTR_CONTACT_PYRAMIDAL, i)) return;
      for (int q = 0; q < NV; q++) row[q] = jac[q] + jac[k * NV + q] * (-c->friction[k - 1]);
      if (!add_row(d, row, c->dist, 0.0, ORC_CNSTR_CONTACT_PYRAMIDAL, i)) return;
    }
    d->flops_constraint += 1100;
  }
}

static void diag_approx(const hmj_model_t* m, orc_data_t* d) {
  for (int i = 0; i < d->nefc; i++) {
    if (d->efc_type[i] == ORC_CNSTR_LIMIT_JOINT) {
      d->efc_diagApprox[i] = m->dof_invweight0[m->jnt_dofadr[d->efc_id[i]]];
    } else {
      const hmj_contact_t* c = &d->contact[d->efc_id[i]];
      int b1 = m->geom_bodyid[c->g1], b2 = m->geom_bodyid[c->g2];
      double tran = m->body_invweight0[b1][0] + m->body_invweight0[b2][0];
      double rot = m->body_invweight0[b1][1] + m->body_invweight0[b2][1];
      for (int j = 0; j < 10; j++) {
        double fri = c->friction[j / 2];
        d->efc_diagApprox[i + j] = tran + fri * fri * (j < 4 ? tran : rot);
      }
      i += 9;
    }
  }
}

static void get_impedance(const double* solimp, double pos, double margin, double* imp, double* impP) {
  if (solimp[0] == solimp[1] || solimp[2] <= MJ_MINVAL) {
    *imp = 0.5 * (solimp[0] + solimp[1]);
    *impP = 0;
    return;
  }
  double x = (pos - margin) / solimp[2], sgn = 1, y, yP;
  if (x < 0) { x = -x; sgn = -1; }
  if (x >= 1 || x <= 0) {
    *imp = (x >= 1 ? solimp[1] : solimp[0]);
    *impP = 0;
    return;
  }
  if (solimp[4] == 1) {
    y = x; yP = 1;
  } else if (x <= solimp[3]) {
    double a = 1 / pow(solimp[3], solimp[4] - 1);
    y = a * pow(x, solimp[4]);
    yP = solimp[4] * a * pow(x, solimp[4] - 1);
  } else {
    double b = 1 / pow(1 - solimp[3], solimp[4] - 1);
    y = 1 - b * pow(1 - x, solimp[4]);
    yP = solimp[4] * b * pow(1 - x, solimp[4] - 1);
  }
  *imp = solimp[0] + y * (solimp[1] - solimp[0]);
  *impP = yP * sgn * (solimp[1] - solimp[0]) / solimp[2];
}

static void make_impedance(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  for (int i = 0; i < d->nefc; i++) {
    double solref[2], solimp[5], imp, impP;
    if (d->efc_type[i] == ORC_CNSTR_LIMIT_JOINT) {
      memcpy(solref, m->jnt_solref[d->efc_id[i]], sizeof(solref));
      memcpy(solimp, m->jnt_solimp[d->efc_id[i]], sizeof(solimp));
    } else {
      memcpy(solref, d->contact[d->efc_id[i]].solref, sizeof(solref));
      memcpy(solimp, d->contact[d->efc_id[i]].solimp, sizeof(solimp));
    }
    /* mj_assignRef / mj_assignImp sanitisation */
    if (solref[0] > 0) solref[0] = fmax(solref[0], 2 * p->timestep); /* refsafe */
    solimp[0] = fmin(MJ_MAXIMP, fmax(MJ_MINIMP, solimp[0]));
    solimp[1] = fmin(MJ_MAXIMP, fmax(MJ_MINIMP, solimp[1]));
    solimp[2] = fmax(0, solimp[2]);
    solimp[3] = fmin(MJ_MAXIMP, fmax(MJ_MINIMP, solimp[3]));
    solimp[4] = fmax(1, solimp[4]);

    get_impedance(solimp, d->efc_pos[i], d->efc_margin[i], &imp, &impP);
    d->efc_R[i] = fmax(MJ_MINVAL, (1 - imp) * d->efc_diagApprox[i] / imp);
    if (solref[0] > 0) {
      d->efc_KBIP[i][0] = 1 / fmax(MJ_MINVAL, solimp[1] * solimp[1] * solref[0] * solref[0] * solref[1] * solref[1]);
      d->efc_KBIP[i][1] = 2 / fmax(MJ_MINVAL, solimp[1] * solref[0]);
    } else {
      d->efc_KBIP[i][0] = -solref[0] / fmax(MJ_MINVAL, solimp[1] * solimp[1]);
      d->efc_KBIP[i][1] = -solref[1] / fmax(MJ_MINVAL, solimp[1]);
    }
    d->efc_KBIP[i][2] = imp;
    d->efc_KBIP[i][3] = impP;
  }
  /* pyramidal contacts: all 2*(dim-1) edges share Rpy = 2 * mu^2 * R(first edge) */
  for (int i = 0; i < d->nefc; i++) {
    if (d->efc_type[i] == ORC_CNSTR_CONTACT_PYRAMIDAL) {
      const hmj_contact_t* c = &d->contact[d->efc_id[i]];
      double Rpy = 2 * c->friction[0] * c->friction[0] * d->efc_R[i];
      for (int j = 0; j < 10; j++) d->efc_R[i + j] = Rpy;
      i += 9;
    }
  }
  for (int i = 0; i < d->nefc; i++) d->efc_D[i] = 1 / d->efc_R[i];
}

void orc_make_constraint(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  d->nefc = 0;
  instantiate_limits(m, p, d);
  instantiate_contacts(m, d);
  diag_approx(m, d);
  make_impedance(m, p, d);
  d->flops_constraint += 30.0 * d->nefc;
}

static void mul_jac_vec(const orc_data_t* d, double* res, const double* vec) {
  for (int i = 0; i < d->nefc; i++) {
    double s = 0;
    for (int k = 0; k < NV; k++) s += d->efc_J[i][k] * vec[k];
    res[i] = s;
  }
}

static void mul_jac_t_vec(const orc_data_t* d, double* res, const double* vec) {
  memset(res, 0, NV * sizeof(double));
  for (int i = 0; i < d->nefc; i++) {
    double f = vec[i];
    if (f != 0)
      for (int k = 0; k < NV; k++) res[k] += d->efc_J[i][k] * f;
  }
}

void orc_reference_constraint(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  (void)m; (void)p;
  mul_jac_vec(d, d->efc_vel, d->qvel);
  for (int i = 0; i < d->nefc; i++)
    d->efc_aref[i] = -d->efc_KBIP[i][1] * d->efc_vel[i] -
                     d->efc_KBIP[i][0] * d->efc_KBIP[i][2] * (d->efc_pos[i] - d->efc_margin[i]);
  d->flops_constraint += 40.0 * d->nefc;
}

/* ------------------------------------------------------------------ mj_constraintUpdate */
static void constraint_update(orc_data_t* d, const double* jar, double* cost) {
  double s = 0;
  for (int i = 0; i < d->nefc; i++) {
    /* limits and pyramidal contacts are both one-sided quadratic rows */
    if (jar[i] >= 0) {
      d->efc_force[i] = 0;
      d->efc_state[i] = ORC_STATE_SATISFIED;
    } else {
      d->efc_force[i] = -d->efc_D[i] * jar[i];
      d->efc_state[i] = ORC_STATE_QUADRATIC;
      s += 0.5 * d->efc_D[i] * jar[i] * jar[i];
    }
  }
  mul_jac_t_vec(d, d->qfrc_constraint, d->efc_force);
  if (cost) *cost = s;
}

/* ------------------------------------------------------------------ dense Cholesky helpers */
static int chol_factor(double* mat, int n, double mindiag) {
  int rank = n;
  for (int j = 0; j < n; j++) {
    double tmp = mat[j * (n + 1)];
    for (int k = 0; k < j; k++) tmp -= mat[j * n + k] * mat[j * n + k];
    if (tmp < mindiag) { tmp = mindiag; rank--; }
    mat[j * (n + 1)] = sqrt(tmp);
    tmp = 1 / mat[j * (n + 1)];
    for (int i = j + 1; i < n; i++) {
      double s = mat[i * n + j];
      for (int k = 0; k < j; k++) s -= mat[i * n + k] * mat[j * n + k];
      mat[i * n + j] = s * tmp;
    }
  }
  return rank;
}

static void chol_solve(double* res, const double* mat, const double* vec, int n) {
  if (res != vec) memcpy(res, vec, n * sizeof(double));
  for (int i = 0; i < n; i++) {
    for (int k = 0; k < i; k++) res[i] -= mat[i * n + k] * res[k];
    res[i] /= mat[i * (n + 1)];
  }
  for (int i = n - 1; i >= 0; i--) {
    for (int j = i + 1; j < n; j++) res[i] -= mat[j * n + i] * res[j];
    res[i] /= mat[i * (n + 1)];
  }
}

static int chol_update(double* mat, double* x, int n, int flg_plus) {
  int rank = n;
  for (int k = 0; k < n; k++) {
    if (x[k] != 0) {
      double Lkk = mat[k * (n + 1)];
      double tmp = Lkk * Lkk + (flg_plus ? x[k] * x[k] : -x[k] * x[k]);
      if (tmp < MJ_MINVAL) { tmp = MJ_MINVAL; rank--; }
      double r = sqrt(tmp), c = r / Lkk, cinv = 1 / c, s = x[k] / Lkk;
      mat[k * (n + 1)] = r;
      if (flg_plus)
        for (int i = k + 1; i < n; i++) mat[i * n + k] = (mat[i * n + k] + s * x[i]) * cinv;
      else
        for (int i = k + 1; i < n; i++) mat[i * n + k] = (mat[i * n + k] - s * x[i]) * cinv;
      for (int i = k + 1; i < n; i++) x[i] = c * x[i] - s * mat[i * n + k];
    }
  }
  return rank;
}

/* ------------------------------------------------------------------ Newton context */
typedef struct {
  double Jaref[ORC_MAXEFC], Jv[ORC_MAXEFC], quad[ORC_MAXEFC][3];
  double Ma[NV], Mv[NV], grad[NV], Mgrad[NV], search[NV];
  double H[NV * NV];
  double quadGauss[3];
  double cost, gauss;
  int LSiter;
} ctx_t;

static void make_hessian(const hmj_model_t* m, orc_data_t* d, ctx_t* c) {
  memset(c->H, 0, sizeof(c->H));
  for (int r = 0; r < d->nefc; r++) {
    if (d->efc_state[r] != ORC_STATE_QUADRATIC) continue;
    const double* J = d->efc_J[r];
    double D = d->efc_D[r];
    for (int i = 0; i < NV; i++) {
      double t = J[i] * D;
      if (t != 0)
        for (int j = 0; j <= i; j++) c->H[i * NV + j] += t * J[j];
    }
  }
  /* add M (lower triangle) */
  for (int i = 0; i < NV; i++) {
    int adr = m->dof_Madr[i];
    for (int j = i; j >= 0; j = m->dof_parentid[j]) c->H[i * NV + j] += d->qM[adr++];
  }
  for (int i = 0; i < NV; i++)
    for (int j = i + 1; j < NV; j++) c->H[i * NV + j] = c->H[j * NV + i];
  chol_factor(c->H, NV, MJ_MINVAL);
  d->flops_solver += 342.0 * d->nefc + 2600;
}

static void hessian_incremental(const hmj_model_t* m, orc_data_t* d, ctx_t* c, const int* oldstate) {
  for (int i = 0; i < d->nefc; i++) {
    int flag = -1;
    if (oldstate[i] != ORC_STATE_QUADRATIC && d->efc_state[i] == ORC_STATE_QUADRATIC) flag = 1;
    else if (oldstate[i] == ORC_STATE_QUADRATIC && d->efc_state[i] != ORC_STATE_QUADRATIC) flag = 0;
    if (flag != -1) {
      double vec[NV], s = sqrt(d->efc_D[i]);
      for (int k = 0; k < NV; k++) vec[k] = d->efc_J[i][k] * s;
      int rank = chol_update(c->H, vec, NV, flag);
      d->flops_solver += 4.0 * NV * NV;
      if (rank < NV) {
        make_hessian(m, d, c);
        return;
      }
    }
  }
}

static void update_constraint(orc_data_t* d, ctx_t* c) {
  constraint_update(d, c->Jaref, &c->cost);
  double gauss = 0;
  for (int i = 0; i < NV; i++) gauss += 0.5 * (c->Ma[i] - d->qfrc_smooth[i]) * (d->qacc[i] - d->qacc_smooth[i]);
  c->gauss = gauss;
  c->cost += gauss;
  d->flops_solver += 40.0 * d->nefc + 100;
}

static void update_gradient(orc_data_t* d, ctx_t* c) {
  for (int i = 0; i < NV; i++) c->grad[i] = c->Ma[i] - d->qfrc_smooth[i] - d->qfrc_constraint[i];
  chol_solve(c->Mgrad, c->H, c->grad, NV);
  d->flops_solver += 2.0 * NV * NV;
}

/* ------------------------------------------------------------------ exact line search */
typedef struct { double alpha, cost, deriv[2]; } pnt_t;

static void ls_prepare(orc_data_t* d, ctx_t* c) {
  double a = 0, b = 0, cc = 0;
  for (int i = 0; i < NV; i++) {
    a += c->search[i] * c->Ma[i];
    b += d->qfrc_smooth[i] * c->search[i];
    cc += c->search[i] * c->Mv[i];
  }
  c->quadGauss[0] = c->gauss;
  c->quadGauss[1] = a - b;
  c->quadGauss[2] = 0.5 * cc;
  for (int i = 0; i < d->nefc; i++) {
    double DJ0 = d->efc_D[i] * c->Jaref[i];
    c->quad[i][0] = 0.5 * (c->Jaref[i] * DJ0);
    c->quad[i][1] = c->Jv[i] * DJ0;
    c->quad[i][2] = 0.5 * (c->Jv[i] * d->efc_D[i] * c->Jv[i]);
  }
}

static void ls_eval(orc_data_t* d, ctx_t* c, pnt_t* p) {
  double alpha = p->alpha, q0 = c->quadGauss[0], q1 = c->quadGauss[1], q2 = c->quadGauss[2];
  for (int i = 0; i < d->nefc; i++) {
    double x = c->Jaref[i] + alpha * c->Jv[i];
    if (x < 0) { q0 += c->quad[i][0]; q1 += c->quad[i][1]; q2 += c->quad[i][2]; }
  }
  p->cost = alpha * alpha * q2 + alpha * q1 + q0;
  p->deriv[0] = 2 * alpha * q2 + q1;
  p->deriv[1] = 2 * q2;
  if (p->deriv[1] <= 0) p->deriv[1] = MJ_MINVAL;
  c->LSiter++;
  d->flops_solver += 8.0 * d->nefc;
}

static int update_bracket(orc_data_t* d, ctx_t* c, pnt_t* p, const pnt_t cand[3], pnt_t* pnext) {
  int flag = 0;
  for (int i = 0; i < 3; i++) {
    if (p->deriv[0] < 0 && cand[i].deriv[0] < 0 && p->deriv[0] < cand[i].deriv[0]) { *p = cand[i]; flag = 1; }
    else if (p->deriv[0] > 0 && cand[i].deriv[0] > 0 && p->deriv[0] > cand[i].deriv[0]) { *p = cand[i]; flag = 2; }
  }
  if (flag) {
    pnext->alpha = p->alpha - p->deriv[0] / p->deriv[1];
    ls_eval(d, c, pnext);
  }
  return flag;
}

static double line_search(const hmj_model_t* m, orc_data_t* d, ctx_t* c) {
  pnt_t p0, p1, p2, pmid, p1next, p2next;
  c->LSiter = 0;
  double snorm = 0;
  for (int i = 0; i < NV; i++) snorm += c->search[i] * c->search[i];
  snorm = sqrt(snorm);
  double scale = 1 / (m->meaninertia * (NV > 1 ? NV : 1));
  if (snorm < MJ_MINVAL) return 0;
  double gtol = m->tolerance * m->ls_tolerance * snorm / scale;

  orc_mul_m(m, d->qM, c->Mv, c->search);
  mul_jac_vec(d, c->Jv, c->search);
  d->flops_solver += 36.0 * d->nefc + 250;
  ls_prepare(d, c);

  p0.alpha = 0;
  ls_eval(d, c, &p0);
  p1.alpha = p0.alpha - p0.deriv[0] / p0.deriv[1];
  ls_eval(d, c, &p1);
  if (p0.cost < p1.cost) p1 = p0;
  if (fabs(p1.deriv[0]) < gtol) return p1.alpha;

  int dir = (p1.deriv[0] < 0 ? +1 : -1);
  int p2update = 0;
  p2 = p1;
  while (p1.deriv[0] * dir <= -gtol && c->LSiter < m->ls_iterations) {
    p2 = p1;
    p2update = 1;
    p1.alpha -= p1.deriv[0] / p1.deriv[1];
    ls_eval(d, c, &p1);
    if (fabs(p1.deriv[0]) < gtol) return p1.alpha;
  }
  if (c->LSiter >= m->ls_iterations) return p1.alpha;
  if (!p2update) return p1.alpha;

  p2next = p1;
  p1next.alpha = p1.alpha - p1.deriv[0] / p1.deriv[1];
  ls_eval(d, c, &p1next);

  while (c->LSiter < m->ls_iterations) {
    pmid.alpha = 0.5 * (p1.alpha + p2.alpha);
    ls_eval(d, c, &pmid);
    pnt_t cand[3] = {p1next, p2next, pmid};
    double bestcost = 0;
    int bestind = -1;
    for (int i = 0; i < 3; i++)
      if (fabs(cand[i].deriv[0]) < gtol && (bestind == -1 || cand[i].cost < bestcost)) {
        bestcost = cand[i].cost;
        bestind = i;
      }
    if (bestind >= 0) return cand[bestind].alpha;
    int b1 = update_bracket(d, c, &p1, cand, &p1next);
    int b2 = update_bracket(d, c, &p2, cand, &p2next);
    if (!b1 && !b2) return pmid.alpha;
  }
  if (p1.cost <= p2.cost && p1.cost < p0.cost) return p1.alpha;
  if (p2.cost <= p1.cost && p2.cost < p0.cost) return p2.alpha;
  return 0;
}

/* ------------------------------------------------------------------ mj_fwdConstraint */
void orc_fwd_constraint(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  (void)p;
  d->solver_niter = 0;
  d->solver_nls = 0;
  if (d->nefc == 0) {
    memcpy(d->qacc, d->qacc_smooth, sizeof(d->qacc));
    memcpy(d->qacc_warmstart, d->qacc_smooth, sizeof(d->qacc));
    memset(d->qfrc_constraint, 0, sizeof(d->qfrc_constraint));
    return;
  }
  ctx_t c;
  int nefc = d->nefc;
#ifdef ORC_TRACE
  for (int i = 0; i < nefc; i++) fprintf(stderr, "ORC row %d D %.17g aref %.17g vel %.17g pos %.17g\n", i, d->efc_D[i], d->efc_aref[i], d->efc_vel[i], d->efc_pos[i]);
  for (int i = 0; i < d->ncon; i++) fprintf(stderr, "ORC con %d g %d %d dist %.17g fric %g %g %g\n", i, d->contact[i].geom1, d->contact[i].geom2, d->contact[i].dist, d->contact[i].friction[0], d->contact[i].friction[2], d->contact[i].friction[3]);
#endif
  double scale = 1 / (m->meaninertia * (NV > 1 ? NV : 1));

  /* efc_b = J*qacc_smooth - aref */
  mul_jac_vec(d, d->efc_b, d->qacc_smooth);
  for (int i = 0; i < nefc; i++) d->efc_b[i] -= d->efc_aref[i];

  /* warmstart: best of (qacc_warmstart, qacc_smooth) */
  {
    double jar[ORC_MAXEFC], Ma[NV], cost_ws, cost_sm;
    memcpy(d->qacc, d->qacc_warmstart, sizeof(d->qacc));
    mul_jac_vec(d, jar, d->qacc_warmstart);
    for (int i = 0; i < nefc; i++) jar[i] -= d->efc_aref[i];
    constraint_update(d, jar, &cost_ws);
    orc_mul_m(m, d->qM, Ma, d->qacc_warmstart);
    for (int i = 0; i < NV; i++) cost_ws += 0.5 * (Ma[i] - d->qfrc_smooth[i]) * (d->qacc_warmstart[i] - d->qacc_smooth[i]);
    constraint_update(d, d->efc_b, &cost_sm);
    if (cost_ws > cost_sm) memcpy(d->qacc, d->qacc_smooth, sizeof(d->qacc));
    d->flops_solver += 160.0 * nefc + 400;
  }

  /* Newton */
  orc_mul_m(m, d->qM, c.Ma, d->qacc);
  mul_jac_vec(d, c.Jaref, d->qacc);
  for (int i = 0; i < nefc; i++) c.Jaref[i] -= d->efc_aref[i];
  update_constraint(d, &c);
  make_hessian(m, d, &c);
  update_gradient(d, &c);
  for (int i = 0; i < NV; i++) c.search[i] = -c.Mgrad[i];

  int iter = 0;
  int oldstate[ORC_MAXEFC];
  while (iter < m->iterations) {
    double alpha = line_search(m, d, &c);
    d->solver_nls += c.LSiter;
#ifdef ORC_TRACE
    fprintf(stderr, "ORC it %d alpha %.17g lsiter %d cost %.17g\n", iter, alpha, c.LSiter, c.cost);
#endif
    if (alpha == 0) break;
    for (int i = 0; i < NV; i++) { d->qacc[i] += alpha * c.search[i]; c.Ma[i] += alpha * c.Mv[i]; }
    for (int i = 0; i < nefc; i++) c.Jaref[i] += alpha * c.Jv[i];
    double oldcost = c.cost;
    memcpy(oldstate, d->efc_state, nefc * sizeof(int));
    update_constraint(d, &c);
    hessian_incremental(m, d, &c, oldstate);
    update_gradient(d, &c);
    double improvement = scale * (oldcost - c.cost);
    double gn = 0;
    for (int i = 0; i < NV; i++) gn += c.grad[i] * c.grad[i];
    double gradient = scale * sqrt(gn);
    iter++;
#ifdef ORC_TRACE
    fprintf(stderr, "ORC    -> cost %.17g improvement %.6g gradient %.6g\n", c.cost, improvement, gradient);
#endif
    if (improvement < m->tolerance || gradient < m->tolerance) break;
    for (int i = 0; i < NV; i++) c.search[i] = -c.Mgrad[i];
  }
  d->solver_niter = iter;
  memcpy(d->qacc_warmstart, d->qacc, sizeof(d->qacc));
}
