/* hmj_oracle_smooth.c — oracle: kinematics, CoM, CRBA, LDL, velocity, RNE, passive, actuation,
 * smooth acceleration, Euler integration.  TEST INFRASTRUCTURE ONLY (see hmj_oracle.h).
 *
 * Restates the MuJoCo 3.1.4 stages declared in the vendored headers
 * (simulation/mujoco/include/mujoco/mujoco.h): mj_kinematics :310, mj_comPos :313, mj_crb :328,
 * mj_factorM :331, mj_solveM :334, mj_comVel :340, mj_passive :343, mj_subtreeVel :346,
 * mj_rne :349, mj_fwdActuation :253, mj_fwdAcceleration :256, mj_Euler :262.
 * Spatial-vector layout is MuJoCo's: [rotational(3), translational(3)].
 */
#include <math.h>
#include <string.h>

#include "hmj_oracle_internal.h"
#include "hmj_sincos.h"

/* ------------------------------------------------------------------ small math */
double orc_normalize3(double* v) {
  double n = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  if (n < MJ_MINVAL) {
    v[0] = 1; v[1] = 0; v[2] = 0;
  } else {
    double s = 1.0 / n;
    v[0] *= s; v[1] *= s; v[2] *= s;
  }
  return n;
}

double orc_normalize4(double* q) {
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < MJ_MINVAL) {
    q[0] = 1; q[1] = 0; q[2] = 0; q[3] = 0;
  } else if (fabs(n - 1) > MJ_MINVAL) {
    double s = 1.0 / n;
    q[0] *= s; q[1] *= s; q[2] *= s; q[3] *= s;
  }
  return n;
}

void orc_mul_quat(double* res, const double* a, const double* b) {
  double t[4];
  t[0] = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  t[1] = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  t[2] = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  t[3] = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  res[0] = t[0]; res[1] = t[1]; res[2] = t[2]; res[3] = t[3];
}

void orc_quat2mat(double* r, const double* q) {
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

void orc_rot_vec_quat(double* res, const double* v, const double* q) {
  /* res = v + 2*cross(q_xyz, q_w*v + cross(q_xyz, v)) */
  double t[3];
  t[0] = q[0] * v[0] + q[2] * v[2] - q[3] * v[1];
  t[1] = q[0] * v[1] + q[3] * v[0] - q[1] * v[2];
  t[2] = q[0] * v[2] + q[1] * v[1] - q[2] * v[0];
  double r0 = v[0] + 2 * (q[2] * t[2] - q[3] * t[1]);
  double r1 = v[1] + 2 * (q[3] * t[0] - q[1] * t[2]);
  double r2 = v[2] + 2 * (q[1] * t[1] - q[2] * t[0]);
  res[0] = r0; res[1] = r1; res[2] = r2;
}

void orc_rot_vec_mat(double* res, const double* v, const double* m) {
  double r0 = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
  double r1 = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
  double r2 = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  res[0] = r0; res[1] = r1; res[2] = r2;
}

void orc_rot_vec_matT(double* res, const double* v, const double* m) {
  double r0 = m[0] * v[0] + m[3] * v[1] + m[6] * v[2];
  double r1 = m[1] * v[0] + m[4] * v[1] + m[7] * v[2];
  double r2 = m[2] * v[0] + m[5] * v[1] + m[8] * v[2];
  res[0] = r0; res[1] = r1; res[2] = r2;
}

void orc_cross(double* r, const double* a, const double* b) {
  double r0 = a[1] * b[2] - a[2] * b[1];
  double r1 = a[2] * b[0] - a[0] * b[2];
  double r2 = a[0] * b[1] - a[1] * b[0];
  r[0] = r0; r[1] = r1; r[2] = r2;
}

static void axis_angle2quat(double* q, const double* axis, double angle) {
  if (angle == 0) {
    q[0] = 1; q[1] = 0; q[2] = 0; q[3] = 0;
  } else {
    double s, c;
    hmj_sincos(angle * 0.5, &s, &c); /* shared deterministic sin/cos (include/hmj_sincos.h): same bits as the kernels */
    q[0] = c;
    q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
  }
}

/* mju_mulInertVec: res = cinert * v  (cinert = [Ixx Iyy Izz Ixy Ixz Iyz mx my mz m]) */
static void mul_inert_vec(double* res, const double* i, const double* v) {
  res[0] = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
  res[1] = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
  res[2] = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
  res[3] = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
  res[4] = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
  res[5] = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
}

static void cross_motion(double* res, const double* vel, const double* v) {
  res[0] = -vel[2] * v[1] + vel[1] * v[2];
  res[1] = vel[2] * v[0] - vel[0] * v[2];
  res[2] = -vel[1] * v[0] + vel[0] * v[1];
  res[3] = -vel[2] * v[4] + vel[1] * v[5];
  res[4] = vel[2] * v[3] - vel[0] * v[5];
  res[5] = -vel[1] * v[3] + vel[0] * v[4];
  res[3] += -vel[5] * v[1] + vel[4] * v[2];
  res[4] += vel[5] * v[0] - vel[3] * v[2];
  res[5] += -vel[4] * v[0] + vel[3] * v[1];
}

static void cross_force(double* res, const double* vel, const double* f) {
  res[0] = -vel[2] * f[1] + vel[1] * f[2];
  res[1] = vel[2] * f[0] - vel[0] * f[2];
  res[2] = -vel[1] * f[0] + vel[0] * f[1];
  res[3] = -vel[2] * f[4] + vel[1] * f[5];
  res[4] = vel[2] * f[3] - vel[0] * f[5];
  res[5] = -vel[1] * f[3] + vel[0] * f[4];
  res[0] += -vel[5] * f[4] + vel[4] * f[5];
  res[1] += vel[5] * f[3] - vel[3] * f[5];
  res[2] += -vel[4] * f[3] + vel[3] * f[4];
}

static double dot6(const double* a, const double* b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3] + a[4] * b[4] + a[5] * b[5];
}

/* ------------------------------------------------------------------ defaults / reset */
void orc_default_params(const hmj_model_t* m, hmj_params_t* p) {
  memset(p, 0, sizeof(*p));
  p->timestep = m->timestep;
  memcpy(p->body_mass, m->body_mass, sizeof(p->body_mass));
  memcpy(p->dof_armature, m->dof_armature, sizeof(p->dof_armature));
  memcpy(p->jnt_range, m->jnt_range, sizeof(p->jnt_range));
  memcpy(p->jnt_margin, m->jnt_margin, sizeof(p->jnt_margin));
  memcpy(p->jnt_stiffness, m->jnt_stiffness, sizeof(p->jnt_stiffness));
  memcpy(p->act_gainprm0, m->act_gainprm0, sizeof(p->act_gainprm0));
  memcpy(p->act_forcerange, m->act_forcerange, sizeof(p->act_forcerange));
  memcpy(p->floor_friction, m->geom_friction[0], sizeof(p->floor_friction));
  memcpy(p->hfield_data, m->hfield_data, sizeof(p->hfield_data));
}

void orc_reset_data(const hmj_model_t* m, orc_data_t* d) {
  memset(d, 0, sizeof(*d));
  memcpy(d->qpos, m->qpos0, sizeof(d->qpos));
}

int orc_sizeof_data(void) { return (int)sizeof(orc_data_t); }
int orc_sizeof_model(void) { return (int)sizeof(hmj_model_t); }
int orc_sizeof_params(void) { return (int)sizeof(hmj_params_t); }

/* ------------------------------------------------------------------ mj_kinematics */
static void local2global(const orc_data_t* d, double* xpos, double* xmat, const double* pos,
                         const double* quat, int body, int sameframe) {
  if (sameframe) {
    memcpy(xpos, d->xpos[body], 3 * sizeof(double));
    memcpy(xmat, d->xmat[body], 9 * sizeof(double));
  } else {
    double t[4];
    orc_rot_vec_mat(xpos, pos, d->xmat[body]);
    xpos[0] += d->xpos[body][0]; xpos[1] += d->xpos[body][1]; xpos[2] += d->xpos[body][2];
    orc_mul_quat(t, d->xquat[body], quat);
    orc_quat2mat(xmat, t);
  }
}

void orc_kinematics(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  (void)p;
  /* world */
  memset(d->xpos[0], 0, sizeof(d->xpos[0]));
  memset(d->xipos[0], 0, sizeof(d->xipos[0]));
  d->xquat[0][0] = 1; d->xquat[0][1] = d->xquat[0][2] = d->xquat[0][3] = 0;
  memset(d->xmat[0], 0, sizeof(d->xmat[0]));
  d->xmat[0][0] = d->xmat[0][4] = d->xmat[0][8] = 1;
  memcpy(d->ximat[0], d->xmat[0], sizeof(d->xmat[0]));

  for (int i = 1; i < HMJ_NBODY; i++) {
    double xpos[3], xquat[4];
    int jntadr = m->body_jntadr[i], jntnum = m->body_jntnum[i], pid = m->body_parentid[i];
    if (jntnum == 1 && m->jnt_type[jntadr] == HMJ_JNT_FREE) {
      int qadr = m->jnt_qposadr[jntadr];
      memcpy(xpos, d->qpos + qadr, 3 * sizeof(double));
      memcpy(xquat, d->qpos + qadr + 3, 4 * sizeof(double));
      orc_normalize4(xquat); /* normalises a copy; qpos itself is left alone */
      memcpy(d->xanchor[jntadr], xpos, sizeof(xpos));
      memcpy(d->xaxis[jntadr], m->jnt_axis[jntadr], sizeof(xpos));
    } else {
      if (pid) {
        orc_rot_vec_mat(xpos, m->body_pos[i], d->xmat[pid]);
        xpos[0] += d->xpos[pid][0]; xpos[1] += d->xpos[pid][1]; xpos[2] += d->xpos[pid][2];
        orc_mul_quat(xquat, d->xquat[pid], m->body_quat[i]);
      } else {
        memcpy(xpos, m->body_pos[i], sizeof(xpos));
        memcpy(xquat, m->body_quat[i], sizeof(xquat));
      }
      for (int j = 0; j < jntnum; j++) {
        int jid = jntadr + j, qadr = m->jnt_qposadr[jid];
        double qloc[4], vec[3];
        orc_rot_vec_quat(d->xaxis[jid], m->jnt_axis[jid], xquat);
        orc_rot_vec_quat(d->xanchor[jid], m->jnt_pos[jid], xquat);
        d->xanchor[jid][0] += xpos[0]; d->xanchor[jid][1] += xpos[1]; d->xanchor[jid][2] += xpos[2];
        /* hinge */
        axis_angle2quat(qloc, m->jnt_axis[jid], d->qpos[qadr] - m->qpos0[qadr]);
        orc_mul_quat(xquat, xquat, qloc);
        orc_rot_vec_quat(vec, m->jnt_pos[jid], xquat);
        xpos[0] = d->xanchor[jid][0] - vec[0];
        xpos[1] = d->xanchor[jid][1] - vec[1];
        xpos[2] = d->xanchor[jid][2] - vec[2];
      }
    }
    orc_normalize4(xquat);
    memcpy(d->xquat[i], xquat, sizeof(xquat));
    memcpy(d->xpos[i], xpos, sizeof(xpos));
    orc_quat2mat(d->xmat[i], xquat);
  }
  for (int i = 1; i < HMJ_NBODY; i++)
    local2global(d, d->xipos[i], d->ximat[i], m->body_ipos[i], m->body_iquat[i], i, m->body_sameframe[i]);
  for (int g = 0; g < HMJ_NGEOM; g++)
    local2global(d, d->geom_xpos[g], d->geom_xmat[g], m->geom_pos[g], m->geom_quat[g], m->geom_bodyid[g], 0);
  for (int s = 0; s < HMJ_NSITE; s++) {
    int b = m->site_bodyid[s];
    orc_rot_vec_mat(d->site_xpos[s], m->site_pos[s], d->xmat[b]);
    d->site_xpos[s][0] += d->xpos[b][0]; d->site_xpos[s][1] += d->xpos[b][1]; d->site_xpos[s][2] += d->xpos[b][2];
  }
  d->flops_smooth += 2300;
}

/* ------------------------------------------------------------------ mj_comPos */
static void inert_com(double* res, const double* inert, const double* mat, const double* dif, double mass) {
  double tmp[9];
  tmp[0] = mat[0] * inert[0]; tmp[1] = mat[3] * inert[0]; tmp[2] = mat[6] * inert[0];
  tmp[3] = mat[1] * inert[1]; tmp[4] = mat[4] * inert[1]; tmp[5] = mat[7] * inert[1];
  tmp[6] = mat[2] * inert[2]; tmp[7] = mat[5] * inert[2]; tmp[8] = mat[8] * inert[2];
  res[0] = mat[0] * tmp[0] + mat[1] * tmp[3] + mat[2] * tmp[6];
  res[1] = mat[3] * tmp[1] + mat[4] * tmp[4] + mat[5] * tmp[7];
  res[2] = mat[6] * tmp[2] + mat[7] * tmp[5] + mat[8] * tmp[8];
  res[3] = mat[0] * tmp[1] + mat[1] * tmp[4] + mat[2] * tmp[7];
  res[4] = mat[0] * tmp[2] + mat[1] * tmp[5] + mat[2] * tmp[8];
  res[5] = mat[3] * tmp[2] + mat[4] * tmp[5] + mat[5] * tmp[8];
  res[0] += mass * (dif[1] * dif[1] + dif[2] * dif[2]);
  res[1] += mass * (dif[0] * dif[0] + dif[2] * dif[2]);
  res[2] += mass * (dif[0] * dif[0] + dif[1] * dif[1]);
  res[3] -= mass * dif[0] * dif[1];
  res[4] -= mass * dif[0] * dif[2];
  res[5] -= mass * dif[1] * dif[2];
  res[6] = mass * dif[0]; res[7] = mass * dif[1]; res[8] = mass * dif[2];
  res[9] = mass;
}

static void dof_com(double* res, const double* axis, const double* offset) {
  res[0] = axis[0]; res[1] = axis[1]; res[2] = axis[2];
  orc_cross(res + 3, axis, offset);
}

void orc_com_pos(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  memset(d->subtree_com, 0, sizeof(d->subtree_com));
  for (int i = HMJ_NBODY - 1; i >= 0; i--) {
    double mass = p->body_mass[i];
    d->subtree_com[i][0] += d->xipos[i][0] * mass;
    d->subtree_com[i][1] += d->xipos[i][1] * mass;
    d->subtree_com[i][2] += d->xipos[i][2] * mass;
    if (i) {
      int pid = m->body_parentid[i];
      d->subtree_com[pid][0] += d->subtree_com[i][0];
      d->subtree_com[pid][1] += d->subtree_com[i][1];
      d->subtree_com[pid][2] += d->subtree_com[i][2];
    }
    /* body_subtreemass is the COMPILE-TIME constant: the reference edits body_mass
     * without mj_setConst (simulation/cpu_env.py:253-265) */
    if (m->body_subtreemass[i] < MJ_MINVAL) {
      memcpy(d->subtree_com[i], d->xipos[i], 3 * sizeof(double));
    } else {
      double s = 1.0 / fmax(MJ_MINVAL, m->body_subtreemass[i]);
      d->subtree_com[i][0] *= s; d->subtree_com[i][1] *= s; d->subtree_com[i][2] *= s;
    }
  }
  memset(d->cinert[0], 0, sizeof(d->cinert[0]));
  for (int i = 1; i < HMJ_NBODY; i++) {
    double off[3];
    const double* com = d->subtree_com[m->body_rootid[i]];
    off[0] = d->xipos[i][0] - com[0]; off[1] = d->xipos[i][1] - com[1]; off[2] = d->xipos[i][2] - com[2];
    inert_com(d->cinert[i], m->body_inertia[i], d->ximat[i], off, p->body_mass[i]);
  }
  for (int j = 0; j < HMJ_NJNT; j++) {
    int da = m->jnt_dofadr[j], bi = m->jnt_bodyid[j];
    double off[3];
    const double* com = d->subtree_com[m->body_rootid[bi]];
    off[0] = com[0] - d->xanchor[j][0]; off[1] = com[1] - d->xanchor[j][1]; off[2] = com[2] - d->xanchor[j][2];
    if (m->jnt_type[j] == HMJ_JNT_FREE) {
      memset(d->cdof[da], 0, 18 * sizeof(double));
      d->cdof[da][3] = 1; d->cdof[da + 1][4] = 1; d->cdof[da + 2][5] = 1;
      for (int i = 0; i < 3; i++) {
        double ax[3] = {d->xmat[bi][i], d->xmat[bi][i + 3], d->xmat[bi][i + 6]};
        dof_com(d->cdof[da + 3 + i], ax, off);
      }
    } else {
      dof_com(d->cdof[da], d->xaxis[j], off);
    }
  }
  d->flops_smooth += 2300;
}

/* ------------------------------------------------------------------ mj_crb / mj_factorM */
void orc_crb(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  memcpy(d->crb, d->cinert, sizeof(d->crb));
  for (int i = HMJ_NBODY - 1; i > 0; i--) {
    int pid = m->body_parentid[i];
    if (pid > 0)
      for (int k = 0; k < 10; k++) d->crb[pid][k] += d->crb[i][k];
  }
  memset(d->qM, 0, sizeof(d->qM));
  for (int i = 0; i < HMJ_NV; i++) {
    double buf[6];
    int adr = m->dof_Madr[i];
    d->qM[adr] = p->dof_armature[i];
    mul_inert_vec(buf, d->crb[m->dof_bodyid[i]], d->cdof[i]);
    for (int j = i; j >= 0; j = m->dof_parentid[j]) d->qM[adr++] += dot6(d->cdof[j], buf);
  }
  d->flops_smooth += 2500;
}

/* LDL factorisation of a sparse inertia-like matrix in place (mj_factorI) */
void orc_factor_i(const hmj_model_t* m, double* qLD, double* diaginv, int* status) {
  for (int k = HMJ_NV - 1; k >= 0; k--) {
    int Madr_kk = m->dof_Madr[k];
    if (qLD[Madr_kk] < MJ_MINVAL) {
      if (status) *status |= HMJ_STATUS_INERTIA;
      qLD[Madr_kk] = MJ_MINVAL;
    }
    int Madr_ki = Madr_kk + 1;
    int i = m->dof_parentid[k];
    while (i >= 0) {
      double tmp = qLD[Madr_ki] / qLD[Madr_kk];
      int cnt = m->dof_Madr[i + 1] - m->dof_Madr[i];
      double* dst = qLD + m->dof_Madr[i];
      const double* src = qLD + Madr_ki;
      for (int c = 0; c < cnt; c++) dst[c] += -tmp * src[c];
      qLD[Madr_ki] = tmp;
      i = m->dof_parentid[i];
      Madr_ki++;
    }
  }
  for (int i = 0; i < HMJ_NV; i++) diaginv[i] = 1.0 / qLD[m->dof_Madr[i]];
}

void orc_solve_ld(const hmj_model_t* m, double* x, const double* qLD, const double* diaginv) {
  for (int i = HMJ_NV - 1; i >= 0; i--) {
    double tmp = x[i];
    if (tmp != 0) {
      int adr = m->dof_Madr[i] + 1;
      int j = m->dof_parentid[i];
      while (j >= 0) {
        x[j] -= qLD[adr++] * tmp;
        j = m->dof_parentid[j];
      }
    }
  }
  for (int i = 0; i < HMJ_NV; i++) x[i] *= diaginv[i];
  for (int i = 0; i < HMJ_NV; i++) {
    int adr = m->dof_Madr[i] + 1;
    int j = m->dof_parentid[i];
    while (j >= 0) {
      x[i] -= qLD[adr++] * x[j];
      j = m->dof_parentid[j];
    }
  }
}

void orc_mul_m(const hmj_model_t* m, const double* qM, double* res, const double* vec) {
  for (int i = 0; i < HMJ_NV; i++) {
    int adr = m->dof_Madr[i];
    res[i] = qM[adr] * vec[i];
    int j = m->dof_parentid[i];
    while (j >= 0) {
      adr++;
      res[i] += qM[adr] * vec[j];
      res[j] += qM[adr] * vec[i];
      j = m->dof_parentid[j];
    }
  }
}

void orc_factor_m(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  (void)p;
  memcpy(d->qLD, d->qM, sizeof(d->qM));
  orc_factor_i(m, d->qLD, d->qLDiagInv, &d->status);
  d->flops_smooth += 1500;
}

/* ------------------------------------------------------------------ mj_comVel */
static void com_vel(const hmj_model_t* m, orc_data_t* d) {
  memset(d->cvel[0], 0, sizeof(d->cvel[0]));
  for (int i = 1; i < HMJ_NBODY; i++) {
    double cvel[6], tmp[6];
    int bda = m->body_dofadr[i], dn = m->body_dofnum[i];
    memcpy(cvel, d->cvel[m->body_parentid[i]], sizeof(cvel));
    if (dn == 6) { /* free joint */
      memset(d->cdof_dot[bda], 0, 18 * sizeof(double));
      for (int k = 0; k < 6; k++)
        tmp[k] = d->cdof[bda][k] * d->qvel[bda] + d->cdof[bda + 1][k] * d->qvel[bda + 1] +
                 d->cdof[bda + 2][k] * d->qvel[bda + 2];
      for (int k = 0; k < 6; k++) cvel[k] += tmp[k];
      for (int k = 0; k < 3; k++) cross_motion(d->cdof_dot[bda + 3 + k], cvel, d->cdof[bda + 3 + k]);
      for (int k = 0; k < 6; k++)
        tmp[k] = d->cdof[bda + 3][k] * d->qvel[bda + 3] + d->cdof[bda + 4][k] * d->qvel[bda + 4] +
                 d->cdof[bda + 5][k] * d->qvel[bda + 5];
      for (int k = 0; k < 6; k++) cvel[k] += tmp[k];
    } else {
      for (int j = 0; j < dn; j++) {
        cross_motion(d->cdof_dot[bda + j], cvel, d->cdof[bda + j]);
        for (int k = 0; k < 6; k++) cvel[k] += d->cdof[bda + j][k] * d->qvel[bda + j];
      }
    }
    memcpy(d->cvel[i], cvel, sizeof(cvel));
  }
}

/* ------------------------------------------------------------------ mj_passive */
static void passive(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  memset(d->qfrc_spring, 0, sizeof(d->qfrc_spring));
  for (int j = 0; j < HMJ_NJNT; j++) {
    double k = p->jnt_stiffness[j];
    int padr = m->jnt_qposadr[j], dadr = m->jnt_dofadr[j];
    if (k == 0) continue;
    if (m->jnt_type[j] == HMJ_JNT_FREE) {
      for (int i = 0; i < 3; i++) d->qfrc_spring[dadr + i] = -k * (d->qpos[padr + i] - m->qpos_spring[padr + i]);
      /* rotational part: quaternion difference as a 3D velocity (mju_subQuat) */
      double quat[4], qneg[4], qdif[4], axis[3];
      memcpy(quat, d->qpos + padr + 3, sizeof(quat));
      orc_normalize4(quat);
      qneg[0] = m->qpos_spring[padr + 3];
      qneg[1] = -m->qpos_spring[padr + 4]; qneg[2] = -m->qpos_spring[padr + 5]; qneg[3] = -m->qpos_spring[padr + 6];
      orc_mul_quat(qdif, qneg, quat);
      axis[0] = qdif[1]; axis[1] = qdif[2]; axis[2] = qdif[3];
      double sin_a_2 = orc_normalize3(axis);
      double speed = 2 * atan2(sin_a_2, qdif[0]);
      if (speed > MJ_PI) speed -= 2 * MJ_PI;
      for (int i = 0; i < 3; i++) d->qfrc_spring[dadr + 3 + i] = -k * (axis[i] * speed);
    } else {
      d->qfrc_spring[dadr] = -k * (d->qpos[padr] - m->qpos_spring[padr]);
    }
  }
  for (int i = 0; i < HMJ_NV; i++) {
    d->qfrc_damper[i] = -m->dof_damping[i] * d->qvel[i];
    d->qfrc_passive[i] = d->qfrc_spring[i] + d->qfrc_damper[i];
  }
}

/* ------------------------------------------------------------------ mj_rne(flg_acc = 0) */
static void rne(const hmj_model_t* m, orc_data_t* d) {
  double cacc[HMJ_NBODY][6], cfrc[HMJ_NBODY][6];
  cacc[0][0] = cacc[0][1] = cacc[0][2] = 0;
  cacc[0][3] = -m->gravity[0]; cacc[0][4] = -m->gravity[1]; cacc[0][5] = -m->gravity[2];
  memset(cfrc[0], 0, sizeof(cfrc[0]));
  for (int i = 1; i < HMJ_NBODY; i++) {
    double tmp[6], tmp1[6];
    int bda = m->body_dofadr[i], dn = m->body_dofnum[i];
    for (int k = 0; k < 6; k++) {
      double s = 0;
      for (int j = 0; j < dn; j++) s += d->cdof_dot[bda + j][k] * d->qvel[bda + j];
      cacc[i][k] = cacc[m->body_parentid[i]][k] + s;
    }
    mul_inert_vec(cfrc[i], d->cinert[i], cacc[i]);
    mul_inert_vec(tmp, d->cinert[i], d->cvel[i]);
    cross_force(tmp1, d->cvel[i], tmp);
    for (int k = 0; k < 6; k++) cfrc[i][k] += tmp1[k];
  }
  for (int i = HMJ_NBODY - 1; i > 0; i--) {
    int pid = m->body_parentid[i];
    if (pid)
      for (int k = 0; k < 6; k++) cfrc[pid][k] += cfrc[i][k];
  }
  for (int i = 0; i < HMJ_NV; i++) d->qfrc_bias[i] = dot6(d->cdof[i], cfrc[m->dof_bodyid[i]]);
}

void orc_fwd_velocity(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  com_vel(m, d);
  passive(m, p, d);
  orc_reference_constraint(m, p, d);
  rne(m, d);
  d->flops_smooth += 3000;
}

/* ------------------------------------------------------------------ mj_subtreeVel (linear part) */
void orc_subtree_vel(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  for (int i = 0; i < HMJ_NBODY; i++) {
    double dx[3], dv[3];
    const double* com = d->subtree_com[m->body_rootid[i]];
    dx[0] = d->xipos[i][0] - com[0]; dx[1] = d->xipos[i][1] - com[1]; dx[2] = d->xipos[i][2] - com[2];
    orc_cross(dv, d->cvel[i], dx);
    for (int k = 0; k < 3; k++) d->subtree_linvel[i][k] = p->body_mass[i] * (d->cvel[i][3 + k] + dv[k]);
  }
  for (int i = HMJ_NBODY - 1; i >= 0; i--) {
    if (i) {
      int pid = m->body_parentid[i];
      for (int k = 0; k < 3; k++) d->subtree_linvel[pid][k] += d->subtree_linvel[i][k];
    }
    double s = 1.0 / fmax(MJ_MINVAL, m->body_subtreemass[i]);
    for (int k = 0; k < 3; k++) d->subtree_linvel[i][k] *= s;
  }
  d->flops_smooth += 400;
}

/* ------------------------------------------------------------------ mj_fwdActuation */
void orc_fwd_actuation(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  memset(d->qfrc_actuator, 0, sizeof(d->qfrc_actuator));
  for (int i = 0; i < HMJ_NU; i++) {
    double ctrl = d->ctrl[i];
    if (ctrl < m->act_ctrlrange[i][0]) ctrl = m->act_ctrlrange[i][0];
    else if (ctrl > m->act_ctrlrange[i][1]) ctrl = m->act_ctrlrange[i][1];
    double force = p->act_gainprm0[i] * ctrl; /* motor: fixed gain, no bias (biasprm ignored) */
    if (force < p->act_forcerange[i][0]) force = p->act_forcerange[i][0];
    else if (force > p->act_forcerange[i][1]) force = p->act_forcerange[i][1];
    d->actuator_force[i] = force;
    d->qfrc_actuator[m->act_dofadr[i]] += m->act_gear[i] * force;
  }
  d->flops_smooth += 100;
}

/* ------------------------------------------------------------------ mj_jac */
void orc_jac(const hmj_model_t* m, const orc_data_t* d, double* jacp, double* jacr, const double* point, int body) {
  double off[3];
  if (jacp) memset(jacp, 0, 3 * HMJ_NV * sizeof(double));
  if (jacr) memset(jacr, 0, 3 * HMJ_NV * sizeof(double));
  const double* com = d->subtree_com[m->body_rootid[body]];
  off[0] = point[0] - com[0]; off[1] = point[1] - com[1]; off[2] = point[2] - com[2];
  while (body && !m->body_dofnum[body]) body = m->body_parentid[body];
  if (!body) return;
  int da = m->body_dofadr[body] + m->body_dofnum[body] - 1;
  while (da >= 0) {
    const double* cd = d->cdof[da];
    if (jacr) {
      jacr[da] = cd[0]; jacr[da + HMJ_NV] = cd[1]; jacr[da + 2 * HMJ_NV] = cd[2];
    }
    if (jacp) {
      double tmp[3];
      orc_cross(tmp, cd, off);
      jacp[da] = cd[3] + tmp[0];
      jacp[da + HMJ_NV] = cd[4] + tmp[1];
      jacp[da + 2 * HMJ_NV] = cd[5] + tmp[2];
    }
    da = m->dof_parentid[da];
  }
}

/* ------------------------------------------------------------------ mj_fwdAcceleration */
void orc_fwd_acceleration(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  (void)p;
  for (int i = 0; i < HMJ_NV; i++) {
    d->qfrc_smooth[i] = d->qfrc_passive[i] - d->qfrc_bias[i];
    d->qfrc_smooth[i] += 0.0; /* qfrc_applied: never set by the reference */
    d->qfrc_smooth[i] += d->qfrc_actuator[i];
  }
  /* mj_xfrcAccumulate: Cartesian pushes (cpu_env.py:618-654) applied at the body CoM */
  for (int b = 1; b < HMJ_NBODY; b++) {
    const double* f = d->xfrc_applied[b];
    if (f[0] == 0 && f[1] == 0 && f[2] == 0 && f[3] == 0 && f[4] == 0 && f[5] == 0) continue;
    double jacp[3 * HMJ_NV], jacr[3 * HMJ_NV];
    orc_jac(m, d, jacp, jacr, d->xipos[b], b);
    for (int i = 0; i < HMJ_NV; i++) {
      d->qfrc_smooth[i] += jacp[i] * f[0] + jacp[i + HMJ_NV] * f[1] + jacp[i + 2 * HMJ_NV] * f[2];
      d->qfrc_smooth[i] += jacr[i] * f[3] + jacr[i + HMJ_NV] * f[4] + jacr[i + 2 * HMJ_NV] * f[5];
    }
  }
  memcpy(d->qacc_smooth, d->qfrc_smooth, sizeof(d->qacc_smooth));
  orc_solve_ld(m, d->qacc_smooth, d->qLD, d->qLDiagInv);
  d->flops_smooth += 800;
}

/* ------------------------------------------------------------------ mj_Euler (implicit joint damping) */
static void quat_integrate(double* quat, const double* vel, double scale) {
  double tmp[3] = {vel[0], vel[1], vel[2]}, qrot[4];
  double angle = scale * orc_normalize3(tmp);
  axis_angle2quat(qrot, tmp, angle);
  orc_normalize4(quat);
  orc_mul_quat(quat, quat, qrot);
}

void orc_euler(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  double h = p->timestep;
  double qH[HMJ_NM], qHDiagInv[HMJ_NV], qacc[HMJ_NV];
  int damping = 0;
  for (int i = 0; i < HMJ_NV; i++)
    if (m->dof_damping[i] > 0) { damping = 1; break; }
  if (!damping) {
    memcpy(qacc, d->qacc, sizeof(qacc));
  } else {
    memcpy(qH, d->qM, sizeof(qH));
    for (int i = 0; i < HMJ_NV; i++) qH[m->dof_Madr[i]] += h * m->dof_damping[i];
    orc_factor_i(m, qH, qHDiagInv, &d->status);
    for (int i = 0; i < HMJ_NV; i++) qacc[i] = d->qfrc_smooth[i] + d->qfrc_constraint[i];
    orc_solve_ld(m, qacc, qH, qHDiagInv);
  }
  /* mj_advance */
  for (int i = 0; i < HMJ_NV; i++) d->qvel[i] += h * qacc[i];
  for (int j = 0; j < HMJ_NJNT; j++) {
    int padr = m->jnt_qposadr[j], vadr = m->jnt_dofadr[j];
    if (m->jnt_type[j] == HMJ_JNT_FREE) {
      for (int i = 0; i < 3; i++) d->qpos[padr + i] += h * d->qvel[vadr + i];
      quat_integrate(d->qpos + padr + 3, d->qvel + vadr + 3, h);
    } else {
      d->qpos[padr] += h * d->qvel[vadr];
    }
  }
  d->time += h;
  d->flops_smooth += 2200;
}
