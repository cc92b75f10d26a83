/* hmj_oracle_collision.c — oracle: collision driver + convex narrowphase.
 * TEST INFRASTRUCTURE ONLY (see hmj_oracle.h).
 *
 * Restates, for this model, MuJoCo 3.1.4's mj_collision (mujoco.h:355): fixed candidate
 * pair list in body-pair-signature order (hmj_model_t.pair_*), bounding-sphere rejection,
 * mjc_Convex (mesh-mesh) and mjc_ConvexHField (mesh-heightfield prisms), both through
 * libccd's ccdMPRPenetration (Minkowski Portal Refinement; bundled with MuJoCo), then
 * contact parameter mixing (condim 6, friction max, solref/solimp solmix-weighted) and
 * mju_makeFrame.  Tolerances: mjOption.mpr_tolerance 1e-6, mpr_iterations 50
 * (mjmodel.h:413,436).  All geoms have margin = gap = 0.
 *
 * Arithmetic order here is the contract for the CUDA narrowphase: contact COUNT and geom
 * PAIRS must match bit for bit, so every expression is written as an explicit sequence
 * of IEEE operations (the library is built with -ffp-contract=off).
 */
#include <float.h>
#include <math.h>
#include <string.h>

#include "hmj_oracle_internal.h"

#define CCD_EPS DBL_EPSILON

typedef struct { double v[3], v1[3], v2[3]; } supp_t;

typedef struct {
  int kind;             /* 0 = mesh hull, 1 = hfield prism */
  const float* verts;   /* mesh: hull vertices (local frame) */
  int nvert;
  const double* pos;    /* mesh: frame */
  const double* mat;
  double prism[6][3];   /* prism: 3 bottom + 3 top vertices */
  orc_data_t* d;        /* counters */
} cobj_t;

/* ------------------------------------------------------------------ libccd vec3 helpers */
static int is_zero(double v) { return fabs(v) < CCD_EPS; }

static int ccd_eq(double _a, double _b) {
  double ab = fabs(_a - _b);
  if (fabs(ab) < CCD_EPS) return 1;
  double a = fabs(_a), b = fabs(_b);
  if (b > a) return ab < CCD_EPS * b;
  return ab < CCD_EPS * a;
}

static int vec_eq(const double* a, const double* b) {
  return ccd_eq(a[0], b[0]) && ccd_eq(a[1], b[1]) && ccd_eq(a[2], b[2]);
}

static double vdot(const double* a, const double* b) {
  double dot = a[0] * b[0];
  dot += a[1] * b[1];
  dot += a[2] * b[2];
  return dot;
}

static void vsub2(double* d, const double* a, const double* b) {
  d[0] = a[0] - b[0]; d[1] = a[1] - b[1]; d[2] = a[2] - b[2];
}

static void vcross(double* d, const double* a, const double* b) {
  double x = (a[1] * b[2]) - (a[2] * b[1]);
  double y = (a[2] * b[0]) - (a[0] * b[2]);
  double z = (a[0] * b[1]) - (a[1] * b[0]);
  d[0] = x; d[1] = y; d[2] = z;
}

static void vscale(double* d, double k) { d[0] *= k; d[1] *= k; d[2] *= k; }

static void vnormalize(double* d) {
  double k = 1.0 / sqrt(vdot(d, d));
  vscale(d, k);
}

static double vdist2(const double* a, const double* b) {
  double ab[3];
  vsub2(ab, a, b);
  return vdot(ab, ab);
}

/* ------------------------------------------------------------------ support / center */
static void obj_support(const cobj_t* o, const double* dir, double* out) {
  if (o->kind == 0) {
    const double* mat = o->mat;
    double ld[3];
    ld[0] = mat[0] * dir[0] + mat[3] * dir[1] + mat[6] * dir[2];
    ld[1] = mat[1] * dir[0] + mat[4] * dir[1] + mat[7] * dir[2];
    ld[2] = mat[2] * dir[0] + mat[5] * dir[1] + mat[8] * dir[2];
    int ibest = 0;
    double best = -MJ_MAXVAL;
    for (int i = 0; i < o->nvert; i++) {
      double vd = ld[0] * (double)o->verts[3 * i] + ld[1] * (double)o->verts[3 * i + 1] +
                  ld[2] * (double)o->verts[3 * i + 2];
      if (vd > best) { best = vd; ibest = i; }
    }
    double p0 = (double)o->verts[3 * ibest], p1 = (double)o->verts[3 * ibest + 1], p2 = (double)o->verts[3 * ibest + 2];
    out[0] = (mat[0] * p0 + mat[1] * p1 + mat[2] * p2) + o->pos[0];
    out[1] = (mat[3] * p0 + mat[4] * p1 + mat[5] * p2) + o->pos[1];
    out[2] = (mat[6] * p0 + mat[7] * p1 + mat[8] * p2) + o->pos[2];
    if (o->d) { o->d->n_support += 1; o->d->flops_collision += 6.0 * o->nvert + 36; }
  } else {
    int istart = dir[2] < 0 ? 0 : 3;
    int ibest = istart;
    double best = o->prism[istart][0] * dir[0] + o->prism[istart][1] * dir[1] + o->prism[istart][2] * dir[2];
    for (int i = istart + 1; i < istart + 3; i++) {
      double t = o->prism[i][0] * dir[0] + o->prism[i][1] * dir[1] + o->prism[i][2] * dir[2];
      if (t > best) { ibest = i; best = t; }
    }
    out[0] = o->prism[ibest][0]; out[1] = o->prism[ibest][1]; out[2] = o->prism[ibest][2];
    if (o->d) { o->d->n_support += 1; o->d->flops_collision += 18; }
  }
}

static void obj_center(const cobj_t* o, double* c) {
  if (o->kind == 0) {
    c[0] = o->pos[0]; c[1] = o->pos[1]; c[2] = o->pos[2];
  } else {
    c[0] = c[1] = c[2] = 0;
    for (int i = 0; i < 6; i++) { c[0] += o->prism[i][0]; c[1] += o->prism[i][1]; c[2] += o->prism[i][2]; }
    c[0] *= 1.0 / 6.0; c[1] *= 1.0 / 6.0; c[2] *= 1.0 / 6.0;
  }
}

static void mink_support(const cobj_t* o1, const cobj_t* o2, const double* dir, supp_t* s) {
  double nd[3] = {dir[0], dir[1], dir[2]};
  obj_support(o1, dir, s->v1);
  vscale(nd, -1.0);
  obj_support(o2, nd, s->v2);
  vsub2(s->v, s->v1, s->v2);
  if (o1->d) o1->d->flops_collision += 150;
}

/* ------------------------------------------------------------------ libccd MPR (mpr.c) */
static void portal_dir(const supp_t* p, double* dir) {
  double v2v1[3], v3v1[3];
  vsub2(v2v1, p[2].v, p[1].v);
  vsub2(v3v1, p[3].v, p[1].v);
  vcross(dir, v2v1, v3v1);
  vnormalize(dir);
}

static int portal_encapsules_origin(const supp_t* p, const double* dir) {
  double dot = vdot(dir, p[1].v);
  return is_zero(dot) || dot > 0;
}

static int portal_reach_tolerance(const supp_t* p, const supp_t* v4, const double* dir, double tol) {
  double dv1 = vdot(p[1].v, dir), dv2 = vdot(p[2].v, dir), dv3 = vdot(p[3].v, dir), dv4 = vdot(v4->v, dir);
  double dot1 = dv4 - dv1, dot2 = dv4 - dv2, dot3 = dv4 - dv3;
  dot1 = fmin(dot1, dot2);
  dot1 = fmin(dot1, dot3);
  return ccd_eq(dot1, tol) || dot1 < tol;
}

static int portal_can_encapsule_origin(const supp_t* v4, const double* dir) {
  double dot = vdot(v4->v, dir);
  return is_zero(dot) || dot > 0;
}

static void expand_portal(supp_t* p, const supp_t* v4) {
  double v4v0[3];
  vcross(v4v0, v4->v, p[0].v);
  double dot = vdot(p[1].v, v4v0);
  if (dot > 0) {
    dot = vdot(p[2].v, v4v0);
    if (dot > 0) p[1] = *v4; else p[3] = *v4;
  } else {
    dot = vdot(p[3].v, v4v0);
    if (dot > 0) p[2] = *v4; else p[1] = *v4;
  }
}

static int discover_portal(const cobj_t* o1, const cobj_t* o2, supp_t* p) {
  static const double origin[3] = {0, 0, 0};
  double dir[3], va[3], vb[3], dot;
  int size;
  obj_center(o1, p[0].v1);
  obj_center(o2, p[0].v2);
  vsub2(p[0].v, p[0].v1, p[0].v2);
  if (vec_eq(p[0].v, origin)) p[0].v[0] += CCD_EPS * 10.0;

  dir[0] = p[0].v[0]; dir[1] = p[0].v[1]; dir[2] = p[0].v[2];
  vscale(dir, -1.0);
  vnormalize(dir);
  mink_support(o1, o2, dir, &p[1]);
  dot = vdot(p[1].v, dir);
  if (is_zero(dot) || dot < 0) return -1;

  vcross(dir, p[0].v, p[1].v);
  if (is_zero(vdot(dir, dir))) {
    if (vec_eq(p[1].v, origin)) return 1;
    return 2;
  }
  vnormalize(dir);
  mink_support(o1, o2, dir, &p[2]);
  dot = vdot(p[2].v, dir);
  if (is_zero(dot) || dot < 0) return -1;
  size = 3;

  vsub2(va, p[1].v, p[0].v);
  vsub2(vb, p[2].v, p[0].v);
  vcross(dir, va, vb);
  vnormalize(dir);
  dot = vdot(dir, p[0].v);
  if (dot > 0) {
    supp_t t = p[1]; p[1] = p[2]; p[2] = t;
    vscale(dir, -1.0);
  }
  /* NOTE: libccd has no iteration cap here; a large cap guards the oracle against fp cycles */
  int guard = 0;
  while (size < 4) {
    if (++guard > 1000) return -1;
    mink_support(o1, o2, dir, &p[3]);
    dot = vdot(p[3].v, dir);
    if (is_zero(dot) || dot < 0) return -1;
    int cont = 0;
    vcross(va, p[1].v, p[3].v);
    dot = vdot(va, p[0].v);
    if (dot < 0 && !is_zero(dot)) { p[2] = p[3]; cont = 1; }
    if (!cont) {
      vcross(va, p[3].v, p[2].v);
      dot = vdot(va, p[0].v);
      if (dot < 0 && !is_zero(dot)) { p[1] = p[3]; cont = 1; }
    }
    if (cont) {
      vsub2(va, p[1].v, p[0].v);
      vsub2(vb, p[2].v, p[0].v);
      vcross(dir, va, vb);
      vnormalize(dir);
    } else {
      size = 4;
    }
  }
  return 0;
}

static int refine_portal(const cobj_t* o1, const cobj_t* o2, supp_t* p, double tol) {
  double dir[3];
  supp_t v4;
  int guard = 0;
  while (1) {
    if (++guard > 1000) return -1;
    portal_dir(p, dir);
    if (portal_encapsules_origin(p, dir)) return 0;
    mink_support(o1, o2, dir, &v4);
    if (!portal_can_encapsule_origin(&v4, dir) || portal_reach_tolerance(p, &v4, dir, tol)) return -1;
    expand_portal(p, &v4);
  }
}

static double point_seg_dist2(const double* P, const double* x0, const double* b, double* witness) {
  double d[3], a[3], dist, t;
  vsub2(d, b, x0);
  vsub2(a, x0, P);
  t = -1.0 * vdot(a, d);
  t /= vdot(d, d);
  if (t < 0 || is_zero(t)) {
    dist = vdist2(x0, P);
    witness[0] = x0[0]; witness[1] = x0[1]; witness[2] = x0[2];
  } else if (t > 1 || ccd_eq(t, 1.0)) {
    dist = vdist2(b, P);
    witness[0] = b[0]; witness[1] = b[1]; witness[2] = b[2];
  } else {
    witness[0] = d[0]; witness[1] = d[1]; witness[2] = d[2];
    vscale(witness, t);
    witness[0] += x0[0]; witness[1] += x0[1]; witness[2] += x0[2];
    dist = vdist2(witness, P);
  }
  return dist;
}

static double point_tri_dist2(const double* P, const double* x0, const double* B, const double* C, double* witness) {
  double d1[3], d2[3], a[3], v, w, p, q, r, dd, s, t, dist, dist2, witness2[3];
  vsub2(d1, B, x0);
  vsub2(d2, C, x0);
  vsub2(a, x0, P);
  v = vdot(d1, d1);
  w = vdot(d2, d2);
  p = vdot(a, d1);
  q = vdot(a, d2);
  r = vdot(d1, d2);
  dd = w * v - r * r;
  if (is_zero(dd)) {
    s = t = -1.0;
  } else {
    s = (q * r - w * p) / dd;
    t = (-s * r - q) / w;
  }
  if ((is_zero(s) || s > 0) && (ccd_eq(s, 1.0) || s < 1) && (is_zero(t) || t > 0) &&
      (ccd_eq(t, 1.0) || t < 1) && (ccd_eq(t + s, 1.0) || t + s < 1)) {
    vscale(d1, s);
    vscale(d2, t);
    witness[0] = x0[0]; witness[1] = x0[1]; witness[2] = x0[2];
    witness[0] += d1[0]; witness[1] += d1[1]; witness[2] += d1[2];
    witness[0] += d2[0]; witness[1] += d2[1]; witness[2] += d2[2];
    dist = vdist2(witness, P);
  } else {
    dist = point_seg_dist2(P, x0, B, witness);
    dist2 = point_seg_dist2(P, x0, C, witness2);
    if (dist2 < dist) { dist = dist2; memcpy(witness, witness2, sizeof(witness2)); }
    dist2 = point_seg_dist2(P, B, C, witness2);
    if (dist2 < dist) { dist = dist2; memcpy(witness, witness2, sizeof(witness2)); }
  }
  return dist;
}

static void find_pos(const supp_t* p, double* pos) {
  double dir[3], b[4], sum, inv, vec[3], p1[3] = {0, 0, 0}, p2[3] = {0, 0, 0};
  portal_dir(p, dir);
  vcross(vec, p[1].v, p[2].v); b[0] = vdot(vec, p[3].v);
  vcross(vec, p[3].v, p[2].v); b[1] = vdot(vec, p[0].v);
  vcross(vec, p[0].v, p[1].v); b[2] = vdot(vec, p[3].v);
  vcross(vec, p[2].v, p[1].v); b[3] = vdot(vec, p[0].v);
  sum = b[0] + b[1] + b[2] + b[3];
  if (is_zero(sum) || sum < 0) {
    b[0] = 0;
    vcross(vec, p[2].v, p[3].v); b[1] = vdot(vec, dir);
    vcross(vec, p[3].v, p[1].v); b[2] = vdot(vec, dir);
    vcross(vec, p[1].v, p[2].v); b[3] = vdot(vec, dir);
    sum = b[1] + b[2] + b[3];
  }
  inv = 1.0 / sum;
  for (int i = 0; i < 4; i++) {
    for (int k = 0; k < 3; k++) {
      p1[k] += p[i].v1[k] * b[i];
      p2[k] += p[i].v2[k] * b[i];
    }
  }
  vscale(p1, inv);
  vscale(p2, inv);
  pos[0] = p1[0] + p2[0]; pos[1] = p1[1] + p2[1]; pos[2] = p1[2] + p2[2];
  vscale(pos, 0.5);
}

static void find_penetr(const cobj_t* o1, const cobj_t* o2, supp_t* p, double tol, int maxit,
                        double* depth, double* pdir, double* pos) {
  static const double origin[3] = {0, 0, 0};
  double dir[3];
  supp_t v4;
  unsigned long iterations = 0;
  while (1) {
    portal_dir(p, dir);
    mink_support(o1, o2, dir, &v4);
    if (portal_reach_tolerance(p, &v4, dir, tol) || iterations > (unsigned long)maxit) {
      *depth = point_tri_dist2(origin, p[1].v, p[2].v, p[3].v, pdir);
      *depth = sqrt(*depth);
      if (is_zero(*depth)) {
        pdir[0] = pdir[1] = pdir[2] = 0;
      } else {
        vnormalize(pdir);
      }
      find_pos(p, pos);
      return;
    }
    expand_portal(p, &v4);
    iterations++;
  }
}

/* ccdMPRPenetration: 0 = penetrating (depth, dir, pos filled), -1 = separated */
static int mpr_penetration_impl(const cobj_t* o1, const cobj_t* o2, double tol, int maxit,
                                double* depth, double* dir, double* pos) {
  supp_t p[4];
  if (o1->d) o1->d->n_mpr += 1;
  int res = discover_portal(o1, o2, p);
  if (res < 0) return -1;
  if (res == 1) {
    *depth = 0;
    dir[0] = dir[1] = dir[2] = 0;
    for (int k = 0; k < 3; k++) pos[k] = (p[1].v1[k] + p[1].v2[k]) * 0.5;
  } else if (res == 2) {
    for (int k = 0; k < 3; k++) pos[k] = (p[1].v1[k] + p[1].v2[k]) * 0.5;
    dir[0] = p[1].v[0]; dir[1] = p[1].v[1]; dir[2] = p[1].v[2];
    *depth = sqrt(vdot(dir, dir));
    vnormalize(dir);
  } else {
    res = refine_portal(o1, o2, p, tol);
    if (res < 0) return -1;
    find_penetr(o1, o2, p, tol, maxit, depth, dir, pos);
  }
  return 0;
}

/* statistics hook (single-threaded use): histogram of Minkowski support calls per MPR query, split by outcome */
long long orc_mpr_hist[2][128];
static int mpr_penetration(const cobj_t* o1, const cobj_t* o2, double tol, int maxit,
                           double* depth, double* dir, double* pos) {
  const double before = o2->d ? o2->d->n_support : 0;
  const int res = mpr_penetration_impl(o1, o2, tol, maxit, depth, dir, pos);
  if (o2->d) {
    int n = (int)((o2->d->n_support - before) / 2);
    if (n > 127) n = 127;
    orc_mpr_hist[res == 0][n] += 1;
  }
  return res;
}

/* ------------------------------------------------------------------ contact helpers */
static void make_frame(double* frame) {
  double tmp;
  orc_normalize3(frame);
  if (sqrt(frame[3] * frame[3] + frame[4] * frame[4] + frame[5] * frame[5]) < 0.5) {
    frame[3] = frame[4] = frame[5] = 0;
    if (frame[1] < 0.5 && frame[1] > -0.5) frame[4] = 1; else frame[5] = 1;
  }
  tmp = frame[0] * frame[3] + frame[1] * frame[4] + frame[2] * frame[5];
  frame[3] -= tmp * frame[0]; frame[4] -= tmp * frame[1]; frame[5] -= tmp * frame[2];
  orc_normalize3(frame + 3);
  orc_cross(frame + 6, frame, frame + 3);
}

static void set_contact_params(const hmj_model_t* m, const hmj_params_t* p, hmj_contact_t* c, int g1, int g2) {
  const double* f1 = (g1 == 0) ? p->floor_friction : m->geom_friction[g1];
  const double* f2 = (g2 == 0) ? p->floor_friction : m->geom_friction[g2];
  double fr[3];
  for (int k = 0; k < 3; k++) fr[k] = fmax(f1[k], f2[k]);
  c->friction[0] = fr[0]; c->friction[1] = fr[0]; c->friction[2] = fr[1]; c->friction[3] = fr[2]; c->friction[4] = fr[2];
  double mix;
  double s1 = m->geom_solmix[g1], s2 = m->geom_solmix[g2];
  if (s1 >= MJ_MINVAL && s2 >= MJ_MINVAL) mix = s1 / (s1 + s2);
  else if (s1 < MJ_MINVAL && s2 < MJ_MINVAL) mix = 0.5;
  else if (s1 < MJ_MINVAL) mix = 0.0;
  else mix = 1.0;
  if (m->geom_solref[g1][0] > 0 && m->geom_solref[g2][0] > 0) {
    for (int k = 0; k < 2; k++) c->solref[k] = mix * m->geom_solref[g1][k] + (1 - mix) * m->geom_solref[g2][k];
  } else {
    for (int k = 0; k < 2; k++) c->solref[k] = fmin(m->geom_solref[g1][k], m->geom_solref[g2][k]);
  }
  for (int k = 0; k < 5; k++) c->solimp[k] = mix * m->geom_solimp[g1][k] + (1 - mix) * m->geom_solimp[g2][k];
  c->g1 = g1; c->g2 = g2;
  c->geom1 = m->geom_mjid[g1]; c->geom2 = m->geom_mjid[g2];
}

static int add_contact(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d, int g1, int g2,
                       double dist, const double* normal, const double* pos) {
  if (d->ncon >= ORC_MAXCON) { d->status |= HMJ_STATUS_CONTACTFULL; return 0; }
  hmj_contact_t* c = &d->contact[d->ncon++];
  c->dist = dist;
  memcpy(c->pos, pos, 3 * sizeof(double));
  memcpy(c->frame, normal, 3 * sizeof(double));
  memset(c->frame + 3, 0, 6 * sizeof(double));
  set_contact_params(m, p, c, g1, g2);
  make_frame(c->frame);
  d->flops_collision += 100;
  return 1;
}

static void mesh_obj(const hmj_model_t* m, orc_data_t* d, cobj_t* o, int mesh, const double* pos, const double* mat) {
  o->kind = 0;
  o->verts = &m->mesh_vert[m->mesh_vertadr[mesh]][0];
  o->nvert = m->mesh_vertnum[mesh];
  o->pos = pos; o->mat = mat; o->d = d;
}

/* mjc_Convex for two meshes: one MPR query, at most one contact (multiccd disabled) */
static void collide_mesh_mesh(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d, int g1, int g2) {
  cobj_t o1, o2;
  double depth, dir[3], pos[3];
  static const double origin[3] = {0, 0, 0};
  mesh_obj(m, d, &o1, m->geom_meshid[g1], d->geom_xpos[g1], d->geom_xmat[g1]);
  mesh_obj(m, d, &o2, m->geom_meshid[g2], d->geom_xpos[g2], d->geom_xmat[g2]);
  if (mpr_penetration(&o1, &o2, m->mpr_tolerance, m->mpr_iterations, &depth, dir, pos) == 0) {
    if (vec_eq(dir, origin)) return; /* contact found but normal undefined */
    add_contact(m, p, d, g1, g2, 0.0 - depth, dir, pos);
  }
}

/* mjc_ConvexHField: mesh g2 against the prisms of the height field g1 = 0 */
static void collide_hfield_mesh(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d, int g2) {
  const double* pos1 = d->geom_xpos[0];
  const double* mat1 = d->geom_xmat[0];
  const double* size1 = m->hfield_size;
  const int nrow = m->hfield_nrow, ncol = m->hfield_ncol;
  const float* data = p->hfield_data;
  double vec[3], pos[3], mat[9], r2, dx, dy;
  double xmin, xmax, ymin, ymax, zmin, zmax, out[3];
  cobj_t o1, o2;
  const double margin = 0.0;

  vec[0] = d->geom_xpos[g2][0] - pos1[0]; vec[1] = d->geom_xpos[g2][1] - pos1[1]; vec[2] = d->geom_xpos[g2][2] - pos1[2];
  orc_rot_vec_matT(pos, vec, mat1);
  r2 = m->geom_rbound[g2];
  for (int i = 0; i < 2; i++)
    if ((size1[i] < pos[i] - r2 - margin) || (-size1[i] > pos[i] + r2 + margin)) return;
  if (size1[2] < pos[2] - r2 - margin) return;
  if (-size1[3] > pos[2] + r2 + margin) return;

  /* mat = mat1' * mat2 : geom2 expressed in the hfield frame */
  const double* mat2 = d->geom_xmat[g2];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      mat[3 * i + j] = mat1[i] * mat2[j] + mat1[3 + i] * mat2[3 + j] + mat1[6 + i] * mat2[6 + j];

  mesh_obj(m, d, &o2, m->geom_meshid[g2], pos, mat);
  { double dirv[3] = {1, 0, 0};  obj_support(&o2, dirv, out); xmax = out[0]; }
  { double dirv[3] = {-1, 0, 0}; obj_support(&o2, dirv, out); xmin = out[0]; }
  { double dirv[3] = {0, 1, 0};  obj_support(&o2, dirv, out); ymax = out[1]; }
  { double dirv[3] = {0, -1, 0}; obj_support(&o2, dirv, out); ymin = out[1]; }
  { double dirv[3] = {0, 0, 1};  obj_support(&o2, dirv, out); zmax = out[2]; }
  { double dirv[3] = {0, 0, -1}; obj_support(&o2, dirv, out); zmin = out[2]; }
  if ((xmin - margin > size1[0]) || (xmax + margin < -size1[0]) || (ymin - margin > size1[1]) ||
      (ymax + margin < -size1[1]) || (zmin - margin > size1[2]) || (zmax + margin < -size1[3]))
    return;

  int cmin = (int)floor((xmin + size1[0]) / (2 * size1[0]) * (ncol - 1));
  int cmax = (int)ceil((xmax + size1[0]) / (2 * size1[0]) * (ncol - 1));
  int rmin = (int)floor((ymin + size1[1]) / (2 * size1[1]) * (nrow - 1));
  int rmax = (int)ceil((ymax + size1[1]) / (2 * size1[1]) * (nrow - 1));
  if (cmin < 0) cmin = 0;
  if (cmax > ncol - 1) cmax = ncol - 1;
  if (rmin < 0) rmin = 0;
  if (rmax > nrow - 1) rmax = nrow - 1;

  memset(&o1, 0, sizeof(o1));
  o1.kind = 1; o1.d = d;
  dx = (2.0 * size1[0]) / (ncol - 1);
  dy = (2.0 * size1[1]) / (nrow - 1);
  const int dr[2] = {1, 0};
  o1.prism[0][2] = o1.prism[1][2] = o1.prism[2][2] = -size1[3];

  int cnt = 0;
  for (int r = rmin; r < rmax; r++) {
    int nvert = 0;
    for (int c = cmin; c <= cmax; c++) {
      for (int i = 0; i < 2; i++) {
        /* addVert: shift the strip, append bottom/top vertex */
        double x = dx * c - size1[0], y = dy * (r + dr[i]) - size1[1];
        double z = (double)data[(r + dr[i]) * ncol + c] * size1[2] + margin;
        memcpy(o1.prism[0], o1.prism[1], 3 * sizeof(double));
        memcpy(o1.prism[1], o1.prism[2], 3 * sizeof(double));
        memcpy(o1.prism[3], o1.prism[4], 3 * sizeof(double));
        memcpy(o1.prism[4], o1.prism[5], 3 * sizeof(double));
        o1.prism[2][0] = x; o1.prism[2][1] = y;
        o1.prism[5][0] = x; o1.prism[5][1] = y; o1.prism[5][2] = z;
        nvert++;
        if (nvert > 2) {
          if (o1.prism[3][2] < zmin && o1.prism[4][2] < zmin && o1.prism[5][2] < zmin) continue;
          double depth, dirv[3], posv[3];
          if (mpr_penetration(&o1, &o2, m->mpr_tolerance, m->mpr_iterations, &depth, dirv, posv) == 0 &&
              !(isnan(depth) || depth > MJ_MAXVAL || depth < -MJ_MAXVAL)) {
            double nrm[3], cp[3];
            orc_rot_vec_mat(nrm, dirv, mat1);
            orc_rot_vec_mat(cp, posv, mat1);
            cp[0] += pos1[0]; cp[1] += pos1[1]; cp[2] += pos1[2];
            /* a zero direction (touching contact) cannot define a frame: MuJoCo's
             * mju_makeFrame would abort; such contacts are dropped here */
            if (nrm[0] == 0 && nrm[1] == 0 && nrm[2] == 0) continue;
            if (!add_contact(m, p, d, 0, g2, 0.0 - depth, nrm, cp)) return;
            cnt++;
            if (cnt >= HMJ_MAXCONPAIR) return; /* per-pair cap (MuJoCo mjMAXCONPAIR = 50) */
          }
        }
      }
    }
  }
}

void orc_collision(const hmj_model_t* m, const hmj_params_t* p, orc_data_t* d) {
  d->ncon = 0;
  for (int k = 0; k < m->npair; k++) {
    int g1 = m->pair_geom1[k], g2 = m->pair_geom2[k];
    if (g1 == 0) {
      collide_hfield_mesh(m, p, d, g2);
    } else {
      /* bounding-sphere rejection (conservative; margin = 0) */
      double dif[3] = {d->geom_xpos[g1][0] - d->geom_xpos[g2][0], d->geom_xpos[g1][1] - d->geom_xpos[g2][1],
                       d->geom_xpos[g1][2] - d->geom_xpos[g2][2]};
      double bound = m->geom_rbound[g1] + m->geom_rbound[g2];
      d->flops_collision += 10;
      if (dif[0] * dif[0] + dif[1] * dif[1] + dif[2] * dif[2] > bound * bound) continue;
      collide_mesh_mesh(m, p, d, g1, g2);
    }
  }
}

int orc_mpr_mesh_mesh(const hmj_model_t* m, int mesh1, const double pos1[3], const double mat1[9],
                      int mesh2, const double pos2[3], const double mat2[9],
                      double* depth, double dir[3], double pos[3]) {
  cobj_t o1, o2;
  mesh_obj(m, 0, &o1, mesh1, pos1, mat1);
  mesh_obj(m, 0, &o2, mesh2, pos2, mat2);
  return mpr_penetration(&o1, &o2, m->mpr_tolerance, m->mpr_iterations, depth, dir, pos);
}
