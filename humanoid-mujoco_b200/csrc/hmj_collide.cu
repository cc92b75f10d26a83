// hmj_collide.cu — kernel 2 of the step: narrowphase for the 37 candidate geom pairs.
//
// Restates mj_collision (simulation/mujoco/include/mujoco/mujoco.h:355) for this model:
// mjc_Convex (mesh-mesh) and mjc_ConvexHField (mesh vs height-field prisms), both via libccd's
// MPR penetration query (mpr_tolerance 1e-6, mpr_iterations 50: mjmodel.h:413,436).
//
// Mapping: one warp per (environment, candidate pair).  The portal logic is scalar and runs
// redundantly on all 32 lanes (uniform control flow, no divergence); the hull support function
// -- >90 % of the arithmetic -- is lane-parallel: each lane scans vertices lane, lane+32, ...
// (SoA float32, coalesced) and a shuffle arg-max with lowest-index tie-break reproduces the
// serial "first maximum" exactly.
//
// THIS TRANSLATION UNIT IS COMPILED WITH -fmad=false: contact count and geom pairs must be
// bit-identical to the reference path, so every expression is the same sequence of IEEE-754
// operations as the oracle (oracle/hmj_oracle_collision.c, built with -ffp-contract=off).
// Roofline: FP64 add/mul issue bound (no FMA by contract); HBM/L2 traffic is 120 doubles of
// geom frames in and <= 8 doubles per contact slot out.
#include <float.h>

#include "hmj_dev.cuh"

#define CCD_EPS DBL_EPSILON

struct Supp { double v[3], v1[3], v2[3]; };

struct MeshObj {
  const float* vx; const float* vy; const float* vz;
  int npad;
  double pos[3];
  double mat[9];
};

struct PrismObj { double p[6][3]; };

// ------------------------------------------------------------------ libccd vec3 helpers
__device__ __forceinline__ bool c_is_zero(double v) { return fabs(v) < CCD_EPS; }

__device__ __forceinline__ bool c_eq(double _a, double _b) {
  double ab = fabs(_a - _b);
  if (fabs(ab) < CCD_EPS) return true;
  double a = fabs(_a), b = fabs(_b);
  if (b > a) return ab < CCD_EPS * b;
  return ab < CCD_EPS * a;
}

__device__ __forceinline__ bool c_vec_is_origin(const double* a) { return c_eq(a[0], 0.0) && c_eq(a[1], 0.0) && c_eq(a[2], 0.0); }

__device__ __forceinline__ double c_dot(const double* a, const double* b) {
  double dot = a[0] * b[0];
  dot += a[1] * b[1];
  dot += a[2] * b[2];
  return dot;
}

__device__ __forceinline__ void c_sub2(double* d, const double* a, const double* b) {
  d[0] = a[0] - b[0]; d[1] = a[1] - b[1]; d[2] = a[2] - b[2];
}

__device__ __forceinline__ void c_cross(double* d, const double* a, const double* b) {
  double x = (a[1] * b[2]) - (a[2] * b[1]);
  double y = (a[2] * b[0]) - (a[0] * b[2]);
  double z = (a[0] * b[1]) - (a[1] * b[0]);
  d[0] = x; d[1] = y; d[2] = z;
}

__device__ __forceinline__ void c_scale(double* d, double k) { d[0] *= k; d[1] *= k; d[2] *= k; }

__device__ __forceinline__ void c_normalize(double* d) {
  double k = 1.0 / sqrt(c_dot(d, d));
  c_scale(d, k);
}

__device__ __forceinline__ double c_dist2(const double* a, const double* b) {
  double ab[3];
  c_sub2(ab, a, b);
  return c_dot(ab, ab);
}

// ------------------------------------------------------------------ support functions
__device__ __forceinline__ void mesh_support(const MeshObj& o, const double* dir, double* out, int lane) {
  const double* mat = o.mat;
  const double ld0 = mat[0] * dir[0] + mat[3] * dir[1] + mat[6] * dir[2];
  const double ld1 = mat[1] * dir[0] + mat[4] * dir[1] + mat[7] * dir[2];
  const double ld2 = mat[2] * dir[0] + mat[5] * dir[1] + mat[8] * dir[2];
  double best = -HMJ_MAXVAL;
  int ibest = 0x7fffffff;
  for (int i = lane; i < o.npad; i += 32) {
    double vd = ld0 * (double)o.vx[i] + ld1 * (double)o.vy[i] + ld2 * (double)o.vz[i];
    if (vd > best) { best = vd; ibest = i; }
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    double ob = __shfl_xor_sync(0xffffffffu, best, s);
    int oi = __shfl_xor_sync(0xffffffffu, ibest, s);
    if (ob > best || (ob == best && oi < ibest)) { best = ob; ibest = oi; }
  }
  if (ibest == 0x7fffffff) ibest = 0;
  const double p0 = (double)o.vx[ibest], p1 = (double)o.vy[ibest], p2 = (double)o.vz[ibest];
  out[0] = (mat[0] * p0 + mat[1] * p1 + mat[2] * p2) + o.pos[0];
  out[1] = (mat[3] * p0 + mat[4] * p1 + mat[5] * p2) + o.pos[1];
  out[2] = (mat[6] * p0 + mat[7] * p1 + mat[8] * p2) + o.pos[2];
}

__device__ __forceinline__ void prism_support(const PrismObj& o, const double* dir, double* out) {
  const bool top = !(dir[2] < 0);
  double a[3], b[3], c[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    a[k] = top ? o.p[3][k] : o.p[0][k];
    b[k] = top ? o.p[4][k] : o.p[1][k];
    c[k] = top ? o.p[5][k] : o.p[2][k];
  }
  double best = a[0] * dir[0] + a[1] * dir[1] + a[2] * dir[2];
  out[0] = a[0]; out[1] = a[1]; out[2] = a[2];
  double t = b[0] * dir[0] + b[1] * dir[1] + b[2] * dir[2];
  if (t > best) { best = t; out[0] = b[0]; out[1] = b[1]; out[2] = b[2]; }
  t = c[0] * dir[0] + c[1] * dir[1] + c[2] * dir[2];
  if (t > best) { best = t; out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; }
}

template <bool PRISM>
struct Pair {
  PrismObj prism;  // used when PRISM
  MeshObj m1;      // used when !PRISM
  MeshObj m2;
  int lane;

  __device__ __forceinline__ void center1(double* c) const {
    if (PRISM) {
      c[0] = c[1] = c[2] = 0;
#pragma unroll
      for (int i = 0; i < 6; i++) { c[0] += prism.p[i][0]; c[1] += prism.p[i][1]; c[2] += prism.p[i][2]; }
      c[0] *= 1.0 / 6.0; c[1] *= 1.0 / 6.0; c[2] *= 1.0 / 6.0;
    } else {
      c[0] = m1.pos[0]; c[1] = m1.pos[1]; c[2] = m1.pos[2];
    }
  }
  __device__ __forceinline__ void support(const double* dir, Supp& s) const {
    double nd[3] = {dir[0], dir[1], dir[2]};
    if (PRISM) prism_support(prism, dir, s.v1); else mesh_support(m1, dir, s.v1, lane);
    c_scale(nd, -1.0);
    mesh_support(m2, nd, s.v2, lane);
    c_sub2(s.v, s.v1, s.v2);
  }
};

// ------------------------------------------------------------------ MPR (libccd mpr.c)
__device__ __forceinline__ void portal_dir(const Supp& p1, const Supp& p2, const Supp& p3, double* dir) {
  double v2v1[3], v3v1[3];
  c_sub2(v2v1, p2.v, p1.v);
  c_sub2(v3v1, p3.v, p1.v);
  c_cross(dir, v2v1, v3v1);
  c_normalize(dir);
}

__device__ __forceinline__ bool portal_reach_tolerance(const Supp& p1, const Supp& p2, const Supp& p3, const Supp& v4,
                                                       const double* dir, double tol) {
  double dv1 = c_dot(p1.v, dir), dv2 = c_dot(p2.v, dir), dv3 = c_dot(p3.v, dir), dv4 = c_dot(v4.v, dir);
  double dot1 = dv4 - dv1, dot2 = dv4 - dv2, dot3 = dv4 - dv3;
  dot1 = fmin(dot1, dot2);
  dot1 = fmin(dot1, dot3);
  return c_eq(dot1, tol) || dot1 < tol;
}

__device__ __forceinline__ void expand_portal(const Supp& p0, Supp& p1, Supp& p2, Supp& p3, const Supp& v4) {
  double v4v0[3];
  c_cross(v4v0, v4.v, p0.v);
  double dot = c_dot(p1.v, v4v0);
  if (dot > 0) {
    dot = c_dot(p2.v, v4v0);
    if (dot > 0) p1 = v4; else p3 = v4;
  } else {
    dot = c_dot(p3.v, v4v0);
    if (dot > 0) p2 = v4; else p1 = v4;
  }
}

__device__ __forceinline__ double point_seg_dist2(const double* P, const double* x0, const double* b, double* witness) {
  double d[3], a[3], dist, t;
  c_sub2(d, b, x0);
  c_sub2(a, x0, P);
  t = -1.0 * c_dot(a, d);
  t /= c_dot(d, d);
  if (t < 0 || c_is_zero(t)) {
    dist = c_dist2(x0, P);
    witness[0] = x0[0]; witness[1] = x0[1]; witness[2] = x0[2];
  } else if (t > 1 || c_eq(t, 1.0)) {
    dist = c_dist2(b, P);
    witness[0] = b[0]; witness[1] = b[1]; witness[2] = b[2];
  } else {
    witness[0] = d[0]; witness[1] = d[1]; witness[2] = d[2];
    c_scale(witness, t);
    witness[0] += x0[0]; witness[1] += x0[1]; witness[2] += x0[2];
    dist = c_dist2(witness, P);
  }
  return dist;
}

__device__ __forceinline__ double point_tri_dist2(const double* P, const double* x0, const double* Bv, const double* C, double* witness) {
  double d1[3], d2[3], a[3], v, w, p, q, r, dd, s, t, dist, dist2, witness2[3];
  c_sub2(d1, Bv, x0);
  c_sub2(d2, C, x0);
  c_sub2(a, x0, P);
  v = c_dot(d1, d1);
  w = c_dot(d2, d2);
  p = c_dot(a, d1);
  q = c_dot(a, d2);
  r = c_dot(d1, d2);
  dd = w * v - r * r;
  if (c_is_zero(dd)) {
    s = t = -1.0;
  } else {
    s = (q * r - w * p) / dd;
    t = (-s * r - q) / w;
  }
  if ((c_is_zero(s) || s > 0) && (c_eq(s, 1.0) || s < 1) && (c_is_zero(t) || t > 0) && (c_eq(t, 1.0) || t < 1) &&
      (c_eq(t + s, 1.0) || t + s < 1)) {
    c_scale(d1, s);
    c_scale(d2, t);
    witness[0] = x0[0]; witness[1] = x0[1]; witness[2] = x0[2];
    witness[0] += d1[0]; witness[1] += d1[1]; witness[2] += d1[2];
    witness[0] += d2[0]; witness[1] += d2[1]; witness[2] += d2[2];
    dist = c_dist2(witness, P);
  } else {
    dist = point_seg_dist2(P, x0, Bv, witness);
    dist2 = point_seg_dist2(P, x0, C, witness2);
    if (dist2 < dist) { dist = dist2; witness[0] = witness2[0]; witness[1] = witness2[1]; witness[2] = witness2[2]; }
    dist2 = point_seg_dist2(P, Bv, C, witness2);
    if (dist2 < dist) { dist = dist2; witness[0] = witness2[0]; witness[1] = witness2[1]; witness[2] = witness2[2]; }
  }
  return dist;
}

__device__ __forceinline__ void find_pos(const Supp& p0, const Supp& p1, const Supp& p2, const Supp& p3, double* pos) {
  double dir[3], b0, b1, b2, b3, sum, inv, vec[3], q1[3] = {0, 0, 0}, q2[3] = {0, 0, 0};
  portal_dir(p1, p2, p3, dir);
  c_cross(vec, p1.v, p2.v); b0 = c_dot(vec, p3.v);
  c_cross(vec, p3.v, p2.v); b1 = c_dot(vec, p0.v);
  c_cross(vec, p0.v, p1.v); b2 = c_dot(vec, p3.v);
  c_cross(vec, p2.v, p1.v); b3 = c_dot(vec, p0.v);
  sum = b0 + b1 + b2 + b3;
  if (c_is_zero(sum) || sum < 0) {
    b0 = 0;
    c_cross(vec, p2.v, p3.v); b1 = c_dot(vec, dir);
    c_cross(vec, p3.v, p1.v); b2 = c_dot(vec, dir);
    c_cross(vec, p1.v, p2.v); b3 = c_dot(vec, dir);
    sum = b1 + b2 + b3;
  }
  inv = 1.0 / sum;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    q1[k] += p0.v1[k] * b0; q2[k] += p0.v2[k] * b0;
    q1[k] += p1.v1[k] * b1; q2[k] += p1.v2[k] * b1;
    q1[k] += p2.v1[k] * b2; q2[k] += p2.v2[k] * b2;
    q1[k] += p3.v1[k] * b3; q2[k] += p3.v2[k] * b3;
  }
  c_scale(q1, inv);
  c_scale(q2, inv);
  pos[0] = q1[0] + q2[0]; pos[1] = q1[1] + q2[1]; pos[2] = q1[2] + q2[2];
  c_scale(pos, 0.5);
}

// ccdMPRPenetration: returns 0 when penetrating (depth, dir, pos filled), -1 when separated
template <bool PRISM>
__device__ int mpr_penetration(const Pair<PRISM>& pr, double tol, int maxit, double* depth, double* pdir, double* pos) {
  Supp p0, p1, p2, p3, v4;
  double dir[3], va[3], vb[3], dot;
  // ---- discoverPortal
  pr.center1(p0.v1);
  p0.v2[0] = pr.m2.pos[0]; p0.v2[1] = pr.m2.pos[1]; p0.v2[2] = pr.m2.pos[2];
  c_sub2(p0.v, p0.v1, p0.v2);
  if (c_vec_is_origin(p0.v)) p0.v[0] += CCD_EPS * 10.0;
  dir[0] = p0.v[0]; dir[1] = p0.v[1]; dir[2] = p0.v[2];
  c_scale(dir, -1.0);
  c_normalize(dir);
  pr.support(dir, p1);
  dot = c_dot(p1.v, dir);
  if (c_is_zero(dot) || dot < 0) return -1;

  c_cross(dir, p0.v, p1.v);
  if (c_is_zero(c_dot(dir, dir))) {
    if (c_vec_is_origin(p1.v)) {  // touching contact on v1
      *depth = 0;
      pdir[0] = pdir[1] = pdir[2] = 0;
      for (int k = 0; k < 3; k++) pos[k] = (p1.v1[k] + p1.v2[k]) * 0.5;
    } else {  // origin on the v0-v1 segment
      for (int k = 0; k < 3; k++) pos[k] = (p1.v1[k] + p1.v2[k]) * 0.5;
      pdir[0] = p1.v[0]; pdir[1] = p1.v[1]; pdir[2] = p1.v[2];
      *depth = sqrt(c_dot(pdir, pdir));
      c_normalize(pdir);
    }
    return 0;
  }
  c_normalize(dir);
  pr.support(dir, p2);
  dot = c_dot(p2.v, dir);
  if (c_is_zero(dot) || dot < 0) return -1;

  c_sub2(va, p1.v, p0.v);
  c_sub2(vb, p2.v, p0.v);
  c_cross(dir, va, vb);
  c_normalize(dir);
  dot = c_dot(dir, p0.v);
  if (dot > 0) {
    Supp t = p1; p1 = p2; p2 = t;
    c_scale(dir, -1.0);
  }
  int guard = 0;
  bool have4 = false;
  while (!have4) {
    if (++guard > 1000) return -1;
    pr.support(dir, p3);
    dot = c_dot(p3.v, dir);
    if (c_is_zero(dot) || dot < 0) return -1;
    bool cont = false;
    c_cross(va, p1.v, p3.v);
    dot = c_dot(va, p0.v);
    if (dot < 0 && !c_is_zero(dot)) { p2 = p3; cont = true; }
    if (!cont) {
      c_cross(va, p3.v, p2.v);
      dot = c_dot(va, p0.v);
      if (dot < 0 && !c_is_zero(dot)) { p1 = p3; cont = true; }
    }
    if (cont) {
      c_sub2(va, p1.v, p0.v);
      c_sub2(vb, p2.v, p0.v);
      c_cross(dir, va, vb);
      c_normalize(dir);
    } else {
      have4 = true;
    }
  }
  // ---- refinePortal
  guard = 0;
  while (true) {
    if (++guard > 1000) return -1;
    portal_dir(p1, p2, p3, dir);
    dot = c_dot(dir, p1.v);
    if (c_is_zero(dot) || dot > 0) break;  // portal encapsules origin
    pr.support(dir, v4);
    dot = c_dot(v4.v, dir);
    if (!(c_is_zero(dot) || dot > 0) || portal_reach_tolerance(p1, p2, p3, v4, dir, tol)) return -1;
    expand_portal(p0, p1, p2, p3, v4);
  }
  // ---- findPenetr
  const double origin[3] = {0, 0, 0};
  unsigned long iterations = 0;
  while (true) {
    portal_dir(p1, p2, p3, dir);
    pr.support(dir, v4);
    if (portal_reach_tolerance(p1, p2, p3, v4, dir, tol) || iterations > (unsigned long)maxit) {
      *depth = point_tri_dist2(origin, p1.v, p2.v, p3.v, pdir);
      *depth = sqrt(*depth);
      if (c_is_zero(*depth)) {
        pdir[0] = pdir[1] = pdir[2] = 0;
      } else {
        c_normalize(pdir);
      }
      find_pos(p0, p1, p2, p3, pos);
      return 0;
    }
    expand_portal(p0, p1, p2, p3, v4);
    iterations++;
  }
}

__device__ __forceinline__ void set_mesh(MeshObj& o, const hmj_tables_t* T, int mesh) {
  const int adr = T->mesh_adr[mesh];
  o.vx = T->vx + adr; o.vy = T->vy + adr; o.vz = T->vz + adr;
  o.npad = T->mesh_npad[mesh];
}

__device__ __forceinline__ void write_slot(double* slot, double dist, const double* nrm, const double* pos, int lane) {
  if (lane == 0) {
    slot[0] = dist;
    slot[1] = pos[0]; slot[2] = pos[1]; slot[3] = pos[2];
    slot[4] = nrm[0]; slot[5] = nrm[1]; slot[6] = nrm[2];
    slot[7] = 1.0;
  }
}

__global__ void __launch_bounds__(32 * HMJ_WARPS_PER_CTA)
k_collide(const hmj_tables_t* __restrict__ T, hmj_buffers_t B) {
  const int lane = threadIdx.x & 31;
  const long long w = (long long)blockIdx.x * HMJ_WARPS_PER_CTA + (threadIdx.x >> 5);
  const int env = (int)(w / HMJ_NPAIR);
  const int k = (int)(w % HMJ_NPAIR);
  if (env >= B.n) return;
  const hmj_model_t& M = T->m;
  const int g1 = M.pair_geom1[k], g2 = M.pair_geom2[k];
  const int slot0 = T->pair_slot[k];
  const int nslot = T->pair_slot[k + 1] - slot0;
  double* slots = B.slots + ((size_t)env * HMJ_NSLOT + slot0) * HMJ_SLOT_DOUBLES;
  const double* gx = B.geom_xpos + (size_t)env * HMJ_NGEOM * 3;
  const double* gm = B.geom_xmat + (size_t)env * HMJ_NGEOM * 9;
  const double tol = M.mpr_tolerance;
  const int maxit = M.mpr_iterations;
  int cnt = 0;

  if (g1 != 0) {
    // ---------------- mesh-mesh: bounding spheres, then one MPR query (mjc_Convex, multiccd off)
    const double dif0 = gx[g1 * 3] - gx[g2 * 3], dif1 = gx[g1 * 3 + 1] - gx[g2 * 3 + 1], dif2 = gx[g1 * 3 + 2] - gx[g2 * 3 + 2];
    const double bound = M.geom_rbound[g1] + M.geom_rbound[g2];
    if (!(dif0 * dif0 + dif1 * dif1 + dif2 * dif2 > bound * bound)) {
      Pair<false> pr;
      pr.lane = lane;
      set_mesh(pr.m1, T, M.geom_meshid[g1]);
      set_mesh(pr.m2, T, M.geom_meshid[g2]);
      for (int q = 0; q < 3; q++) { pr.m1.pos[q] = gx[g1 * 3 + q]; pr.m2.pos[q] = gx[g2 * 3 + q]; }
      for (int q = 0; q < 9; q++) { pr.m1.mat[q] = gm[g1 * 9 + q]; pr.m2.mat[q] = gm[g2 * 9 + q]; }
      double depth, dir[3], pos[3];
      if (mpr_penetration<false>(pr, tol, maxit, &depth, dir, pos) == 0 && !c_vec_is_origin(dir)) {
        write_slot(slots, 0.0 - depth, dir, pos, lane);
        cnt = 1;
      }
    }
  } else {
    // ---------------- height field (geom 0) vs mesh g2: mjc_ConvexHField
    const double* pos1 = gx;
    const double* mat1 = gm;
    const double* size1 = M.hfield_size;
    const int nrow = M.hfield_nrow, ncol = M.hfield_ncol;
    const float* data = B.params[env].hfield_data;
    Pair<true> pr;
    pr.lane = lane;
    set_mesh(pr.m2, T, M.geom_meshid[g2]);
    double vec[3] = {gx[g2 * 3] - pos1[0], gx[g2 * 3 + 1] - pos1[1], gx[g2 * 3 + 2] - pos1[2]};
    double* pos = pr.m2.pos;
    pos[0] = mat1[0] * vec[0] + mat1[3] * vec[1] + mat1[6] * vec[2];
    pos[1] = mat1[1] * vec[0] + mat1[4] * vec[1] + mat1[7] * vec[2];
    pos[2] = mat1[2] * vec[0] + mat1[5] * vec[1] + mat1[8] * vec[2];
    const double r2 = M.geom_rbound[g2];
    bool ok = true;
    for (int i = 0; i < 2; i++)
      if ((size1[i] < pos[i] - r2 - 0.0) || (-size1[i] > pos[i] + r2 + 0.0)) ok = false;
    if (size1[2] < pos[2] - r2 - 0.0) ok = false;
    if (-size1[3] > pos[2] + r2 + 0.0) ok = false;
    if (ok) {
      const double* mat2 = gm + g2 * 9;
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
          pr.m2.mat[3 * i + j] = mat1[i] * mat2[j] + mat1[3 + i] * mat2[3 + j] + mat1[6 + i] * mat2[6 + j];
      double out[3], xmin, xmax, ymin, ymax, zmin, zmax;
      { double dv[3] = {1, 0, 0};  mesh_support(pr.m2, dv, out, lane); xmax = out[0]; }
      { double dv[3] = {-1, 0, 0}; mesh_support(pr.m2, dv, out, lane); xmin = out[0]; }
      { double dv[3] = {0, 1, 0};  mesh_support(pr.m2, dv, out, lane); ymax = out[1]; }
      { double dv[3] = {0, -1, 0}; mesh_support(pr.m2, dv, out, lane); ymin = out[1]; }
      { double dv[3] = {0, 0, 1};  mesh_support(pr.m2, dv, out, lane); zmax = out[2]; }
      { double dv[3] = {0, 0, -1}; mesh_support(pr.m2, dv, out, lane); zmin = out[2]; }
      if (!((xmin > size1[0]) || (xmax < -size1[0]) || (ymin > size1[1]) || (ymax < -size1[1]) ||
            (zmin > size1[2]) || (zmax < -size1[3]))) {
        int cmin = (int)floor((xmin + size1[0]) / (2 * size1[0]) * (ncol - 1));
        int cmax = (int)ceil((xmax + size1[0]) / (2 * size1[0]) * (ncol - 1));
        int rmin = (int)floor((ymin + size1[1]) / (2 * size1[1]) * (nrow - 1));
        int rmax = (int)ceil((ymax + size1[1]) / (2 * size1[1]) * (nrow - 1));
        cmin = max(0, cmin); cmax = min(ncol - 1, cmax);
        rmin = max(0, rmin); rmax = min(nrow - 1, rmax);
        const double dx = (2.0 * size1[0]) / (ncol - 1);
        const double dy = (2.0 * size1[1]) / (nrow - 1);
        for (int i = 0; i < 6; i++) pr.prism.p[i][0] = pr.prism.p[i][1] = pr.prism.p[i][2] = 0.0;
        pr.prism.p[0][2] = pr.prism.p[1][2] = pr.prism.p[2][2] = -size1[3];
        bool full = false;
        for (int r = rmin; r < rmax && !full; r++) {
          int nvert = 0;
          for (int c = cmin; c <= cmax && !full; c++) {
            for (int i = 0; i < 2 && !full; i++) {
              const int rr = r + (i == 0 ? 1 : 0);
              const double x = dx * c - size1[0], y = dy * rr - size1[1];
              const double z = (double)data[rr * ncol + c] * size1[2] + 0.0;
#pragma unroll
              for (int q = 0; q < 3; q++) {
                pr.prism.p[0][q] = pr.prism.p[1][q]; pr.prism.p[1][q] = pr.prism.p[2][q];
                pr.prism.p[3][q] = pr.prism.p[4][q]; pr.prism.p[4][q] = pr.prism.p[5][q];
              }
              pr.prism.p[2][0] = x; pr.prism.p[2][1] = y;
              pr.prism.p[5][0] = x; pr.prism.p[5][1] = y; pr.prism.p[5][2] = z;
              nvert++;
              if (nvert > 2) {
                if (pr.prism.p[3][2] < zmin && pr.prism.p[4][2] < zmin && pr.prism.p[5][2] < zmin) continue;
                double depth, dirv[3], posv[3];
                if (mpr_penetration<true>(pr, tol, maxit, &depth, dirv, posv) == 0 && !d_is_bad(depth)) {
                  double nrm[3], cp[3];
                  d_rot_vec_mat(nrm, dirv, mat1);
                  d_rot_vec_mat(cp, posv, mat1);
                  cp[0] += pos1[0]; cp[1] += pos1[1]; cp[2] += pos1[2];
                  if (nrm[0] == 0 && nrm[1] == 0 && nrm[2] == 0) continue;
                  write_slot(slots + (size_t)cnt * HMJ_SLOT_DOUBLES, 0.0 - depth, nrm, cp, lane);
                  cnt++;
                  if (cnt >= nslot) full = true;  // per-pair cap HMJ_MAXCONPAIR
                }
              }
            }
          }
        }
      }
    }
  }
  // clear the unused slots of this pair
  for (int s = cnt + lane; s < nslot; s += 32) slots[(size_t)s * HMJ_SLOT_DOUBLES + 7] = 0.0;
}

void hmj_launch_collide(const hmj_tables_t* tab, hmj_buffers_t buf, cudaStream_t stream) {
  const long long nwarp = (long long)buf.n * HMJ_NPAIR;
  const int grid = (int)((nwarp + HMJ_WARPS_PER_CTA - 1) / HMJ_WARPS_PER_CTA);
  k_collide<<<grid, 32 * HMJ_WARPS_PER_CTA, 0, stream>>>(tab, buf);
}
