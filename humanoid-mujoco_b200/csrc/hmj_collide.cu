// hmj_collide.cu — stage 2 of the step: broadphase + narrowphase for the 37 candidate geom pairs.
//
// Restates mj_collision (simulation/mujoco/include/mujoco/mujoco.h:355) for this model:
// mjc_Convex (mesh-mesh) and mjc_ConvexHField (mesh vs height-field prisms), both via libccd's
// MPR penetration query (mpr_tolerance 1e-6, mpr_iterations 50: mjmodel.h:413,436).
//
// Two kernels:
//   k_broad   thread per (pair, env), pair warp-uniform.  Mesh-mesh: bounding spheres.  Hfield-mesh: box test,
//             hull AABB, strip walk over the prisms under the AABB, and the FIRST support call of every prism's
//             MPR query (where 64 % of them end).  Survivors are appended to task lists grouped by hull
//             (warp-aggregated atomics); the per-env slot bitmask (zeroed before the launch) marks the slots k_narrow fills.
//   k_narrow  persistent warps; a work item is up to 32 tasks OF THE SAME GROUP, ONE MPR QUERY PER LANE.  MPR is
//             written as a state machine with exactly one support call per trip, so lanes in different phases
//             (discover / refine / penetration) meet at the support call; the portal logic runs per lane.
//             A contact's slot is fixed by (pair, prism ordinal), so the contact order k_solve compacts is MuJoCo's
//             regardless of which warp produced it.
// Support calls do not scan the hull: every hull carries, per cell of a 6 x 8 x 8 cube map of directions, the list of
// vertices that can be the support point for some direction of the cell (hmj_api.cu: build_support_cells), and the
// arg-max runs over that list (2-4 vertices on average) with the scan's own expression and tie rule: same vertex, same
// bits.  Hull vertices sit in shared memory (fp64 SoA), brought in by the bulk-copy engine while the CTA sets up.
//
// THIS TRANSLATION UNIT IS COMPILED WITH -fmad=false: contact count and geom pairs must be
// bit-identical to the reference path, so every expression is the same sequence of IEEE-754
// operations as the oracle (oracle/hmj_oracle_collision.c, built with -ffp-contract=off).
// The six axis-aligned support calls of mjc_ConvexHField collapse into one min/max pass: for
// dir = +-e_r the scanned value IS the returned coordinate minus pos[r], and negating the direction
// negates every product exactly, so (max, min) of three dot products reproduce all six results.
// Roofline: FP64 add/mul issue bound (no FMA by contract); HBM/L2 traffic is 120 doubles of
// geom frames in and <= 8 doubles per contact slot out.
#include <float.h>
#include <stdlib.h>

#include "hmj_dev.cuh"

#define CCD_EPS DBL_EPSILON
#define NARROW_THREADS 128
#ifndef BROAD_THREADS
#define BROAD_THREADS 128
#endif
#ifndef NARROW_MIN_CTAS
#define NARROW_MIN_CTAS 2
#endif

// A support point of the Minkowski difference.  Its two witnesses (the support points of the two shapes) are NOT carried
// through the iteration: only the vertex indices are, and the points are recomputed — with the very expression that
// produced them — in the one place that needs them (the contact position at the end of a query).  18 doubles of per-lane
// state less for the three portal vertices, 6 more for the interior point's witnesses.
struct Supp { double v[3]; int i1, i2; };
#define NARROW_SMEM_BYTES ((3 * HMJ_NVERT_SMEM + 24 * NARROW_THREADS) * sizeof(double))

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

__device__ __forceinline__ void cp3(double* d, const double* s) { d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; }
__device__ __forceinline__ void cps(Supp& d, const Supp& s) { cp3(d.v, s.v); d.i1 = s.i1; d.i2 = s.i2; }

// ------------------------------------------------------------------ support functions
// Hull vertices of all meshes live in shared memory as fp64 SoA (sx|sy|sz), each mesh starting at an even
// index and padded to a multiple of 4 with copies of its vertex 0 (a copy can never win: ties go to the
// lowest index).  `base`/`n4` are warp-uniform, so every LDS.128 is a broadcast.
struct Hull { int base, n4; };

struct Frame {
  double pos[3]; double mat[9];
  __device__ __forceinline__ double P(int k) const { return pos[k]; }
  __device__ __forceinline__ double M(int k) const { return mat[k]; }
};
// The same frame kept in shared memory, element k of this lane's frame at b[k * NARROW_THREADS] (conflict-free): k_narrow
// holds its per-query frames there instead of in 24-48 registers per lane, which is what lets a third CTA fit on an SM.
struct SFrame {
  double* b;
  __device__ __forceinline__ double P(int k) const { return b[k * NARROW_THREADS]; }
  __device__ __forceinline__ double M(int k) const { return b[(3 + k) * NARROW_THREADS]; }
  __device__ __forceinline__ void setP(int k, double v) { b[k * NARROW_THREADS] = v; }
  __device__ __forceinline__ void setM(int k, double v) { b[(3 + k) * NARROW_THREADS] = v; }
};

#ifdef HMJ_PHASE_PROF
__device__ unsigned long long hmj_sup_stat[10];   // support calls, candidates scanned, histogram of list lengths by 8
extern "C" int hmj_debug_sup_stat(unsigned long long* out) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, hmj_sup_stat, sizeof(unsigned long long) * 10);
  return 0;
}
#endif
// ---- support through the direction-cell candidate lists (hmj_tables_t::sup_start / sup_vert, built in hmj_api.cu)
struct SupTab { const uint16_t* start; const uint16_t* vert; };   // start == nullptr: no tables, scan the hull
__device__ __forceinline__ SupTab sup_tab(const hmj_tables_t* __restrict__ T, int mesh) {
  SupTab t;
  t.start = T->sup_ok ? T->sup_start[mesh] : nullptr;
  t.vert = T->sup_vert;
  return t;
}

// Cell of the direction (d0, d1, d2) (any length), -1 when the direction is degenerate (zero, huge, not finite: the
// caller scans the whole hull).  The face comes from exact fp64 comparisons; the position on the face is found in fp32,
// which the overlap of the cells (HMJ_SUP_DELTA = 1e-3, fp32 error < 1e-6) makes safe.
__device__ __forceinline__ int sup_cell(double d0, double d1, double d2) {
  const double a0 = fabs(d0), a1 = fabs(d1), a2 = fabs(d2);
  int ax = 0;
  double am = a0;
  if (a1 > am) { ax = 1; am = a1; }
  if (a2 > am) { ax = 2; am = a2; }
  if (!(am > 1e-30 && am < 1e8)) return -1;   // |dot| stays far below HMJ_MAXVAL, the scans' start value
  const double dm = ax == 0 ? d0 : (ax == 1 ? d1 : d2);
  const double da = ax == 0 ? d1 : (ax == 1 ? d2 : d0);
  const double db = ax == 0 ? d2 : (ax == 1 ? d0 : d1);
  const float inv = 1.0f / (float)am;
  const float u = (float)da * inv, v = (float)db * inv;
  const int iu = min(max((int)((u + 1.0f) * (0.5f * HMJ_SUP_G)), 0), HMJ_SUP_G - 1);
  const int iv = min(max((int)((v + 1.0f) * (0.5f * HMJ_SUP_G)), 0), HMJ_SUP_G - 1);
  return ((ax << 1) | (dm < 0.0 ? 1 : 0)) * (HMJ_SUP_G * HMJ_SUP_G) + iu * HMJ_SUP_G + iv;
}

// arg-max of ld . v over the whole hull, first maximum (the definition the candidate lists are a shortcut for)
__device__ __noinline__ int hull_argmax(const double* __restrict__ sx, const double* __restrict__ sy,
                                        const double* __restrict__ sz, int base, int n4, double ld0, double ld1, double ld2) {
  double best = -HMJ_MAXVAL;
  int ib = 0;
  for (int i = 0; i < n4; i++) {
    const double va = ld0 * sx[base + i] + ld1 * sy[base + i] + ld2 * sz[base + i];
    if (va > best) { best = va; ib = i; }
  }
  return ib;
}

// obj_support of a mesh geom: arg-max over the candidates of the direction's cell (ascending vertex order, strict >:
// the first maximum, as the scan over the whole hull finds it — same expression per vertex, same bits).
template <typename F>
__device__ __forceinline__ void mesh_support(const double* __restrict__ sx, const double* __restrict__ sy,
                                             const double* __restrict__ sz, Hull h, SupTab tab, const F& f, const double* dir,
                                             double* out, int* idx = nullptr) {
  const double ld0 = f.M(0) * dir[0] + f.M(3) * dir[1] + f.M(6) * dir[2];
  const double ld1 = f.M(1) * dir[0] + f.M(4) * dir[1] + f.M(7) * dir[2];
  const double ld2 = f.M(2) * dir[0] + f.M(5) * dir[1] + f.M(8) * dir[2];
  const int c = tab.start ? sup_cell(ld0, ld1, ld2) : -1;
  int i0 = 0;
  if (c >= 0) {
    int k = tab.start[c];
    const int k1 = tab.start[c + 1];
#ifdef HMJ_PHASE_PROF
    atomicAdd(&hmj_sup_stat[0], 1ull); atomicAdd(&hmj_sup_stat[1], (unsigned long long)(k1 - k));
    atomicAdd(&hmj_sup_stat[2 + min((k1 - k) >> 3, 7)], 1ull);
#endif
    double best = -HMJ_MAXVAL;
    // four candidates per round, their loads and products independent of one another (the lanes of a warp wait for the
    // longest list among them); a short round repeats its last candidate, which cannot win (strict >, same value)
#pragma unroll 1
    for (; k < k1; k += 4) {
      const int e = k1 - 1;
      const int ia = tab.vert[k], ib = tab.vert[min(k + 1, e)], ic = tab.vert[min(k + 2, e)], id = tab.vert[min(k + 3, e)];
      const double va = ld0 * sx[h.base + ia] + ld1 * sy[h.base + ia] + ld2 * sz[h.base + ia];
      const double vb = ld0 * sx[h.base + ib] + ld1 * sy[h.base + ib] + ld2 * sz[h.base + ib];
      const double vc = ld0 * sx[h.base + ic] + ld1 * sy[h.base + ic] + ld2 * sz[h.base + ic];
      const double vd = ld0 * sx[h.base + id] + ld1 * sy[h.base + id] + ld2 * sz[h.base + id];
      if (va > best) { best = va; i0 = ia; }
      if (vb > best) { best = vb; i0 = ib; }
      if (vc > best) { best = vc; i0 = ic; }
      if (vd > best) { best = vd; i0 = id; }
    }
  } else {
    i0 = hull_argmax(sx, sy, sz, h.base, h.n4, ld0, ld1, ld2);
  }
  if (idx) *idx = i0;
  const double p0 = sx[h.base + i0], p1 = sy[h.base + i0], p2 = sz[h.base + i0];
  out[0] = (f.M(0) * p0 + f.M(1) * p1 + f.M(2) * p2) + f.P(0);
  out[1] = (f.M(3) * p0 + f.M(4) * p1 + f.M(5) * p2) + f.P(1);
  out[2] = (f.M(6) * p0 + f.M(7) * p1 + f.M(8) * p2) + f.P(2);
}

// extreme values of row . v over the hull: min through the cell of -row, max through the cell of +row (the extreme
// vertex is in the list and every value is computed by the same expression as in a scan of the whole hull)
__device__ __forceinline__ void hull_extent(const double* __restrict__ sx, const double* __restrict__ sy,
                                            const double* __restrict__ sz, Hull h, SupTab tab, double r0, double r1, double r2,
                                            double* mn_out, double* mx_out) {
  double mn = HMJ_MAXVAL, mx = -HMJ_MAXVAL;
  const int cp = tab.start ? sup_cell(r0, r1, r2) : -1;
  if (cp >= 0) {
    const int cm = sup_cell(-r0, -r1, -r2);
    // both lists two candidates per round, loads and products independent (a short round repeats its last candidate:
    // min / max do not care)
    int kp = tab.start[cp], km = tab.start[cm];
    const int ep = tab.start[cp + 1] - 1, em = tab.start[cm + 1] - 1;
#pragma unroll 1
    while (kp <= ep || km <= em) {
      const int ia = tab.vert[min(kp, ep)], ib = tab.vert[min(kp + 1, ep)];
      const int ic = tab.vert[min(km, em)], id = tab.vert[min(km + 1, em)];
      const double va = r0 * sx[h.base + ia] + r1 * sy[h.base + ia] + r2 * sz[h.base + ia];
      const double vb = r0 * sx[h.base + ib] + r1 * sy[h.base + ib] + r2 * sz[h.base + ib];
      const double vc = r0 * sx[h.base + ic] + r1 * sy[h.base + ic] + r2 * sz[h.base + ic];
      const double vd = r0 * sx[h.base + id] + r1 * sy[h.base + id] + r2 * sz[h.base + id];
      mx = fmax(mx, fmax(va, vb));
      mn = fmin(mn, fmin(vc, vd));
      kp += 2; km += 2;
    }
  } else {
    for (int i = 0; i < h.n4; i++) {
      const double d = r0 * sx[h.base + i] + r1 * sy[h.base + i] + r2 * sz[h.base + i];
      mn = fmin(mn, d); mx = fmax(mx, d);
    }
  }
  *mn_out = mn; *mx_out = mx;
}

// vertex i of hull h in the frame f: the tail of mesh_support, for recomputing a witness from its index (same bits)
template <typename F>
__device__ __forceinline__ void mesh_point(const double* __restrict__ sx, const double* __restrict__ sy,
                                           const double* __restrict__ sz, Hull h, const F& f, int i, double* out) {
  const double p0 = sx[h.base + i], p1 = sy[h.base + i], p2 = sz[h.base + i];
  out[0] = (f.M(0) * p0 + f.M(1) * p1 + f.M(2) * p2) + f.P(0);
  out[1] = (f.M(3) * p0 + f.M(4) * p1 + f.M(5) * p2) + f.P(1);
  out[2] = (f.M(6) * p0 + f.M(7) * p1 + f.M(8) * p2) + f.P(2);
}

// Axis-aligned bounding box of a hull in the frame `f` == the six obj_support(+-e_r) calls of
// mjc_ConvexHField (see the header comment for why min/max of three dot products is bit-identical).
__device__ __forceinline__ void mesh_aabb(const double* __restrict__ sx, const double* __restrict__ sy,
                                          const double* __restrict__ sz, Hull h, SupTab tab, const Frame& f, double* lo, double* hi) {
#pragma unroll
  for (int r = 0; r < 3; r++) {
    double mn, mx;
    hull_extent(sx, sy, sz, h, tab, f.mat[3 * r], f.mat[3 * r + 1], f.mat[3 * r + 2], &mn, &mx);
    lo[r] = mn + f.pos[r];
    hi[r] = mx + f.pos[r];
  }
}

// Height-field prism: three top vertices (strip order) over a flat bottom at z = zb with the same (x, y).
struct Prism {
  double t[3][3]; double zb;
  __device__ __forceinline__ double TT(int i, int k) const { return t[i][k]; }
  __device__ __forceinline__ double ZB() const { return zb; }
};
// the same prism in shared memory (k_narrow), element e of this lane's prism at b[e * NARROW_THREADS]
struct SPrism {
  double* b;
  __device__ __forceinline__ double TT(int i, int k) const { return b[(3 * i + k) * NARROW_THREADS]; }
  __device__ __forceinline__ double ZB() const { return b[9 * NARROW_THREADS]; }
};

// returns the vertex code: column i (0..2) + 3 when it is a top vertex
template <typename PR>
__device__ __forceinline__ int prism_support(const PR& o, const double* dir, double* out) {
  const bool top = !(dir[2] < 0);
  const double ozb = o.ZB();
  const double za = top ? o.TT(0, 2) : ozb, zb = top ? o.TT(1, 2) : ozb, zc = top ? o.TT(2, 2) : ozb;
  const double x0 = o.TT(0, 0), y0 = o.TT(0, 1), x1 = o.TT(1, 0), y1 = o.TT(1, 1), x2 = o.TT(2, 0), y2 = o.TT(2, 1);
  double best = x0 * dir[0] + y0 * dir[1] + za * dir[2];
  int code = 0;
  out[0] = x0; out[1] = y0; out[2] = za;
  double t = x1 * dir[0] + y1 * dir[1] + zb * dir[2];
  if (t > best) { best = t; code = 1; out[0] = x1; out[1] = y1; out[2] = zb; }
  t = x2 * dir[0] + y2 * dir[1] + zc * dir[2];
  if (t > best) { best = t; code = 2; out[0] = x2; out[1] = y2; out[2] = zc; }
  return code + (top ? 3 : 0);
}
template <typename PR>
__device__ __forceinline__ void prism_point(const PR& o, int code, double* out) {
  const int i = code >= 3 ? code - 3 : code;
  out[0] = o.TT(i, 0); out[1] = o.TT(i, 1); out[2] = code >= 3 ? o.TT(i, 2) : o.ZB();
}

// Recomputation of witnesses from their indices: the two shapes of a query
struct WitMesh {   // mesh vs mesh
  const double *sx, *sy, *sz; const SFrame *f1, *f2; Hull h1, h2;
  __device__ __forceinline__ void get(int i1, int i2, double* v1, double* v2) const {
    mesh_point(sx, sy, sz, h1, *f1, i1, v1);
    mesh_point(sx, sy, sz, h2, *f2, i2, v2);
  }
  __device__ __forceinline__ void centers(double* c1, double* c2) const {
    for (int k = 0; k < 3; k++) { c1[k] = f1->P(k); c2[k] = f2->P(k); }
  }
};

template <typename PR>
__device__ __forceinline__ void prism_center(const PR& o, double* c) {
  c[0] = c[1] = c[2] = 0;
#pragma unroll
  for (int i = 0; i < 3; i++) { c[0] += o.TT(i, 0); c[1] += o.TT(i, 1); c[2] += o.ZB(); }  // bottom vertices 0..2
#pragma unroll
  for (int i = 0; i < 3; i++) { c[0] += o.TT(i, 0); c[1] += o.TT(i, 1); c[2] += o.TT(i, 2); }  // top vertices 3..5
  c[0] *= 1.0 / 6.0; c[1] *= 1.0 / 6.0; c[2] *= 1.0 / 6.0;
}

struct WitPrism {   // height-field prism vs mesh
  const double *sx, *sy, *sz; const SPrism* p; const SFrame* f2; Hull h2;
  __device__ __forceinline__ void get(int i1, int i2, double* v1, double* v2) const {
    prism_point(*p, i1, v1);
    mesh_point(sx, sy, sz, h2, *f2, i2, v2);
  }
  __device__ __forceinline__ void centers(double* c1, double* c2) const {
    prism_center(*p, c1);
    for (int k = 0; k < 3; k++) c2[k] = f2->P(k);
  }
};

// ------------------------------------------------------------------ MPR pieces (libccd mpr.c)
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

__device__ __forceinline__ void expand_portal(const double* p0v, Supp& p1, Supp& p2, Supp& p3, const Supp& v4) {
  double v4v0[3];
  c_cross(v4v0, v4.v, p0v);
  double dot = c_dot(p1.v, v4v0);
  if (dot > 0) {
    dot = c_dot(p2.v, v4v0);
    if (dot > 0) cps(p1, v4); else cps(p3, v4);
  } else {
    dot = c_dot(p3.v, v4v0);
    if (dot > 0) cps(p2, v4); else cps(p1, v4);
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
    cp3(witness, x0);
  } else if (t > 1 || c_eq(t, 1.0)) {
    dist = c_dist2(b, P);
    cp3(witness, b);
  } else {
    cp3(witness, d);
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
    cp3(witness, x0);
    witness[0] += d1[0]; witness[1] += d1[1]; witness[2] += d1[2];
    witness[0] += d2[0]; witness[1] += d2[1]; witness[2] += d2[2];
    dist = c_dist2(witness, P);
  } else {
    dist = point_seg_dist2(P, x0, Bv, witness);
    dist2 = point_seg_dist2(P, x0, C, witness2);
    if (dist2 < dist) { dist = dist2; cp3(witness, witness2); }
    dist2 = point_seg_dist2(P, Bv, C, witness2);
    if (dist2 < dist) { dist = dist2; cp3(witness, witness2); }
  }
  return dist;
}

// p0v: the interior point of the Minkowski difference; its witnesses (the two centres) and those of the portal vertices
// are recomputed from W
template <typename W>
__device__ __forceinline__ void find_pos(const double* p0v, const W& wit, const Supp& p1, const Supp& p2, const Supp& p3,
                                         double* pos) {
  double dir[3], b0, b1, b2, b3, sum, inv, vec[3], q1[3] = {0, 0, 0}, q2[3] = {0, 0, 0};
  portal_dir(p1, p2, p3, dir);
  c_cross(vec, p1.v, p2.v); b0 = c_dot(vec, p3.v);
  c_cross(vec, p3.v, p2.v); b1 = c_dot(vec, p0v);
  c_cross(vec, p0v, p1.v); b2 = c_dot(vec, p3.v);
  c_cross(vec, p2.v, p1.v); b3 = c_dot(vec, p0v);
  sum = b0 + b1 + b2 + b3;
  if (c_is_zero(sum) || sum < 0) {
    b0 = 0;
    c_cross(vec, p2.v, p3.v); b1 = c_dot(vec, dir);
    c_cross(vec, p3.v, p1.v); b2 = c_dot(vec, dir);
    c_cross(vec, p1.v, p2.v); b3 = c_dot(vec, dir);
    sum = b1 + b2 + b3;
  }
  inv = 1.0 / sum;
  double a1[3], a2[3];
  wit.centers(a1, a2);
#pragma unroll
  for (int k = 0; k < 3; k++) { q1[k] += a1[k] * b0; q2[k] += a2[k] * b0; }
  wit.get(p1.i1, p1.i2, a1, a2);
#pragma unroll
  for (int k = 0; k < 3; k++) { q1[k] += a1[k] * b1; q2[k] += a2[k] * b1; }
  wit.get(p2.i1, p2.i2, a1, a2);
#pragma unroll
  for (int k = 0; k < 3; k++) { q1[k] += a1[k] * b2; q2[k] += a2[k] * b2; }
  wit.get(p3.i1, p3.i2, a1, a2);
#pragma unroll
  for (int k = 0; k < 3; k++) { q1[k] += a1[k] * b3; q2[k] += a2[k] * b3; }
  c_scale(q1, inv);
  c_scale(q2, inv);
  pos[0] = q1[0] + q2[0]; pos[1] = q1[1] + q2[1]; pos[2] = q1[2] + q2[2];
  c_scale(pos, 0.5);
}

// ------------------------------------------------------------------ MPR as a per-lane state machine
// ccdMPRPenetration split at its support calls.  `step` consumes the support point `s` computed for the
// current `dir` and returns 1 (continue: new dir set), 0 (penetration found: depth/pdir/pos set) or -1.
enum { ST_D1 = 0, ST_D2, ST_D3, ST_REFINE, ST_PENETR };

struct Mpr {
  Supp p1, p2, p3;
  double p0v[3];
  double dir[3];
  int st, guard, iterations;
  double depth, pdir[3], pos[3];
#ifdef HMJ_PHASE_PROF
  double sepgap;   // developer build: -dot(s.v, dir) at a no-contact exit that proves separation along dir (else 0)
#define MPR_SEP(d) (sepgap = -(d))
#else
#define MPR_SEP(d) ((void)0)
#endif

  // discoverPortal, up to the first support call
  __device__ __forceinline__ void begin(const double* center1, const double* center2) {
    c_sub2(p0v, center1, center2);
    if (c_vec_is_origin(p0v)) p0v[0] += CCD_EPS * 10.0;
    cp3(dir, p0v);
    c_scale(dir, -1.0);
    c_normalize(dir);
    st = ST_D1;
  }

  // head of refinePortal's loop: leaves st = ST_REFINE or ST_PENETR with dir set; -1 on the guard
  __device__ __forceinline__ int refine_head() {
    if (++guard > 1000) return -1;
    portal_dir(p1, p2, p3, dir);
    const double dot = c_dot(dir, p1.v);
    if (c_is_zero(dot) || dot > 0) {  // portal encapsules origin -> findPenetr (same dir)
      st = ST_PENETR;
      iterations = 0;
    } else {
      st = ST_REFINE;
    }
    return 1;
  }

  template <typename W>
  __device__ __forceinline__ int step(const Supp& s, double tol, int maxit, const W& wit) {
    double va[3], vb[3], dot;
    switch (st) {
      case ST_D1: {
        cps(p1, s);
        dot = c_dot(p1.v, dir);
        if (c_is_zero(dot) || dot < 0) { MPR_SEP(dot); return -1; }
        c_cross(dir, p0v, p1.v);
        if (c_is_zero(c_dot(dir, dir))) {
          double w1[3], w2[3];
          wit.get(p1.i1, p1.i2, w1, w2);
          if (c_vec_is_origin(p1.v)) {  // touching contact on v1
            depth = 0;
            pdir[0] = pdir[1] = pdir[2] = 0;
            for (int k = 0; k < 3; k++) pos[k] = (w1[k] + w2[k]) * 0.5;
          } else {  // origin on the v0-v1 segment
            for (int k = 0; k < 3; k++) pos[k] = (w1[k] + w2[k]) * 0.5;
            cp3(pdir, p1.v);
            depth = sqrt(c_dot(pdir, pdir));
            c_normalize(pdir);
          }
          return 0;
        }
        c_normalize(dir);
        st = ST_D2;
        return 1;
      }
      case ST_D2: {
        cps(p2, s);
        dot = c_dot(p2.v, dir);
        if (c_is_zero(dot) || dot < 0) { MPR_SEP(dot); return -1; }
        c_sub2(va, p1.v, p0v);
        c_sub2(vb, p2.v, p0v);
        c_cross(dir, va, vb);
        c_normalize(dir);
        dot = c_dot(dir, p0v);
        if (dot > 0) {
          Supp t;
          cps(t, p1); cps(p1, p2); cps(p2, t);
          c_scale(dir, -1.0);
        }
        guard = 1;  // first trip of the while(!have4) loop
        st = ST_D3;
        return 1;
      }
      case ST_D3: {
        cps(p3, s);
        dot = c_dot(p3.v, dir);
        if (c_is_zero(dot) || dot < 0) { MPR_SEP(dot); return -1; }
        bool cont = false;
        c_cross(va, p1.v, p3.v);
        dot = c_dot(va, p0v);
        if (dot < 0 && !c_is_zero(dot)) { cps(p2, p3); cont = true; }
        if (!cont) {
          c_cross(va, p3.v, p2.v);
          dot = c_dot(va, p0v);
          if (dot < 0 && !c_is_zero(dot)) { cps(p1, p3); cont = true; }
        }
        if (cont) {
          c_sub2(va, p1.v, p0v);
          c_sub2(vb, p2.v, p0v);
          c_cross(dir, va, vb);
          c_normalize(dir);
          if (++guard > 1000) return -1;
          return 1;
        }
        guard = 0;
        return refine_head();
      }
      default: {  // ST_REFINE / ST_PENETR: one body for both (refinePortal's and findPenetr's loops run the same portal
                  // test, expansion and new direction; lanes of a warp are in either state, so a shared body halves the
                  // divergent work of a trip).  Per lane this is the same operation sequence as the two separate loops.
        const bool ref = st == ST_REFINE;
        if (ref) {
          dot = c_dot(s.v, dir);
          if (!(c_is_zero(dot) || dot > 0)) { MPR_SEP(dot); return -1; }
        }
        const bool reach = portal_reach_tolerance(p1, p2, p3, s, dir, tol);
        if (ref) {
          if (reach) return -1;
        } else if (reach || iterations > maxit) {
          return 2;   // penetration found: depth / direction / position are computed by finalize(), once per warp
        }
        expand_portal(p0v, p1, p2, p3, s);
        if (ref) {
          if (++guard > 1000) return -1;
        } else {
          iterations++;
        }
        portal_dir(p1, p2, p3, dir);
        if (ref) {
          dot = c_dot(dir, p1.v);
          if (c_is_zero(dot) || dot > 0) {  // portal encapsules origin -> findPenetr (same dir)
            st = ST_PENETR;
            iterations = 0;
          }
        }
        return 1;
      }
    }
  }

  // tail of findPenetr (depth, penetration direction, contact position), deferred so that all queries of a warp that
  // end in a penetration run it together after the trip loop instead of one after the other at their own last trips
  template <typename W>
  __device__ __forceinline__ void finalize(const W& wit) {
    const double origin[3] = {0, 0, 0};
    depth = point_tri_dist2(origin, p1.v, p2.v, p3.v, pdir);
    depth = sqrt(depth);
    if (c_is_zero(depth)) {
      pdir[0] = pdir[1] = pdir[2] = 0;
    } else {
      c_normalize(pdir);
    }
    find_pos(p0v, wit, p1, p2, p3, pos);
  }
};

// ------------------------------------------------------------------ height-field strip walk
// mjc_ConvexHField walks, for every grid row r in [rmin, rmax), the strip of vertices
// (c, r+1), (c, r), (c+1, r+1), (c+1, r), ... ; every three consecutive strip vertices are the top of a prism.
__device__ __forceinline__ void strip_vertex(double* v, int r, int cmin, int s, const float* __restrict__ data,
                                             const double* size1, int ncol, double dx, double dy) {
  const int c = cmin + (s >> 1), i = s & 1;
  const int rr = r + (i == 0 ? 1 : 0);
  v[0] = dx * c - size1[0];
  v[1] = dy * rr - size1[1];
  v[2] = (double)data[rr * ncol + c] * size1[2] + 0.0;
}

struct PrismWalk {
  int r, s, nvert, rmax, cmin, smax;  // smax = strip vertices per row
  int pr, ps;                          // row / strip index of the newest vertex of the prism just produced
  // advance until a prism that must be tested is in `p`; false when exhausted
  __device__ __forceinline__ bool next(Prism& p, const float* __restrict__ data, const double* size1, int ncol, double dx,
                                       double dy, double zmin) {
    while (r < rmax) {
      cp3(p.t[0], p.t[1]);
      cp3(p.t[1], p.t[2]);
      strip_vertex(p.t[2], r, cmin, s, data, size1, ncol, dx, dy);
      nvert++;
      const bool test = nvert > 2 && !(p.t[0][2] < zmin && p.t[1][2] < zmin && p.t[2][2] < zmin);
      pr = r; ps = s;
      if (++s == smax) { s = 0; r++; nvert = 0; }
      if (test) return true;
    }
    return false;
  }
};

// mesh g2 expressed in the frame of the height field (geom 0)
__device__ __forceinline__ void hfield_rel_frame(Frame& f2, const double* gx, const double* gm, int g2) {
  const double* pos1 = gx;
  const double* mat1 = gm;
  const double vec[3] = {gx[g2 * 3] - pos1[0], gx[g2 * 3 + 1] - pos1[1], gx[g2 * 3 + 2] - pos1[2]};
  f2.pos[0] = mat1[0] * vec[0] + mat1[3] * vec[1] + mat1[6] * vec[2];
  f2.pos[1] = mat1[1] * vec[0] + mat1[4] * vec[1] + mat1[7] * vec[2];
  f2.pos[2] = mat1[2] * vec[0] + mat1[5] * vec[1] + mat1[8] * vec[2];
  const double* mat2 = gm + g2 * 9;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      f2.mat[3 * i + j] = mat1[i] * mat2[j] + mat1[3 + i] * mat2[3 + j] + mat1[6 + i] * mat2[6 + j];
}

// task word: pair k (6 bits) | slot ordinal (3) | strip row (3) | cmin (3) | strip index of the newest vertex (5)
__device__ __forceinline__ int pack_task(int k, int ord, int r, int cmin, int s) {
  return k | (ord << 6) | (r << 9) | (cmin << 12) | (s << 15);
}

// ---- hull vertices into shared memory with the bulk-copy engine (TMA, 1-D): one thread programs the copies against an
// mbarrier and the CTA goes on with work that does not need the vertices (frames, bounding tests, fetching its first work
// item); every thread waits on the barrier right before its first support call.  Sources and sizes are multiples of 16
// bytes: every mesh starts at a multiple of 4 vertices and is padded to a multiple of 4 (hmj_api.cu: build_tables).
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(bar)));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst)),
               "l"(src), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "HMJ_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra HMJ_DONE;\n"
      "bra HMJ_WAIT;\n"
      "HMJ_DONE:\n"
      "}\n" ::"r"(smem_addr(bar)),
      "r"(phase)
      : "memory");
}
// vertices [first, first + count) of the fp64 tables -> sx / sy / sz at offset 0; call from ONE thread after mbar_init
__device__ __forceinline__ void stage_hulls_async(const hmj_tables_t* __restrict__ T, double* sx, double* sy, double* sz, int first,
                                                  int count, uint64_t* bar) {
  const uint32_t bytes = (uint32_t)count * 8u;
  mbar_expect(bar, 3u * bytes);
  bulk_load(sx, T->dvx + first, bytes, bar);
  bulk_load(sy, T->dvy + first, bytes, bar);
  bulk_load(sz, T->dvz + first, bytes, bar);
}

// warp-aggregated append of (env, word) to task group g for the lanes with `hit`
__device__ __forceinline__ void push_tasks(const hmj_tables_t* __restrict__ T, const hmj_buffers_t& B, int g, bool hit,
                                           int env, int word, int lane) {
  const unsigned bal = __ballot_sync(0xffffffffu, hit);
  if (!bal) return;
  int base = 0;
  if (lane == __ffs(bal) - 1) base = atomicAdd(&B.wq_count[g], __popc(bal));
  base = __shfl_sync(0xffffffffu, base, __ffs(bal) - 1);
  if (hit) B.wq_task[(size_t)T->group_base[g] * B.n + base + __popc(bal & ((1u << lane) - 1u))] = make_int2(env, word);
}

// Mesh-mesh candidate pair k of one env (lane <-> env, the pair is warp-uniform): bounding spheres, then a conservative
// pre-screen of MPR's first exit; survivors become one narrowphase task.
__device__ __forceinline__ void broad_mesh_mesh(const hmj_tables_t* __restrict__ T, const hmj_buffers_t& B, int k, int env,
                                                bool in, int lane) {
  const hmj_model_t& M = T->m;
  const int g1 = M.pair_geom1[k], g2 = M.pair_geom2[k];
  const double* gx = B.geom_xpos + (size_t)env * HMJ_NGEOM * 3;
  const double dif0 = gx[g1 * 3] - gx[g2 * 3], dif1 = gx[g1 * 3 + 1] - gx[g2 * 3 + 1], dif2 = gx[g1 * 3 + 2] - gx[g2 * 3 + 2];
  const double bound = M.geom_rbound[g1] + M.geom_rbound[g2];
  bool hit = in && !(dif0 * dif0 + dif1 * dif1 + dif2 * dif2 > bound * bound);
  if (hit) {
    // Conservative pre-screen of ccdMPRPenetration's first exit (dot(support(dir), dir) <= 0 along the
    // centre line): the hulls' local bounding boxes bound both supports, so a box gap larger than the
    // rounding noise proves the exact test fails too.  Never rejects a query MPR would continue.
    const double* gm = B.geom_xmat + (size_t)env * HMJ_NGEOM * 9;
    const double n2 = dif0 * dif0 + dif1 * dif1 + dif2 * dif2;
    if (n2 > 1e-12) {
      const double inv = rsqrt(n2);
      const double d[3] = {-dif0 * inv, -dif1 * inv, -dif2 * inv};  // from geom1's centre towards geom2's
      const double* m1 = gm + g1 * 9;
      const double* m2 = gm + g2 * 9;
      const double* lo1 = T->mesh_lo[M.geom_meshid[g1]];
      const double* hi1 = T->mesh_hi[M.geom_meshid[g1]];
      const double* lo2 = T->mesh_lo[M.geom_meshid[g2]];
      const double* hi2 = T->mesh_hi[M.geom_meshid[g2]];
      double max1 = d[0] * gx[g1 * 3] + d[1] * gx[g1 * 3 + 1] + d[2] * gx[g1 * 3 + 2];
      double min2 = d[0] * gx[g2 * 3] + d[1] * gx[g2 * 3 + 1] + d[2] * gx[g2 * 3 + 2];
#pragma unroll
      for (int q = 0; q < 3; q++) {
        const double l1 = m1[q] * d[0] + m1[3 + q] * d[1] + m1[6 + q] * d[2];
        const double l2 = m2[q] * d[0] + m2[3 + q] * d[1] + m2[6 + q] * d[2];
        max1 += fmax(l1 * lo1[q], l1 * hi1[q]);
        min2 += fmin(l2 * lo2[q], l2 * hi2[q]);
      }
      if (max1 - min2 < -1e-9) hit = false;
    }
  }
  push_tasks(T, B, T->pair_group[k], hit, env, pack_task(k, 0, 0, 0, 0), lane);
}

// ------------------------------------------------------------------ broadphase + midphase
// Thread t = k * npad + e (npad = N rounded up to 32, so the pair k is warp-uniform).
//  * mesh-mesh pair: bounding spheres; survivors become one MPR task.
//  * hfield-mesh pair (head of mjc_ConvexHField): box test, hull AABB in the hfield frame, strip walk; every
//    prism that must be tested gets its slot ordinal and is pre-screened with the FIRST support call of
//    ccdMPRPenetration's discoverPortal (64 % of all queries end there: dot(v1, dir) <= 0 -> no contact);
//    survivors become MPR tasks, which restart the query from the top in k_narrow (same arithmetic).
// The per-env slot bitmask is zeroed before this kernel; k_narrow sets the bits of the slots it fills.
__global__ void __launch_bounds__(BROAD_THREADS)
k_broad(const hmj_tables_t* __restrict__ T, hmj_buffers_t B, int npad, int kpos0) {
  __shared__ __align__(16) double sx[HMJ_NVERT_SMEM];
  __shared__ __align__(16) double sy[HMJ_NVERT_SMEM];
  __shared__ __align__(16) double sz[HMJ_NVERT_SMEM];
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int kpos = kpos0 + t / npad;  // warp-uniform; whole CTAs can be past the last pair
  const int e = t - (t / npad) * npad;
  const int lane = threadIdx.x & 31;
  const hmj_model_t& M = T->m;
  if (kpos >= HMJ_NPAIR) return;  // npad * NPAIR is a multiple of 32: whole warps leave together (before any barrier use)
  const int k = T->broad_order[kpos];   // expensive pairs first (hmj_api.cu: build of the tables)
  const bool in = e < B.n && (!B.active || B.active[e]);
  const int env = e < B.n ? e : B.n - 1;
  const int g1 = M.pair_geom1[k], g2 = M.pair_geom2[k];
  const double* gx = B.geom_xpos + (size_t)env * HMJ_NGEOM * 3;
  const double* gm = B.geom_xmat + (size_t)env * HMJ_NGEOM * 9;
  const int nslot = T->pair_slot[k + 1] - T->pair_slot[k];

  if (g1 != 0) {
    broad_mesh_mesh(T, B, k, env, in, lane);
    return;
  }
  // ---- hfield pair.  NOTE: k_broad is launched with hfield pairs in their own CTAs (npad is a multiple of
  // BROAD_THREADS), so the barrier below is reached by whole CTAs only.
  // this CTA works on ONE pair, i.e. one hull: only that hull is staged (at offset 0)
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) mbar_init(&bar);
  __syncthreads();
  if (threadIdx.x == 0) stage_hulls_async(T, sx, sy, sz, T->mesh_adr[M.geom_meshid[g2]], T->mesh_npad[M.geom_meshid[g2]], &bar);
  const double* size1 = M.hfield_size;
  const int nrow = M.hfield_nrow, ncol = M.hfield_ncol;
  const float* data = B.params[env].hfield_data;
  const Hull h2 = {0, T->mesh_npad[M.geom_meshid[g2]]};
  const SupTab tab2 = sup_tab(T, M.geom_meshid[g2]);
  Frame f2;
  hfield_rel_frame(f2, gx, gm, g2);
  bool active = in;
  {
    const double r2 = M.geom_rbound[g2];
    for (int i = 0; i < 2; i++)
      if ((size1[i] < f2.pos[i] - r2 - 0.0) || (-size1[i] > f2.pos[i] + r2 + 0.0)) active = false;
    if (size1[2] < f2.pos[2] - r2 - 0.0) active = false;
    if (-size1[3] > f2.pos[2] + r2 + 0.0) active = false;
  }
  mbar_wait(&bar, 0);   // the hull is in shared memory (also before leaving: no CTA exits with a copy in flight)
  if (!__any_sync(0xffffffffu, active)) return;
  double lo[3], hi[3];
  mesh_aabb(sx, sy, sz, h2, tab2, f2, lo, hi);
  if ((lo[0] > size1[0]) || (hi[0] < -size1[0]) || (lo[1] > size1[1]) || (hi[1] < -size1[1]) || (lo[2] > size1[2]) ||
      (hi[2] < -size1[3]))
    active = false;
  PrismWalk w;
  {
    int cmin = (int)floor((lo[0] + size1[0]) / (2 * size1[0]) * (ncol - 1));
    int cmax = (int)ceil((hi[0] + size1[0]) / (2 * size1[0]) * (ncol - 1));
    int rmin = (int)floor((lo[1] + size1[1]) / (2 * size1[1]) * (nrow - 1));
    int rmax = (int)ceil((hi[1] + size1[1]) / (2 * size1[1]) * (nrow - 1));
    cmin = max(0, cmin); cmax = min(ncol - 1, cmax);
    w.cmin = cmin;
    w.smax = 2 * (cmax - cmin + 1);
    w.r = max(0, rmin); w.rmax = min(nrow - 1, rmax);
    w.s = 0; w.nvert = 0; w.pr = w.ps = 0;
    if (cmin > cmax) active = false;  // empty column range: the strip loops never produce a prism
  }
  const double zmin = lo[2];
  const double dx = (2.0 * size1[0]) / (ncol - 1);
  const double dy = (2.0 * size1[1]) / (nrow - 1);
  Prism p;
  for (int a = 0; a < 3; a++) p.t[a][0] = p.t[a][1] = p.t[a][2] = 0.0;
  p.zb = -size1[3];
  const int g = T->pair_group[k];
  int ord = 0;
  while (true) {
    const bool has = active && w.next(p, data, size1, ncol, dx, dy, zmin);
    if (!__any_sync(0xffffffffu, has)) break;
    active = has;
    // discoverPortal up to its first exit test
    double c1[3], p0v[3], dir[3], v1[3], v2[3], v[3], nd[3];
    prism_center(p, c1);
    c_sub2(p0v, c1, f2.pos);
    if (c_vec_is_origin(p0v)) p0v[0] += CCD_EPS * 10.0;
    cp3(dir, p0v);
    c_scale(dir, -1.0);
    c_normalize(dir);
    prism_support(p, dir, v1);
    cp3(nd, dir);
    c_scale(nd, -1.0);
    mesh_support(sx, sy, sz, h2, tab2, f2, nd, v2);
    c_sub2(v, v1, v2);
    const double dot = c_dot(v, dir);
    bool survive = has && !(c_is_zero(dot) || dot < 0);
    if (has && ord >= nslot) {  // more tested prisms than slots: cannot happen for this model's geometry (<= 8)
      survive = false;
      active = false;
      atomicOr(&B.status[env], HMJ_STATUS_CONTACTFULL);
    }
    push_tasks(T, B, g, survive, env, pack_task(k, ord, w.pr, w.cmin, w.ps), lane);
    if (has) ord++;
  }
}

// ------------------------------------------------------------------ broadphase of the height-field pairs, small batches
// The same work as the height-field branch of k_broad with ONE WARP per (pair, env) instead of one thread: lane p owns
// prism p of the strip walk (k_broad's thread walks them one after the other: pure latency when the batch is too small
// to hide it, which is the planner's case); its slot ordinal is its rank among the tested prisms, so tasks and slots are
// exactly k_broad's.  The bounding box is computed by every lane alike (a handful of candidates, broadcast loads), the
// first support call of a prism by its lane.  Vertices are read from the fp64 tables in global memory (25 KB,
// L1-resident); 4 warps per CTA.
#define BROAD_COOP_MAX_N 512    /* launcher uses this kernel up to this batch size (measured with the candidate lists: +4 % at 10-64 samples, even at 256) */
__global__ void __launch_bounds__(128)
k_broad_hf_coop(const hmj_tables_t* __restrict__ T, hmj_buffers_t B, int nhf) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int wid = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (wid >= nhf * B.n) {   // trailing warps: the mesh-mesh pairs, lane <-> env as in k_broad (one launch for the whole stage)
    const int w2 = wid - nhf * B.n, per = (B.n + 31) >> 5;
    const int kp = nhf + w2 / per;
    if (kp >= HMJ_NPAIR) return;
    const int e = (w2 - (w2 / per) * per) * 32 + lane;
    const bool in = e < B.n && (!B.active || B.active[e]);
    broad_mesh_mesh(T, B, T->broad_order[kp], e < B.n ? e : B.n - 1, in, lane);
    return;
  }
  const int kpos = wid / B.n, env = wid - kpos * B.n;
  if (B.active && !B.active[env]) return;
  const hmj_model_t& M = T->m;
  const int k = T->broad_order[kpos];
  const int g2 = M.pair_geom2[k];
  const double* gx = B.geom_xpos + (size_t)env * HMJ_NGEOM * 3;
  const double* gm = B.geom_xmat + (size_t)env * HMJ_NGEOM * 9;
  const int nslot = T->pair_slot[k + 1] - T->pair_slot[k];
  const double* sx = T->dvx;
  const double* sy = T->dvy;
  const double* sz = T->dvz;
  const double* size1 = M.hfield_size;
  const int nrow = M.hfield_nrow, ncol = M.hfield_ncol;
  const float* data = B.params[env].hfield_data;
  const Hull h2 = {T->mesh_adr[M.geom_meshid[g2]], T->mesh_npad[M.geom_meshid[g2]]};
  Frame f2;
  hfield_rel_frame(f2, gx, gm, g2);
  {
    const double r2 = M.geom_rbound[g2];
    bool active = true;
    for (int i = 0; i < 2; i++)
      if ((size1[i] < f2.pos[i] - r2 - 0.0) || (-size1[i] > f2.pos[i] + r2 + 0.0)) active = false;
    if (size1[2] < f2.pos[2] - r2 - 0.0) active = false;
    if (-size1[3] > f2.pos[2] + r2 + 0.0) active = false;
    if (!active) return;   // warp-uniform
  }
  // ---- bounding box of the hull in the height field's frame (every lane the same few candidates: broadcast loads)
  const SupTab tab2 = sup_tab(T, M.geom_meshid[g2]);
  double lo[3], hi[3];
  mesh_aabb(sx, sy, sz, h2, tab2, f2, lo, hi);
  if ((lo[0] > size1[0]) || (hi[0] < -size1[0]) || (lo[1] > size1[1]) || (hi[1] < -size1[1]) || (lo[2] > size1[2]) ||
      (hi[2] < -size1[3]))
    return;
  int cmin = (int)floor((lo[0] + size1[0]) / (2 * size1[0]) * (ncol - 1));
  int cmax = (int)ceil((hi[0] + size1[0]) / (2 * size1[0]) * (ncol - 1));
  int rmin = (int)floor((lo[1] + size1[1]) / (2 * size1[1]) * (nrow - 1));
  int rmax = (int)ceil((hi[1] + size1[1]) / (2 * size1[1]) * (nrow - 1));
  cmin = max(0, cmin); cmax = min(ncol - 1, cmax);
  rmin = max(0, rmin); rmax = min(nrow - 1, rmax);
  if (cmin > cmax) return;
  const int smax = 2 * (cmax - cmin + 1);       // strip vertices per row
  const int per_row = smax - 2;                  // prisms per row (every three consecutive strip vertices)
  const int nprism = (rmax > rmin ? rmax - rmin : 0) * per_row;
  const double zmin = lo[2];
  const double dx = (2.0 * size1[0]) / (ncol - 1);
  const double dy = (2.0 * size1[1]) / (nrow - 1);
  const int g = T->pair_group[k];
  int ord_base = 0;
  for (int p0 = 0; p0 < nprism; p0 += 32) {
    const int pi = p0 + lane;
    const bool exists = pi < nprism;
    const int pr = rmin + (exists ? pi / per_row : 0);
    const int ps = 2 + (exists ? pi % per_row : 0);        // strip index of the prism's newest vertex
    Prism p;
    p.zb = -size1[3];
#pragma unroll
    for (int a = 0; a < 3; a++) strip_vertex(p.t[a], pr, cmin, ps - 2 + a, data, size1, ncol, dx, dy);
    const bool test = exists && !(p.t[0][2] < zmin && p.t[1][2] < zmin && p.t[2][2] < zmin);
    const unsigned tm = __ballot_sync(FULL, test);
    const int ord = ord_base + __popc(tm & ((1u << lane) - 1u));
    // discoverPortal up to its first exit test (as in k_broad)
    double c1[3], p0v[3], dir[3], v1[3], v2[3] = {0, 0, 0}, v[3], nd[3];
    prism_center(p, c1);
    c_sub2(p0v, c1, f2.pos);
    if (c_vec_is_origin(p0v)) p0v[0] += CCD_EPS * 10.0;
    cp3(dir, p0v);
    c_scale(dir, -1.0);
    c_normalize(dir);
    prism_support(p, dir, v1);
    cp3(nd, dir);
    c_scale(nd, -1.0);
    if (test) mesh_support(sx, sy, sz, h2, tab2, f2, nd, v2);
    c_sub2(v, v1, v2);
    const double dot = c_dot(v, dir);
    bool survive = test && !(c_is_zero(dot) || dot < 0);
    const bool over = test && ord >= nslot;   // more tested prisms than slots: k_broad drops this one and stops the walk
    const unsigned om = __ballot_sync(FULL, over);
    if (om) {
      if (lane == 0) atomicOr(&B.status[env], HMJ_STATUS_CONTACTFULL);
      survive = survive && !over;
    }
    push_tasks(T, B, g, survive, env, pack_task(k, ord, pr, cmin, ps), lane);
    if (om) return;
    ord_base += __popc(tm);
  }
}

#ifdef HMJ_PHASE_PROF
// developer build: per-group item statistics of k_narrow {items, sum cycles, max cycles, sum trips, max trips} and the
// per-warp busy time {max cycles, sum cycles, warps, staging cycles}
__device__ unsigned long long hmj_narrow_prof[64][5];
__device__ unsigned long long hmj_narrow_warp[12];   // [4 + shift]: launches that ran with item size 1 << shift; [10]: sum of total items, [11]: sum of queries
__device__ unsigned long long hmj_narrow_q[12];   // queries: [0..2] contact {n, trips, max trips}, [3..5] no contact, [6..9] no-contact exits with a separation gap > 1e-9 / 1e-6 / 1e-4 / 1e-3
__device__ int hmj_narrow_trips;   // scratch written by the narrow_* routines (lane 0 of the running warp reads its own)
extern "C" int hmj_debug_narrow_prof(unsigned long long* out, unsigned long long* warp, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, hmj_narrow_prof, sizeof(unsigned long long) * 64 * 5);
  cudaMemcpyFromSymbol(warp, hmj_narrow_warp, sizeof(unsigned long long) * 12);
  if (reset) {
    void* p = nullptr;
    cudaGetSymbolAddress(&p, hmj_narrow_prof); cudaMemset(p, 0, sizeof(unsigned long long) * 64 * 5);
    cudaGetSymbolAddress(&p, hmj_narrow_warp); cudaMemset(p, 0, sizeof(unsigned long long) * 12);
  }
  return 0;
}
extern "C" int hmj_debug_narrow_q(unsigned long long* out, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, hmj_narrow_q, sizeof(unsigned long long) * 12);
  if (reset) { void* p = nullptr; cudaGetSymbolAddress(&p, hmj_narrow_q); cudaMemset(p, 0, sizeof(unsigned long long) * 12); }
  return 0;
}
#define NP_QUERY(valid, mt, found, gap) do { if (valid) { const int o_ = (found) ? 0 : 3; atomicAdd(&hmj_narrow_q[o_], 1ull); \
  atomicAdd(&hmj_narrow_q[o_ + 1], (unsigned long long)(mt)); atomicMax(&hmj_narrow_q[o_ + 2], (unsigned long long)(mt)); \
  if (!(found)) { if ((gap) > 1e-9) atomicAdd(&hmj_narrow_q[6], 1ull); if ((gap) > 1e-6) atomicAdd(&hmj_narrow_q[7], 1ull); \
  if ((gap) > 1e-4) atomicAdd(&hmj_narrow_q[8], 1ull); if ((gap) > 1e-3) atomicAdd(&hmj_narrow_q[9], 1ull); } } } while (0)
#define NP_TRIP() do { trips_++; } while (0)
#define NP_MINE() do { if (active) my_trips_++; } while (0)
#else
#define NP_TRIP() do { } while (0)
#define NP_MINE() do { } while (0)
#define NP_QUERY(valid, mt, found, gap) do { } while (0)
#endif
// ------------------------------------------------------------------ narrowphase
__device__ __forceinline__ void write_slot(const hmj_buffers_t& B, int env, int s, double dist, const double* nrm, const double* pos) {
  double* slot = B.slots + ((size_t)env * HMJ_NSLOT + s) * HMJ_SLOT_DOUBLES;
  slot[0] = dist;
  slot[1] = pos[0]; slot[2] = pos[1]; slot[3] = pos[2];
  slot[4] = nrm[0]; slot[5] = nrm[1]; slot[6] = nrm[2];
  atomicOr(&B.slot_mask[(size_t)env * 4 + (s >> 5)], 1u << (s & 31));
}

// 32 mesh-mesh tasks of one pair (mjc_Convex, multiccd off): lane <-> environment.
__device__ __forceinline__ void narrow_mesh_mesh(const hmj_tables_t* __restrict__ T, const hmj_buffers_t& B,
                                                 const double* sx, const double* sy, const double* sz, double* sfr, int2 task,
                                                 bool active, int lane, int& trips_) {
  const hmj_model_t& M = T->m;
  const int env = task.x, k = task.y & 63;
  const int g1 = M.pair_geom1[k], g2 = M.pair_geom2[k];
  const Hull h1 = {T->mesh_adr[M.geom_meshid[g1]], T->mesh_npad[M.geom_meshid[g1]]};
  const Hull h2 = {T->mesh_adr[M.geom_meshid[g2]], T->mesh_npad[M.geom_meshid[g2]]};
  const double* gx = B.geom_xpos + (size_t)env * HMJ_NGEOM * 3;
  const double* gm = B.geom_xmat + (size_t)env * HMJ_NGEOM * 9;
  SFrame f1 = {sfr}, f2 = {sfr + 12 * NARROW_THREADS};   // this lane's two frames, in shared memory
  for (int q = 0; q < 3; q++) { f1.setP(q, gx[g1 * 3 + q]); f2.setP(q, gx[g2 * 3 + q]); }
  for (int q = 0; q < 9; q++) { f1.setM(q, gm[g1 * 9 + q]); f2.setM(q, gm[g2 * 9 + q]); }
  const double tol = M.mpr_tolerance;
  const int maxit = M.mpr_iterations;
  Mpr m;
  {
    const double c1[3] = {f1.P(0), f1.P(1), f1.P(2)}, c2[3] = {f2.P(0), f2.P(1), f2.P(2)};
    m.begin(c1, c2);
  }
  const WitMesh wit = {sx, sy, sz, &f1, &f2, h1, h2};
  bool found = false, fin = false;
  const SupTab tab1 = sup_tab(T, M.geom_meshid[g1]), tab2 = sup_tab(T, M.geom_meshid[g2]);
#ifdef HMJ_PHASE_PROF
  const bool valid_ = active; int my_trips_ = 0; m.sepgap = 0;
#endif
  while (true) {
    const unsigned am = __ballot_sync(0xffffffffu, active);
    if (!am) break;
    NP_TRIP(); NP_MINE();
    Supp s;
    double nd[3] = {m.dir[0], m.dir[1], m.dir[2]};
    c_scale(nd, -1.0);
    double v1[3], v2[3];
    if (active) {
      mesh_support(sx, sy, sz, h1, tab1, f1, m.dir, v1, &s.i1);
      mesh_support(sx, sy, sz, h2, tab2, f2, nd, v2, &s.i2);
      c_sub2(s.v, v1, v2);
      const int res = m.step(s, tol, maxit, wit);
      if (res != 1) {
        active = false;
        found = res == 0;
        fin = res == 2;
      }
    }
  }
  if (fin) { m.finalize(wit); found = true; }
  NP_QUERY(valid_, my_trips_, found, m.sepgap);
  if (found && !c_vec_is_origin(m.pdir)) write_slot(B, env, T->pair_slot[k], 0.0 - m.depth, m.pdir, m.pos);
}

// 32 prism-vs-hull tasks of one hull (pairs may differ per lane): lane <-> one pre-screened prism of
// mjc_ConvexHField; the query restarts from the top (same arithmetic as the screen in k_broad).
__device__ __forceinline__ void narrow_hfield(const hmj_tables_t* __restrict__ T, const hmj_buffers_t& B, const double* sx,
                                              const double* sy, const double* sz, double* sfr, Hull h2, int2 task, bool active, int lane,
                                              int& trips_) {
  const hmj_model_t& M = T->m;
  const int env = task.x;
  const int k = task.y & 63, ord = (task.y >> 6) & 7, r = (task.y >> 9) & 7, cmin = (task.y >> 12) & 7, ps = (task.y >> 15) & 31;
  const int g2 = M.pair_geom2[k];
  const double* gx = B.geom_xpos + (size_t)env * HMJ_NGEOM * 3;
  const double* gm = B.geom_xmat + (size_t)env * HMJ_NGEOM * 9;
  const double* size1 = M.hfield_size;
  const int nrow = M.hfield_nrow, ncol = M.hfield_ncol;
  const float* data = B.params[env].hfield_data;
  const double tol = M.mpr_tolerance;
  const int maxit = M.mpr_iterations;
  SFrame f2 = {sfr};   // this lane's frame of the mesh in the height field's frame, in shared memory
  double c2[3];
  {
    Frame fr;
    hfield_rel_frame(fr, gx, gm, g2);
    for (int q = 0; q < 3; q++) { f2.setP(q, fr.pos[q]); c2[q] = fr.pos[q]; }
    for (int q = 0; q < 9; q++) f2.setM(q, fr.mat[q]);
  }
  const double dx = (2.0 * size1[0]) / (ncol - 1);
  const double dy = (2.0 * size1[1]) / (nrow - 1);
  SPrism p = {sfr + 12 * NARROW_THREADS};   // this lane's prism, in shared memory behind its frame
  {
    Prism pr;
    pr.zb = -size1[3];
#pragma unroll
    for (int a = 0; a < 3; a++) strip_vertex(pr.t[a], r, cmin, max(ps - 2 + a, 0), data, size1, ncol, dx, dy);
#pragma unroll
    for (int e = 0; e < 9; e++) p.b[e * NARROW_THREADS] = pr.t[e / 3][e % 3];
    p.b[9 * NARROW_THREADS] = pr.zb;
  }
  Mpr m;
  {
    double c1[3];
    prism_center(p, c1);
    m.begin(c1, c2);
  }
  const WitPrism wit = {sx, sy, sz, &p, &f2, h2};
  bool found = false, fin = false;
  const SupTab tab2 = sup_tab(T, M.geom_meshid[g2]);
#ifdef HMJ_PHASE_PROF
  const bool valid_ = active; int my_trips_ = 0; m.sepgap = 0;
#endif
  while (true) {
    const unsigned am = __ballot_sync(0xffffffffu, active);
    if (!am) break;
    NP_TRIP(); NP_MINE();
    Supp s;
    double nd[3] = {m.dir[0], m.dir[1], m.dir[2]};
    double v1[3], v2[3];
    s.i1 = prism_support(p, m.dir, v1);
    c_scale(nd, -1.0);
    if (active) {
      mesh_support(sx, sy, sz, h2, tab2, f2, nd, v2, &s.i2);
      c_sub2(s.v, v1, v2);
      const int res = m.step(s, tol, maxit, wit);
      if (res != 1) {
        active = false;
        found = res == 0;
        fin = res == 2;
      }
    }
  }
  if (fin) { m.finalize(wit); found = true; }
  NP_QUERY(valid_, my_trips_, found, m.sepgap);
  if (found && !d_is_bad(m.depth)) {
    const double* pos1 = gx;
    const double* mat1 = gm;
    double nrm[3], cp[3];
    d_rot_vec_mat(nrm, m.pdir, mat1);
    d_rot_vec_mat(cp, m.pos, mat1);
    cp[0] += pos1[0]; cp[1] += pos1[1]; cp[2] += pos1[2];
    if (!(nrm[0] == 0 && nrm[1] == 0 && nrm[2] == 0)) {
      write_slot(B, env, T->pair_slot[k] + ord, 0.0 - m.depth, nrm, cp);
    }
  }
}

// Persistent warps over the task groups (7 hfield groups keyed by hull, 28 mesh-mesh groups keyed by pair);
// a work item is up to 32 consecutive tasks of one group, so the hull(s) a warp scans are uniform.
__global__ void __launch_bounds__(NARROW_THREADS, NARROW_MIN_CTAS)
k_narrow(const hmj_tables_t* __restrict__ T, hmj_buffers_t B) {
  // dynamic shared memory: hull vertices (fp64 SoA) | per-lane query frames (SFrame: [element][thread])
  extern __shared__ __align__(16) double narrow_smem[];
  double* sx = narrow_smem;
  double* sy = sx + HMJ_NVERT_SMEM;
  double* sz = sy + HMJ_NVERT_SMEM;
  double* sfr = sz + HMJ_NVERT_SMEM;
  const int lane = threadIdx.x & 31;
  // work items: group g contributes ceil(count[g] / isz); groups are numbered across TWO registers per lane.
  // The item size isz (tasks per warp) is the smallest of 4 / 8 / 16 / 32 that still gives every item its own warp:
  // a step has a few thousand to tens of thousands of queries against 1184 resident warps, and a trip costs a warp the
  // more the more of its lanes are in different MPR phases (3-6 k cycles at 16 lanes), so short items cut the
  // query-chain latency that bounds this kernel.  Every query is independent: the item size never changes a result bit.
  // Lane l holds schedule positions l and l + 32; position p is group T->group_order[p] (most expensive first).
  const int ga = T->group_order[lane];
  const int gb = (lane + 32 < HMJ_NGROUP) ? T->group_order[lane + 32] : 0;
  const int c0 = B.wq_count[ga];
  const int c1 = (lane + 32 < HMJ_NGROUP) ? B.wq_count[gb] : 0;
  const int cap0 = T->group_shiftcap[ga], cap1 = T->group_shiftcap[gb];
  const int nwarp = gridDim.x * (blockDim.x >> 5);
  // base item size: the smallest of 4 / 8 / 16 / 32 that gives every item its own warp (small batches: every query
  // chain as short as it can be).  When even 32 is not enough the kernel is in its throughput regime: groups with big
  // hulls keep short items (their cap), so that no warp is stuck with a 32-query item of 400-vertex scans while the
  // others idle, and the items are handed out dynamically, expensive groups first.
  int shift = 2;
  for (; shift < 5; shift++) {
    int items = ((c0 + (1 << shift) - 1) >> shift) + ((c1 + (1 << shift) - 1) >> shift);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) items += __shfl_xor_sync(0xffffffffu, items, o);
    if (items <= nwarp) break;
  }
  const int s0 = min(shift, cap0), s1 = min(shift, cap1);
  const int n0 = (c0 + (1 << s0) - 1) >> s0, n1 = (c1 + (1 << s1) - 1) >> s1;
  int inc0 = n0;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc0, o); if (lane >= o) inc0 += v; }
  const int tot0 = __shfl_sync(0xffffffffu, inc0, 31);
  int inc1 = n1;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc1, o); if (lane >= o) inc1 += v; }
  inc1 += tot0;
  const int total = __shfl_sync(0xffffffffu, inc1, 31);
  if (blockIdx.x >= total) return;  // more CTAs than items: the first `total` CTAs hold enough warps
#ifdef HMJ_PHASE_PROF
  if (blockIdx.x == 0 && threadIdx.x < 32) {
    int qsum = c0 + c1;
    for (int o = 16; o > 0; o >>= 1) qsum += __shfl_xor_sync(0xffffffffu, qsum, o);
    if (threadIdx.x == 0) { atomicAdd(&hmj_narrow_warp[4 + shift], 1ull); atomicAdd(&hmj_narrow_warp[10], (unsigned long long)total); atomicAdd(&hmj_narrow_warp[11], (unsigned long long)qsum); }
  }
#endif
#ifdef HMJ_PHASE_PROF
  const long long w_t0 = clock64();
#endif
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) mbar_init(&bar);
  __syncthreads();
  if (threadIdx.x == 0) stage_hulls_async(T, sx, sy, sz, 0, HMJ_NVERT_SMEM, &bar);   // dvx | dvy | dvz whole: 3 x 8.4 KB
  bool staged = false;
#ifdef HMJ_PHASE_PROF
  const long long w_t1 = clock64();
#endif
  int trips_ = 0;
  for (;;) {
    // dynamic hand-out (wq_count[63] is the item counter of this launch, zeroed with the group counts)
    int item = 0;
    if (lane == 0) item = atomicAdd(&B.wq_count[63], 1);
    item = __shfl_sync(0xffffffffu, item, 0);
    if (item >= total) break;
    if (!staged) { mbar_wait(&bar, 0); staged = true; }   // the copy ran beside the schedule look-up and the hand-out
#ifdef HMJ_PHASE_PROF
    const long long i_t0 = clock64();
    trips_ = 0;
#endif
    // schedule position = number of positions whose inclusive prefix is <= item
    const int pos = __popc(__ballot_sync(0xffffffffu, inc0 <= item)) + __popc(__ballot_sync(0xffffffffu, inc1 <= item));
    const int excl = (pos == 0) ? 0 : (pos <= 32 ? __shfl_sync(0xffffffffu, inc0, (pos - 1) & 31) : __shfl_sync(0xffffffffu, inc1, (pos - 1) & 31));
    const int cnt = (pos < 32) ? __shfl_sync(0xffffffffu, c0, pos & 31) : __shfl_sync(0xffffffffu, c1, pos & 31);
    const int g = (pos < 32) ? __shfl_sync(0xffffffffu, ga, pos & 31) : __shfl_sync(0xffffffffu, gb, pos & 31);
    const int ishift = (pos < 32) ? __shfl_sync(0xffffffffu, s0, pos & 31) : __shfl_sync(0xffffffffu, s1, pos & 31);
    const int isz = 1 << ishift;
    const int q0 = (item - excl) << ishift;
    const int q = q0 + lane;
    const bool valid = lane < isz && q < cnt;
    const int2 task = B.wq_task[(size_t)T->group_base[g] * B.n + (valid ? q : q0)];  // idle lanes shadow the item's first task
    if (g < HMJ_NMESH) {
      const Hull h2 = {T->mesh_adr[g], T->mesh_npad[g]};
      narrow_hfield(T, B, sx, sy, sz, sfr + threadIdx.x, h2, task, valid, lane, trips_);
    } else {
      narrow_mesh_mesh(T, B, sx, sy, sz, sfr + threadIdx.x, task, valid, lane, trips_);
    }
#ifdef HMJ_PHASE_PROF
    if (lane == 0) {
      const unsigned long long dt = (unsigned long long)(clock64() - i_t0);
      atomicAdd(&hmj_narrow_prof[g][0], 1ull); atomicAdd(&hmj_narrow_prof[g][1], dt); atomicMax(&hmj_narrow_prof[g][2], dt);
      atomicAdd(&hmj_narrow_prof[g][3], (unsigned long long)trips_); atomicMax(&hmj_narrow_prof[g][4], (unsigned long long)trips_);
    }
#endif
  }
  if (!staged) mbar_wait(&bar, 0);   // no warp leaves with the copy into its CTA's shared memory in flight
#ifdef HMJ_PHASE_PROF
  if (lane == 0) {
    const unsigned long long dt = (unsigned long long)(clock64() - w_t0);
    atomicMax(&hmj_narrow_warp[0], dt); atomicAdd(&hmj_narrow_warp[1], dt); atomicAdd(&hmj_narrow_warp[2], 1ull);
    atomicMax(&hmj_narrow_warp[3], (unsigned long long)(w_t1 - w_t0));
  }
#endif
  (void)trips_;
}

void hmj_launch_collide(const hmj_tables_t* tab, hmj_buffers_t buf, cudaStream_t stream, bool zero) {
  static int narrow_grids[HMJ_MAX_DEVICES] = {0};   // persistent grid = SMs x resident CTAs of the current device
  const int dev = hmj_current_device();
  if (!narrow_grids[dev]) {
    int sms = 148, per_sm = 2;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaFuncSetAttribute(k_narrow, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NARROW_SMEM_BYTES);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_narrow, NARROW_THREADS, NARROW_SMEM_BYTES);
    narrow_grids[dev] = sms * (per_sm > 0 ? per_sm : 1);
  }
  const int narrow_grid = narrow_grids[dev];
  if (zero) {   // callers whose k_smooth ran just before on this stream skip this: k_smooth zeroes both
    cudaMemsetAsync(buf.wq_count, 0, 64 * sizeof(int32_t), stream);
    cudaMemsetAsync(buf.slot_mask, 0, (size_t)buf.n * 4 * sizeof(uint32_t), stream);
  }
  const int npad = (buf.n + BROAD_THREADS - 1) / BROAD_THREADS * BROAD_THREADS;  // CTAs are pair-uniform
  const long long nthread = (long long)npad * HMJ_NPAIR;
  const char* force = getenv("HMJ_BROAD_COOP");   // developer override: 0 / 1
  const bool coop = force ? atoi(force) != 0 : buf.n <= BROAD_COOP_MAX_N;
  if (coop) {
    // small batch: the 9 height-field pairs (first in broad_order) go warp-per-(pair, env), the mesh-mesh pairs ride in the
    // trailing warps of the same launch
    const int nhf = HMJ_NPAIR - 28;
    const int warps = nhf * buf.n + (HMJ_NPAIR - nhf) * ((buf.n + 31) / 32);
    k_broad_hf_coop<<<(warps + 3) / 4, 128, 0, stream>>>(tab, buf, nhf);
  } else {
    k_broad<<<(int)(nthread / BROAD_THREADS), BROAD_THREADS, 0, stream>>>(tab, buf, npad, 0);
  }
  k_narrow<<<narrow_grid, NARROW_THREADS, NARROW_SMEM_BYTES, stream>>>(tab, buf);
}
