// hmj_solve.cu — last stage of the step: contacts -> constraints -> Newton solver -> Euler integrator.
//
// Restates mj_fwdAcceleration's solve (mujoco.h:256), mj_makeConstraint (:358),
// mj_referenceConstraint (:367), mj_fwdConstraint with the Newton solver (:259), mj_checkAcc (:307)
// and mj_Euler with implicit joint damping (:262) for this model (MuJoCo 3.1.4 semantics, defaults of
// mjmodel.h:403-444: pyramidal cones, condim 6 -> 10 rows per contact, tolerance 1e-8, 100 iterations,
// exact line search with ls_tolerance 0.01 / 50 iterations, refsafe, warmstart).
//
// Launch structure (one step):
//   k_classify   warp per env: counts contact slots and active joint limits, appends the env to one of three
//                work lists — class 0: no constraint rows, class 1: <= SOLVE_SMALL_CON contacts, class 2: the rest;
//   k_solve<C>   one warp (= one CTA) per listed env, one instantiation per class.  The classes differ only in
//                the shared-memory footprint (contact basis + per-row arrays sized for the class), which is what
//                bounds residency: class 0 keeps ~6 KB per env, class 1 ~15 KB, class 2 ~32 KB.
// Mapping inside k_solve: lane d owns dof d for all 18-vectors; constraint rows are strided over lanes
// (row r -> lane r % 32) with their scalars in shared memory; reductions are xor-butterfly shuffles, which give
// bit-identical sums on all lanes, so all solver control flow is warp-uniform.
//   * dense 18x18 M and Hessian factor in shared memory (row stride 19), rolled right-looking warp Cholesky;
//   * the 10 pyramid-edge rows of a contact are never materialised: only the 6 contact-frame basis rows
//     Jb[c][6][18] are stored (J_edge = Jn +- mu_k Jk), so J.v, J'.f and J'DJ all work on the basis.
// Every loop is rolled and the line-search / evaluation bodies exist once: the whole kernel is a few thousand
// SASS instructions and stays in the instruction cache (the unrolled predecessor was 300 KB of code and spent
// more than half of its stall cycles on instruction fetch: profiles/r1b_ncu_full_summary.csv).
// Roofline: FP64 latency (Cholesky pivots, line-search reductions); per-env HBM/L2 traffic is ~2.6 KB in
// (qM, cdof, slots, state) and ~0.6 KB out.
#include <cstdio>
#include "hmj_dev.cuh"

#define NV HMJ_NV
#define LD 19                   /* padded row stride of the dense 18x18 matrices */
#define NLIMMAX 24
#define NTRI (NV * (NV + 1) / 2)
#define SOLVE_SMALL_CON 4       /* class 1 holds envs with at most this many contacts */
#define SOLVE_MID_CON 8         /* class 2: up to this many; class 3: the rest (up to HMJ_MAXCON) */
#define HMJ_SOLVE_SELFCLASS 256  /* flag in k_solve's `integrate` argument */
#define SOLVE_SELFCLASS_MAX_N 512
#define SOLVE_CLASS_OF(MAXC) ((MAXC) == 0 ? 0 : ((MAXC) <= SOLVE_SMALL_CON ? 1 : ((MAXC) <= SOLVE_MID_CON ? 2 : 3)))

// staging rows of chol_factor_solve (below)
struct CholScratch {
  double col[2][32 + NV];   // staged column of the current pivot, double-buffered by pivot parity (tail: read padding)
  double xb[2][32];         // staged right-hand side
  double invd[NV];
};

template <int MAXC>
struct SolveSmem {
  static constexpr int MC = MAXC > 0 ? MAXC : 1;
  static constexpr int NL = MAXC > 0 ? NLIMMAX : 1;
  static constexpr int MR = MAXC > 0 ? NLIMMAX + 10 * MAXC : 1;
  double M[NV * LD];
  double H[NV * LD];
  double vec[NV];               // vector being multiplied (search / qacc / qvel)
  double qfs[NV];               // qfrc_smooth
  double com[3];
  CholScratch chol;
  // ---- constrained classes only
  double Jb[MC][6][LD];
  double jbv[MC][6];
  double W[MC][12];             // per-contact 6x6 weight in the basis: Wnn, Wnk[5], Wkk[5]
  double rowbuf[MR];
  double rD[MR], rAref[MR], rJaref[MR], rJv[MR];
  double cpos[MC][3];
  double cframe[MC][9];
  double cfric[MC][5];
  double cprm[MC][8];            // dist, solref[2], solimp[5]
  double lim_sign[NL], lim_pos[NL], lim_margin[NL];
  int32_t cslot[MC], cb1[MC], cb2[MC];
  int32_t lim_dof[NL];   // the hinge joint of a limit row is dof - 5 (free joint first)
};

// ---------------------------------------------------------------- dense helpers (warp cooperative)
// x <- H^-1 x for the SPD matrix whose lower triangle sits in shared memory (H = L L').  One fused routine:
//   * right-looking Cholesky with lane i holding row i in registers.  The pivot loop is ROLLED: the register row is
//     rotated by one column per pivot (the rank-1 update writes a[m-1] <- a[m] - l_ij l_kj), so the pivot column is
//     always a[0] and one ~110-instruction body serves all 18 pivots and stays in the L0 instruction cache (the fully
//     unrolled form was 20 KB of straight-line code and lost ~40 % of its issue slots to instruction fetch:
//     profiles/r1g_solve_lines.txt).  Late pivots update columns past 17, which only touch garbage registers.
//   * the scaled column j is broadcast through a shared-memory staging row (one store + 17 broadcast loads per pivot
//     instead of 34 shuffles);
//   * the forward substitution rides along (column j of L is final when pivot j is done); L is published to shared
//     memory for the backward substitution, which walks columns.
// On return H holds L (lower triangle; entries above the diagonal are garbage), S.invd the reciprocal diagonal.
// one pivot of chol_factor_solve; MC = how many of the rotated row entries are still alive (NV - 1 - j at pivot j, rounded
// up to the phase's bound): the rank-1 update touches only those
template <int MC>
__device__ __forceinline__ void chol_pivot(double (&a)[NV], double& dj, double& x, double& yv, double* H, CholScratch* C, int j,
                                           int lane) {
  const unsigned FULL = 0xffffffffu;
  if (dj < HMJ_MINVAL) dj = HMJ_MINVAL;
  const double inv = rsqrt(dj);
  const double lij = (lane == j) ? dj * inv : a[0] * inv;   // a[0] is this row's entry of column j
  // next pivot's diagonal, formed by its owner lane (j + 1) ahead of the column broadcast: this shuffle -> rsqrt ->
  // scale -> FMA is the only loop-carried chain; the staged broadcast below runs beside the next rsqrt
  const double dnext = a[1] - lij * lij;
  dj = __shfl_sync(FULL, dnext, (j + 1) & 31);
  double* col = C->col[j & 1];
  double* xb = C->xb[j & 1];
  col[lane] = lij;
  xb[lane] = x;
  if (lane >= j && lane < NV) H[lane * LD + j] = lij;
  if (lane == 0) C->invd[j] = inv;
  __syncwarp();
  const double* cj = col + j;                                // cj[m] = l(j + m, j)
  double cm[MC];                                             // all broadcast loads in flight before the first FMA
#pragma unroll
  for (int m = 1; m <= MC; m++) cm[m - 1] = cj[m];
#pragma unroll
  for (int m = 1; m <= MC; m++) a[m - 1] = a[m] - lij * cm[m - 1];
  // forward substitution step j: y_j = x_j / l_jj, then x_i -= l_ij y_j for i > j
  const double xj = xb[j] * inv;
  yv = (lane == j) ? xj : yv;
  const double lm = (lane > j) ? lij : 0.0;
  x = x - lm * xj;
}

__device__ __forceinline__ double chol_factor_solve(double* H, CholScratch* C, double x, int lane) {
  const unsigned FULL = 0xffffffffu;
  const int row = lane < NV ? lane : NV - 1;
  double a[NV];
  __syncwarp();
#pragma unroll
  for (int k = 0; k < NV; k++) a[k] = H[row * LD + k];
  double yv = 0.0;
  double dj = __shfl_sync(FULL, a[0], 0);
  // three phases of six pivots: the remaining matrix shrinks, and so does the update (17, 11, 5 entries per row instead
  // of 17 throughout: a third of the loop's broadcast loads and FMAs gone; same operations on the live entries)
  static_assert(NV == 18, "phases are written for 18 dofs");
#pragma unroll 1
  for (int j = 0; j < 6; j++) chol_pivot<17>(a, dj, x, yv, H, C, j, lane);
#pragma unroll 1
  for (int j = 6; j < 12; j++) chol_pivot<11>(a, dj, x, yv, H, C, j, lane);
#pragma unroll 1
  for (int j = 12; j < 18; j++) chol_pivot<5>(a, dj, x, yv, H, C, j, lane);
  __syncwarp();
  // backward substitution: lane k publishes x_k = y_k / l_kk, lanes below it eliminate column k
  x = yv;
  double res = 0.0;
#pragma unroll 3
  for (int k = NV - 1; k >= 0; k--) {
    const double xk = __shfl_sync(FULL, x, k) * C->invd[k];
    res = (lane == k) ? xk : res;
    const double l = (lane < k) ? H[k * LD + lane] : 0.0;
    x = x - l * xk;
  }
  return lane < NV ? res : 0.0;
}

// ---------------------------------------------------------------- Cholesky along the kinematic tree
// The robot is a free root (dofs 0-5) carrying four serial chains of three hinges (dofs 6-8, 9-11, 12-14, 15-17;
// checked against dof_parentid when the tables are built: hmj_tables_t::tree_chol).  M couples a chain only with itself
// and the root: the 18 x 18 matrix is  [ R  B' ; B  C ]  with C block-diagonal.  Eliminating leaf-first (MuJoCo's own
// order for M, mj_factorM) creates no fill, and the four chains are independent of one another, so the 18 dependent
// pivots of the dense routine become 3 steps (one dof of every chain at a time) + 6 root pivots.  Used for the two
// factorisations of k_factor (M and M + h diag(damping)) and mj_Euler's solve; the Newton Hessian keeps the dense routine
// (a contact between two chains fills C, and carrying both routines in k_solve cost more than the tree one saved).
// Lane i holds row i of the lower triangle: r[j] = H(i, j) for the root columns j (root lanes: j <= i), c[t] = H(i, first
// dof of its chain + t) for t <= its position.  Pivot k scales its own row (u = row / sqrt(d)), publishes it through a
// shared staging row, and every ancestor lane i subtracts u_i u_j from its entries: H = sum_k u_k u_k' = L' L with row k
// of L = u_k.  L' y = b rides along (b_j -= u_j y_k), L x = y walks root first, then down the chains.
struct TreeRow { double r[6]; double c[3]; double inv; };
#define TREE_STAGE 40   /* doubles per staging buffer: 4 chains x (6 root + 3 chain entries + y) */

__device__ __forceinline__ void tree_load(TreeRow& R, const double* __restrict__ H, int lane, double diag_add) {
  const int i = lane < NV ? lane : 0;   // idle lanes shadow row 0 and never publish
  const bool chain = lane >= 6 && lane < NV;
  const int pos = chain ? (lane - 6) % 3 : 0, first = i - pos;
#pragma unroll
  for (int j = 0; j < 6; j++) {
    R.r[j] = (chain || j <= i) ? H[i * LD + j] : 0.0;
    if (!chain && j == i) R.r[j] += diag_add;
  }
#pragma unroll
  for (int t = 0; t < 3; t++) {
    R.c[t] = (chain && t <= pos) ? H[i * LD + first + t] : 0.0;
    if (chain && t == pos) R.c[t] += diag_add;
  }
  R.inv = 0.0;
}

// factor (R becomes row `lane` of L and 1 / L(lane, lane)) and solve: returns x = H^-1 b for lanes < NV
__device__ __forceinline__ double tree_factor_solve(TreeRow& R, double* stage, double b, int lane) {
  const unsigned FULL = 0xffffffffu;
  const bool chain = lane >= 6 && lane < NV;
  const int ci = chain ? (lane - 6) / 3 : 0, pos = chain ? (lane - 6) - 3 * ci : -1;
  double y = 0.0;
  __syncwarp();
#pragma unroll
  for (int s = 0; s < 3; s++) {   // position q of every chain, deepest first
    const int q = 2 - s;
    double* st = stage + (s & 1) * TREE_STAGE;
    if (pos == q) {
      double d = R.c[q];
      if (d < HMJ_MINVAL) d = HMJ_MINVAL;
      const double inv = rsqrt(d);
      R.inv = inv;
#pragma unroll
      for (int j = 0; j < 6; j++) { R.r[j] *= inv; st[ci * 10 + j] = R.r[j]; }
#pragma unroll
      for (int t = 0; t < q; t++) { R.c[t] *= inv; st[ci * 10 + 6 + t] = R.c[t]; }
      R.c[q] = d * inv;
      y = b * inv;
      st[ci * 10 + 9] = y;
    }
    __syncwarp();
    if (chain && pos < q) {
      const double* u = st + ci * 10;
      const double ui = u[6 + pos];
#pragma unroll
      for (int j = 0; j < 6; j++) R.r[j] -= ui * u[j];
#pragma unroll
      for (int t = 0; t < q; t++) R.c[t] -= ui * u[6 + t];   // entries past this lane's position are never read
      b -= ui * u[9];
    } else if (lane < 6) {
#pragma unroll
      for (int cc = 0; cc < 4; cc++) {
        const double* u = st + cc * 10;
        const double ui = u[lane];
#pragma unroll
        for (int j = 0; j < 6; j++) R.r[j] -= ui * u[j];      // columns past the diagonal are never read
        b -= ui * u[9];
      }
    }
  }
#pragma unroll
  for (int k = 5; k >= 0; k--) {   // the root block
    double* st = stage + (k & 1) * TREE_STAGE;
    if (lane == k) {
      double d = R.r[k];
      if (d < HMJ_MINVAL) d = HMJ_MINVAL;
      const double inv = rsqrt(d);
      R.inv = inv;
#pragma unroll
      for (int j = 0; j < k; j++) { R.r[j] *= inv; st[j] = R.r[j]; }
      R.r[k] = d * inv;
      y = b * inv;
      st[6] = y;
    }
    __syncwarp();
    if (lane < k) {
      const double ui = st[lane];
#pragma unroll
      for (int j = 0; j < k; j++) R.r[j] -= ui * st[j];
      b -= ui * st[6];
    }
  }
  // L x = y: root rows in order, then down the chains
  double acc = y, x = 0.0;
#pragma unroll
  for (int k = 0; k < 6; k++) {
    const double xk = __shfl_sync(FULL, acc * R.inv, k);
    if (lane == k) x = xk;
    else if (lane > k && lane < NV) acc -= R.r[k] * xk;
  }
#pragma unroll
  for (int q = 0; q < 3; q++) {
    const double xq = __shfl_sync(FULL, acc * R.inv, chain ? 6 + 3 * ci + q : lane);
    if (pos == q) x = xq;
    else if (pos > q) acc -= R.c[q] * xq;
  }
  return lane < NV ? x : 0.0;
}

// x = (L' L)^-1 b with the factor rows already in R (k_factor -> mj_Euler)
__device__ __forceinline__ double tree_solve(const TreeRow& R, double* stage, double b, int lane) {
  const unsigned FULL = 0xffffffffu;
  const bool chain = lane >= 6 && lane < NV;
  const int ci = chain ? (lane - 6) / 3 : 0, pos = chain ? (lane - 6) - 3 * ci : -1;
  double y = 0.0;
  __syncwarp();
#pragma unroll
  for (int s = 0; s < 3; s++) {
    const int q = 2 - s;
    double* st = stage + (s & 1) * TREE_STAGE;
    if (pos == q) {
      y = b * R.inv;
#pragma unroll
      for (int j = 0; j < 6; j++) st[ci * 10 + j] = R.r[j] * y;
#pragma unroll
      for (int t = 0; t < q; t++) st[ci * 10 + 6 + t] = R.c[t] * y;
    }
    __syncwarp();
    if (chain && pos < q) b -= st[ci * 10 + 6 + pos];
    else if (lane < 6) b -= ((st[lane] + st[10 + lane]) + (st[20 + lane] + st[30 + lane]));
  }
#pragma unroll
  for (int k = 5; k >= 0; k--) {
    double* st = stage + (k & 1) * TREE_STAGE;
    if (lane == k) {
      y = b * R.inv;
#pragma unroll
      for (int j = 0; j < k; j++) st[j] = R.r[j] * y;
    }
    __syncwarp();
    if (lane < k) b -= st[lane];
  }
  double acc = y, x = 0.0;
#pragma unroll
  for (int k = 0; k < 6; k++) {
    const double xk = __shfl_sync(FULL, acc * R.inv, k);
    if (lane == k) x = xk;
    else if (lane > k && lane < NV) acc -= R.r[k] * xk;
  }
#pragma unroll
  for (int q = 0; q < 3; q++) {
    const double xq = __shfl_sync(FULL, acc * R.inv, chain ? 6 + 3 * ci + q : lane);
    if (pos == q) x = xq;
    else if (pos > q) acc -= R.c[q] * xq;
  }
  return lane < NV ? x : 0.0;
}

// the same with the rows read from k_factor's record ([10][NV]: r[0..5], c[0..2], 1 / diagonal).  Out of line: once per
// env-step, and its ten row registers stay out of the solver's allocation.
__device__ __noinline__ double tree_solve_stored(const double* __restrict__ F, double* stage, double b, int lane) {
  TreeRow R;
  const int li = lane < NV ? lane : 0;
#pragma unroll
  for (int j = 0; j < 6; j++) R.r[j] = F[j * NV + li];
#pragma unroll
  for (int t = 0; t < 3; t++) R.c[t] = F[(6 + t) * NV + li];
  R.inv = F[9 * NV + li];
  return tree_solve(R, stage, b, lane);
}

__device__ __noinline__ double mat_vec(const double* A, const double* v, int lane) {
  double s = 0.0;
  if (lane < NV) {
#pragma unroll 6
    for (int j = 0; j < NV; j++) s += A[lane * LD + j] * v[j];
  }
  return s;
}

// pow for the impedance curve: the exponents that occur are solimp[4] = 2 (default) and solimp[4] - 1 = 1, where pow is
// exactly x * x and x; anything else goes through an out-of-line call (the inlined pow was 5 KB of code in the set-up
// stream of every solver warp, four times over)
__device__ __noinline__ double pow_general(double x, double p) { return pow(x, p); }
__device__ __forceinline__ double pow_imp(double x, double p) { return p == 2.0 ? x * x : (p == 1.0 ? x : pow_general(x, p)); }

__device__ __forceinline__ void get_impedance(const double* solimp, double pos, double margin, double* imp) {
  if (solimp[0] == solimp[1] || solimp[2] <= HMJ_MINVAL) { *imp = 0.5 * (solimp[0] + solimp[1]); return; }
  double x = (pos - margin) / solimp[2];
  if (x < 0) x = -x;
  if (x >= 1 || x <= 0) { *imp = (x >= 1 ? solimp[1] : solimp[0]); return; }
  double y;
  if (solimp[4] == 1) {
    y = x;
  } else if (x <= solimp[3]) {
    double a = 1 / pow_imp(solimp[3], solimp[4] - 1);
    y = a * pow_imp(x, solimp[4]);
  } else {
    double b = 1 / pow_imp(1 - solimp[3], solimp[4] - 1);
    y = 1 - b * pow_imp(1 - x, solimp[4]);
  }
  *imp = solimp[0] + y * (solimp[1] - solimp[0]);
}

// R, K*imp and B of one constraint row (mj_makeImpedance)
__device__ __forceinline__ void row_impedance(const double* solref_in, const double* solimp_in, double pos, double margin,
                                              double diagApprox, double timestep, double* R, double* KI, double* Bc) {
  double solref[2] = {solref_in[0], solref_in[1]};
  double solimp[5];
  if (solref[0] > 0) solref[0] = fmax(solref[0], 2 * timestep);  // refsafe
  solimp[0] = fmin(0.9999, fmax(0.0001, solimp_in[0]));
  solimp[1] = fmin(0.9999, fmax(0.0001, solimp_in[1]));
  solimp[2] = fmax(0.0, solimp_in[2]);
  solimp[3] = fmin(0.9999, fmax(0.0001, solimp_in[3]));
  solimp[4] = fmax(1.0, solimp_in[4]);
  double imp;
  get_impedance(solimp, pos, margin, &imp);
  *R = fmax(HMJ_MINVAL, (1 - imp) * diagApprox / imp);
  double K, Bv;
  if (solref[0] > 0) {
    K = 1 / fmax(HMJ_MINVAL, solimp[1] * solimp[1] * solref[0] * solref[0] * solref[1] * solref[1]);
    Bv = 2 / fmax(HMJ_MINVAL, solimp[1] * solref[0]);
  } else {
    K = -solref[0] / fmax(HMJ_MINVAL, solimp[1] * solimp[1]);
    Bv = -solref[1] / fmax(HMJ_MINVAL, solimp[1]);
  }
  *KI = K * imp;
  *Bc = Bv;
}

// mju_makeFrame.  Written with __dmul_rn/__dadd_rn (never contracted into FMA) so the contact frame is the
// same IEEE operation sequence as the oracle's (-ffp-contract=off): contact records compare bit-exact.
__device__ __forceinline__ double nf_dot3(const double* a, const double* b) {
  return __dadd_rn(__dadd_rn(__dmul_rn(a[0], b[0]), __dmul_rn(a[1], b[1])), __dmul_rn(a[2], b[2]));
}
__device__ __forceinline__ void nf_normalize3(double* v) {
  const double n = sqrt(nf_dot3(v, v));
  if (n < HMJ_MINVAL) {
    v[0] = 1; v[1] = 0; v[2] = 0;
  } else {
    const double s = 1.0 / n;
    v[0] = __dmul_rn(v[0], s); v[1] = __dmul_rn(v[1], s); v[2] = __dmul_rn(v[2], s);
  }
}
__device__ __forceinline__ void make_frame(double* frame) {
  nf_normalize3(frame);
  if (sqrt(nf_dot3(frame + 3, frame + 3)) < 0.5) {
    frame[3] = frame[4] = frame[5] = 0;
    if (frame[1] < 0.5 && frame[1] > -0.5) frame[4] = 1; else frame[5] = 1;
  }
  const double tmp = nf_dot3(frame, frame + 3);
  frame[3] = __dsub_rn(frame[3], __dmul_rn(tmp, frame[0]));
  frame[4] = __dsub_rn(frame[4], __dmul_rn(tmp, frame[1]));
  frame[5] = __dsub_rn(frame[5], __dmul_rn(tmp, frame[2]));
  nf_normalize3(frame + 3);
  frame[6] = __dsub_rn(__dmul_rn(frame[1], frame[5]), __dmul_rn(frame[2], frame[4]));
  frame[7] = __dsub_rn(__dmul_rn(frame[2], frame[3]), __dmul_rn(frame[0], frame[5]));
  frame[8] = __dsub_rn(__dmul_rn(frame[0], frame[4]), __dmul_rn(frame[1], frame[3]));
}

struct LsPnt { double alpha, cost, d0, d1; };

// HMJ_PHASE_PROF: per-phase latency accounting (developer build only): lane 0 adds the cycles since the previous
// mark to hmj_phase_cycles[phase]; read back with cudaMemcpyFromSymbol through hmj_debug_phase_cycles.
#ifdef HMJ_PHASE_PROF
__device__ unsigned long long hmj_phase_cycles[16];
__device__ unsigned int hmj_phase_env[16384][16];   // the same, per env (the tail is what bounds a single-wave launch)
#define PH(i) do { const long long t_ = clock64(); if (lane == 0) { atomicAdd(&hmj_phase_cycles[i], (unsigned long long)(t_ - ph_t0)); \
    hmj_phase_env[env & 16383][i] += (unsigned int)(t_ - ph_t0); } ph_t0 = clock64(); } while (0)
#else
#define PH(i) do { } while (0)
#endif

// Geometry of the constraint rows of one env: limit rows first, then 10 pyramid edges per contact.
struct RowMap {
  int nlim, nefc, ncon;
  const double* lim_sign;
  const int32_t* lim_dof;
  const double (*cfric)[5];
  const double (*jbv)[6];
  const double* vec;
};

// (J v)_r from the basis products jbv = Jb v (contacts) or the single +-1 entry (limits)
__device__ __forceinline__ double row_value(const RowMap& R, int r) {
  if (r < R.nlim) return R.lim_sign[r] * R.vec[R.lim_dof[r]];
  const int q = r - R.nlim, c = q / 10, e = q - 10 * c, kk = 1 + (e >> 1);
  const double mu = (e & 1) ? -R.cfric[c][kk - 1] : R.cfric[c][kk - 1];
  return R.jbv[c][0] + mu * R.jbv[c][kk];
}

// jbv[c][k] = Jb[c][k] . vec
__device__ __noinline__ void basis_dot(const double (*Jb)[6][LD], const double* vec, double (*jbv)[6], int ncon, int lane) {
  __syncwarp();
#pragma unroll 1
  for (int t = lane; t < ncon * 6; t += 32) {
    const int c = t / 6, kk = t - c * 6;
    double s = 0.0;
#pragma unroll 6
    for (int d = 0; d < NV; d++) s += Jb[c][kk][d] * vec[d];
    jbv[c][kk] = s;
  }
  __syncwarp();
}

// One evaluation of the line-search objective and its derivatives at p.alpha (mj_solver PrimalEval): the quadratic
// of every row that is active at alpha, summed over the warp, plus the Gauss term (qg).
__device__ __forceinline__ void ls_eval(LsPnt& p, const double* rD, const double* rJaref, const double* rJv, int nefc,
                                     double qg0, double qg1, double qg2, int lane) {
  double a0 = 0, a1 = 0, a2 = 0;
#pragma unroll 1
  for (int r = lane; r < nefc; r += 32) {
    const double ja = rJaref[r], jv = rJv[r];
    if (ja + p.alpha * jv < 0) {
      const double D = rD[r], DJ0 = D * ja;
      a0 += 0.5 * (ja * DJ0);
      a1 += jv * DJ0;
      a2 += 0.5 * (jv * D * jv);
    }
  }
  const double q0 = qg0 + warp_sum(a0), q1 = qg1 + warp_sum(a1), q2 = qg2 + warp_sum(a2);
  p.cost = p.alpha * p.alpha * q2 + p.alpha * q1 + q0;
  p.d0 = 2 * p.alpha * q2 + q1;
  p.d1 = 2 * q2;
  if (p.d1 <= 0) p.d1 = HMJ_MINVAL;
}

// x <- (L L')^-1 x for a factor already in shared memory (lower triangle of H, reciprocal diagonal in invd)
__device__ __noinline__ double chol_solve_only(const double* H, const double* invd, double x, int lane) {
  const unsigned FULL = 0xffffffffu;
  double yv = 0.0;
#pragma unroll 3
  for (int j = 0; j < NV; j++) {
    const double xj = __shfl_sync(FULL, x, j) * invd[j];
    yv = (lane == j) ? xj : yv;
    const double l = (lane > j && lane < NV) ? H[lane * LD + j] : 0.0;
    x = x - l * xj;
  }
  x = yv;
  double res = 0.0;
#pragma unroll 3
  for (int k = NV - 1; k >= 0; k--) {
    const double xk = __shfl_sync(FULL, x, k) * invd[k];
    res = (lane == k) ? xk : res;
    const double l = (lane < k) ? H[k * LD + lane] : 0.0;
    x = x - l * xk;
  }
  return lane < NV ? res : 0.0;
}

// ---------------------------------------------------------------- factorisations that do not depend on contacts
// Warp per env, launched on a side stream right after k_smooth so it runs beside the collision kernels:
//   qacc_smooth = M^-1 qfrc_smooth  (dense Cholesky instead of the sparse L'DL: same result to rounding), and
//   the factor of M + h diag(damping) that mj_Euler's implicit-damping solve needs after the constraint solve.
// Both were the first / last thing every k_solve warp did; hoisting them removes two factorisations from each env's
// serial chain and the start-up burst in which every resident warp factorised at once.
#define FAC_DOUBLES HMJ_FAC_DOUBLES   /* lower triangle of L (tri order), reciprocal diagonal, padding */
static_assert(FAC_DOUBLES >= NTRI + NV, "Euler factor record too small");
struct FactorSmem {
  double M[NV * LD];
  double H[NV * LD];
  CholScratch chol;
};
__global__ void __launch_bounds__(32) k_factor(const hmj_tables_t* __restrict__ T, hmj_buffers_t B) {
  __shared__ FactorSmem S;
  const int lane = threadIdx.x & 31;
  const int env = blockIdx.x;
  if (env >= B.n) return;
  if (B.active && !B.active[env]) return;
  const hmj_model_t& M = T->m;
  const double h = B.params[env].timestep;
  for (int t = lane; t < NV * LD; t += 32) S.M[t] = 0.0;
  __syncwarp();
  for (int e = lane; e < HMJ_NM; e += 32) {
    const double v = B.qM[(size_t)env * HMJ_NM + e];
    const int i = T->M_row[e], j = T->M_col[e];
    S.M[i * LD + j] = v;
    S.M[j * LD + i] = v;
  }
  const double qfs = lane < NV ? B.qfrc_smooth[(size_t)env * NV + lane] : 0.0;
  __syncwarp();
  if (T->tree_chol) {   // M and M + h diag(damping) have the tree's sparsity: 9 pivot steps each instead of 18
    TreeRow R;
    tree_load(R, S.M, lane, 0.0);
    const double qa = tree_factor_solve(R, &S.chol.col[0][0], qfs, lane);
    if (lane < NV) B.qacc_smooth[(size_t)env * NV + lane] = qa;
    tree_load(R, S.M, lane, lane < NV ? h * M.dof_damping[lane] : 0.0);
    tree_factor_solve(R, &S.chol.col[0][0], 0.0, lane);
    double* F = B.euler_factor + (size_t)env * FAC_DOUBLES;   // [10][NV]: r[0..5], c[0..2], 1 / diagonal of every row
    if (lane < NV) {
#pragma unroll
      for (int j = 0; j < 6; j++) F[j * NV + lane] = R.r[j];
#pragma unroll
      for (int t = 0; t < 3; t++) F[(6 + t) * NV + lane] = R.c[t];
      F[9 * NV + lane] = R.inv;
    }
    return;
  }
  for (int t = lane; t < NV * LD; t += 32) S.H[t] = S.M[t];
  const double qas = chol_factor_solve(S.H, &S.chol, qfs, lane);
  if (lane < NV) B.qacc_smooth[(size_t)env * NV + lane] = qas;
  __syncwarp();
  for (int t = lane; t < NV * LD; t += 32) S.H[t] = S.M[t];
  __syncwarp();
  if (lane < NV) S.H[lane * LD + lane] += h * M.dof_damping[lane];
  chol_factor_solve(S.H, &S.chol, 0.0, lane);
  __syncwarp();
  double* F = B.euler_factor + (size_t)env * FAC_DOUBLES;
  for (int t = lane; t < NTRI; t += 32) F[t] = S.H[T->tri_i[t] * LD + T->tri_j[t]];
  if (lane < NV) F[NTRI + lane] = S.chol.invd[lane];
}

void hmj_launch_factor(const hmj_tables_t* tab, hmj_buffers_t buf, cudaStream_t stream, const hmj_solve_ctx_t* ctx) {
  if (!ctx) {
    k_factor<<<buf.n, 32, 0, stream>>>(tab, buf);
    return;
  }
  cudaEventRecord(ctx->fac_fork, stream);
  cudaStreamWaitEvent(ctx->side[0], ctx->fac_fork, 0);
  k_factor<<<buf.n, 32, 0, ctx->side[0]>>>(tab, buf);
  cudaEventRecord(ctx->fac_done, ctx->side[0]);
}

// ---------------------------------------------------------------- classification
// Warp per env: number of valid contact slots and whether any joint limit is active -> work class.
// Class lists are filled with atomics (order is irrelevant: envs are independent); counters live in
// B.solve_count[0..2] and are zeroed by hmj_launch_solve before this kernel.
__global__ void __launch_bounds__(128) k_classify(const hmj_tables_t* __restrict__ T, hmj_buffers_t B, int force_rows) {
  const int lane = threadIdx.x & 31;
  const int env = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (env >= B.n) return;
  if (B.active && !B.active[env]) return;
  const unsigned FULL = 0xffffffffu;
  const hmj_model_t& M = T->m;
  const hmj_params_t* P = B.params + env;
  const uint32_t mw = lane < 4 ? B.slot_mask[(size_t)env * 4 + lane] : 0u;
  int ncon = __popc(mw);
#pragma unroll
  for (int o = 2; o > 0; o >>= 1) ncon += __shfl_xor_sync(FULL, ncon, o);
  ncon = __shfl_sync(FULL, ncon, 0);
  bool lim = false;
  const int j = lane + 1;
  if (j < HMJ_NJNT && M.jnt_limited[j]) {
    const double margin = P->jnt_margin[j];
    const double value = B.qpos[(size_t)env * HMJ_NQ + M.jnt_qposadr[j]];
    lim = (-(P->jnt_range[j][0] - value) < margin) || (P->jnt_range[j][1] - value < margin);
  }
  const bool anylim = __any_sync(FULL, lim);
  if (lane == 0) {
    // force_rows (mj_inverse): the unconstrained class does not carry the dense M the inverse needs
    const int cls = (ncon == 0 && !anylim && !force_rows) ? 0 : (ncon <= SOLVE_SMALL_CON ? 1 : (ncon <= SOLVE_MID_CON ? 2 : 3));
    const int at = atomicAdd(&B.solve_count[cls], 1);
    B.solve_list[(size_t)cls * B.n + at] = env;
  }
}

// ---------------------------------------------------------------- the solver
template <int MAXC, bool INV = false>
#ifndef HMJ_SOLVE4_WARPS
#define HMJ_SOLVE4_WARPS 14
#endif
__global__ void __launch_bounds__(32, MAXC == 0 ? 24 : (MAXC <= SOLVE_SMALL_CON ? HMJ_SOLVE4_WARPS : (MAXC <= SOLVE_MID_CON ? 10 : 6)))
k_solve(const hmj_tables_t* __restrict__ T, hmj_buffers_t B, int integrate) {
  using Smem = SolveSmem<MAXC>;
  extern __shared__ double smem_raw[];
  const int lane = threadIdx.x & 31;
  int env;
  if (integrate & HMJ_SOLVE_SELFCLASS) {
    // small batches: no k_classify launch; CTA b looks at env b and leaves unless the env belongs to this class
    // (one kernel and one dependency edge less on the critical path of a step; the test is k_classify's)
    env = blockIdx.x;
    if (env >= B.n || (B.active && !B.active[env])) return;
    const hmj_model_t& Mm = T->m;
    const hmj_params_t* Pp = B.params + env;
    const uint32_t mw = lane < 4 ? B.slot_mask[(size_t)env * 4 + lane] : 0u;
    int nc = __popc(mw);
#pragma unroll
    for (int o = 2; o > 0; o >>= 1) nc += __shfl_xor_sync(0xffffffffu, nc, o);
    nc = __shfl_sync(0xffffffffu, nc, 0);
    bool lim = false;
    const int j = lane + 1;
    if (j < HMJ_NJNT && Mm.jnt_limited[j]) {
      const double margin = Pp->jnt_margin[j];
      const double value = B.qpos[(size_t)env * HMJ_NQ + Mm.jnt_qposadr[j]];
      lim = (-(Pp->jnt_range[j][0] - value) < margin) || (Pp->jnt_range[j][1] - value < margin);
    }
    const bool anylim = __any_sync(0xffffffffu, lim);
    const int cls = (nc == 0 && !anylim && !INV) ? 0 : (nc <= SOLVE_SMALL_CON ? 1 : (nc <= SOLVE_MID_CON ? 2 : 3));
    if (cls != SOLVE_CLASS_OF(MAXC)) return;
  } else {
    if ((int)blockIdx.x >= B.solve_count[SOLVE_CLASS_OF(MAXC)]) return;
    env = B.solve_list[(size_t)SOLVE_CLASS_OF(MAXC) * B.n + blockIdx.x];
  }
  integrate &= 1;
  Smem& S = *reinterpret_cast<Smem*>(smem_raw);
  const hmj_model_t& M = T->m;
  const hmj_params_t* P = B.params + env;
  const unsigned FULL = 0xffffffffu;
  const double h = P->timestep;
#ifdef HMJ_PHASE_PROF
  long long ph_t0 = clock64();
#endif

  // ---------------- load: dense M (constrained classes only), qfrc_smooth, state, qacc_smooth from k_factor
  if constexpr (MAXC > 0) {
    for (int t = lane; t < NV * LD; t += 32) S.M[t] = 0.0;
    __syncwarp();
    for (int e = lane; e < HMJ_NM; e += 32) {
      const double v = B.qM[(size_t)env * HMJ_NM + e];
      const int i = T->M_row[e], j = T->M_col[e];
      S.M[i * LD + j] = v;
      S.M[j * LD + i] = v;
    }
  }
  double qfs = 0, qvel = 0, qws = 0, qas = 0;
  if (lane < NV) {
    qfs = B.qfrc_smooth[(size_t)env * NV + lane];
    qvel = B.qvel[(size_t)env * NV + lane];
    qws = B.warmstart[(size_t)env * NV + lane];
    qas = B.qacc_smooth[(size_t)env * NV + lane];
    S.qfs[lane] = qfs;
  }
  if (lane < 3) S.com[lane] = B.com[(size_t)env * 3 + lane];
  __syncwarp();
  PH(1);

  double qacc = qas;  // lanes < NV
  if constexpr (INV) qacc = lane < NV ? B.qacc[(size_t)env * NV + lane] : 0.0;   // mj_inverse: the acceleration is an input
  double qfc = 0.0;
  int niter = 0, ncon = 0, nefc = 0, status = 0;

  if constexpr (MAXC > 0) {
    // ---------------- contacts: ordered compaction of the narrowphase slots (slot order == MuJoCo contact order)
    {
      const uint32_t mw = lane < 4 ? B.slot_mask[(size_t)env * 4 + lane] : 0u;
#pragma unroll
      for (int w = 0; w < 4; w++) {
        const unsigned bal = __shfl_sync(FULL, mw, w);   // slots 32 w .. 32 w + 31
        const bool valid = (bal >> lane) & 1u;
        const int idx = ncon + __popc(bal & ((1u << lane) - 1u));
        if (valid && idx < MAXC) S.cslot[idx] = 32 * w + lane;
        ncon += __popc(bal);
      }
      if (ncon > MAXC) { ncon = MAXC; status |= HMJ_STATUS_CONTACTFULL; }  // only reachable for MAXC == HMJ_MAXCON
    }
    __syncwarp();
    if (lane < ncon) {  // lane c owns contact c: parameters, frame (mj_contactParam, mju_makeFrame)
      const int s = S.cslot[lane];
      const double* sl = B.slots + ((size_t)env * HMJ_NSLOT + s) * HMJ_SLOT_DOUBLES;
      const int k = T->slot_pair[s];
      const int g1 = M.pair_geom1[k], g2 = M.pair_geom2[k];
      hmj_contact_t* C = B.contact + (size_t)env * HMJ_MAXCON + lane;
      double frame[9] = {sl[4], sl[5], sl[6], 0, 0, 0, 0, 0, 0};
      make_frame(frame);
      const double* f1 = (g1 == 0) ? P->floor_friction : M.geom_friction[g1];
      const double* f2 = (g2 == 0) ? P->floor_friction : M.geom_friction[g2];
      const double fr0 = fmax(f1[0], f2[0]), fr1 = fmax(f1[1], f2[1]), fr2 = fmax(f1[2], f2[2]);
      const double fric[5] = {fr0, fr0, fr1, fr2, fr2};
      double mix;
      const double s1 = M.geom_solmix[g1], s2 = M.geom_solmix[g2];
      if (s1 >= HMJ_MINVAL && s2 >= HMJ_MINVAL) mix = s1 / (s1 + s2);
      else if (s1 < HMJ_MINVAL && s2 < HMJ_MINVAL) mix = 0.5;
      else if (s1 < HMJ_MINVAL) mix = 0.0;
      else mix = 1.0;
      for (int q = 0; q < 2; q++)
        C->solref[q] = (M.geom_solref[g1][0] > 0 && M.geom_solref[g2][0] > 0)
                           ? mix * M.geom_solref[g1][q] + (1 - mix) * M.geom_solref[g2][q]
                           : fmin(M.geom_solref[g1][q], M.geom_solref[g2][q]);
      for (int q = 0; q < 5; q++) C->solimp[q] = mix * M.geom_solimp[g1][q] + (1 - mix) * M.geom_solimp[g2][q];
      C->dist = sl[0];
      S.cprm[lane][0] = sl[0];
      for (int q = 0; q < 2; q++) S.cprm[lane][1 + q] = C->solref[q];
      for (int q = 0; q < 5; q++) S.cprm[lane][3 + q] = C->solimp[q];
      for (int q = 0; q < 3; q++) { C->pos[q] = sl[1 + q]; S.cpos[lane][q] = sl[1 + q]; }
      for (int q = 0; q < 9; q++) { C->frame[q] = frame[q]; S.cframe[lane][q] = frame[q]; }
      for (int q = 0; q < 5; q++) { C->friction[q] = fric[q]; S.cfric[lane][q] = fric[q]; }
      C->g1 = g1; C->g2 = g2;
      C->geom1 = M.geom_mjid[g1]; C->geom2 = M.geom_mjid[g2];
      S.cb1[lane] = M.geom_bodyid[g1]; S.cb2[lane] = M.geom_bodyid[g2];
    }

    // ---------------- joint-limit rows (mj_instantiateLimit): lane j-1 tests hinge joint j, lower then upper side
    int nlim = 0;
    {
      bool act0 = false, act1 = false;
      double dist0 = 0, dist1 = 0, margin = 0;
      const int j = lane + 1;
      if (j < HMJ_NJNT && M.jnt_limited[j]) {
        margin = P->jnt_margin[j];
        const double value = B.qpos[(size_t)env * HMJ_NQ + M.jnt_qposadr[j]];
        dist0 = -(P->jnt_range[j][0] - value);  // side = -1
        dist1 = P->jnt_range[j][1] - value;     // side = +1
        act0 = dist0 < margin;
        act1 = dist1 < margin;
      }
      const unsigned b0 = __ballot_sync(FULL, act0), b1 = __ballot_sync(FULL, act1);
      const unsigned below = (1u << lane) - 1u;
      const int before = __popc(b0 & below) + __popc(b1 & below);
      if (act0) {
        S.lim_dof[before] = M.jnt_dofadr[j]; S.lim_sign[before] = 1.0; S.lim_pos[before] = dist0;
        S.lim_margin[before] = margin;
      }
      if (act1) {
        const int r = before + (act0 ? 1 : 0);
        S.lim_dof[r] = M.jnt_dofadr[j]; S.lim_sign[r] = -1.0; S.lim_pos[r] = dist1;
        S.lim_margin[r] = margin;
      }
      nlim = __popc(b0) + __popc(b1);
    }
    nefc = nlim + 10 * ncon;
    __syncwarp();
    PH(2);

    if (nefc > 0) {
      // ---------------- contact-frame Jacobian basis Jb[c][k][d] (mj_jacDifPair rotated into the contact frame)
#pragma unroll 1
      for (int t = lane; t < ncon * NV; t += 32) {
        const int c = t / NV, d = t - c * NV;
        const double sgn = (double)((T->body_dofmask[S.cb2[c]] >> d) & 1u) - (double)((T->body_dofmask[S.cb1[c]] >> d) & 1u);
        double jp[3] = {0, 0, 0}, jr[3] = {0, 0, 0};
        if (sgn != 0.0) {
          const double* cd = B.cdof + ((size_t)env * NV + d) * 6;
          const double cdv[6] = {cd[0], cd[1], cd[2], cd[3], cd[4], cd[5]};
          const double off[3] = {S.cpos[c][0] - S.com[0], S.cpos[c][1] - S.com[1], S.cpos[c][2] - S.com[2]};
          double tmp[3];
          d_cross(tmp, cdv, off);
          jp[0] = sgn * (cdv[3] + tmp[0]); jp[1] = sgn * (cdv[4] + tmp[1]); jp[2] = sgn * (cdv[5] + tmp[2]);
          jr[0] = sgn * cdv[0]; jr[1] = sgn * cdv[1]; jr[2] = sgn * cdv[2];
        }
        const double* fr = S.cframe[c];
#pragma unroll
        for (int r = 0; r < 3; r++) {
          S.Jb[c][r][d] = fr[3 * r] * jp[0] + fr[3 * r + 1] * jp[1] + fr[3 * r + 2] * jp[2];
          S.Jb[c][3 + r][d] = fr[3 * r] * jr[0] + fr[3 * r + 1] * jr[1] + fr[3 * r + 2] * jr[2];
        }
      }
      __syncwarp();

      RowMap R;
      R.nlim = nlim; R.nefc = nefc; R.ncon = ncon;
      R.lim_sign = S.lim_sign; R.lim_dof = S.lim_dof; R.cfric = S.cfric; R.jbv = S.jbv; R.vec = S.vec;

      // ---------------- efc_vel = J qvel -> aref ; D  (mj_makeImpedance, mj_referenceConstraint)
      if (lane < NV) S.vec[lane] = qvel;
      basis_dot(S.Jb, S.vec, S.jbv, ncon, lane);
#pragma unroll 1
      for (int r = lane; r < nefc; r += 32) {
        double Rr, KI, Bc, pos, margin, diag, scaleR;
        const double *solref, *solimp;
        if (r < nlim) {
          const int j = M.dof_jntid[S.lim_dof[r]];
          pos = S.lim_pos[r]; margin = S.lim_margin[r];
          solref = M.jnt_solref[j]; solimp = M.jnt_solimp[j];
          diag = M.dof_invweight0[S.lim_dof[r]];
          scaleR = 1.0;
        } else {
          const int c = (r - nlim) / 10;
          pos = S.cprm[c][0]; margin = 0.0;
          solref = &S.cprm[c][1]; solimp = &S.cprm[c][3];
          const double tran = M.body_invweight0[S.cb1[c]][0] + M.body_invweight0[S.cb2[c]][0];
          const double mu0 = S.cfric[c][0];
          // all edges share Rpy = 2 mu0^2 R(first edge); first edge: diagApprox = tran + mu0^2 tran
          diag = tran + mu0 * mu0 * tran;
          scaleR = 2 * mu0 * mu0;
        }
        row_impedance(solref, solimp, pos, margin, diag, h, &Rr, &KI, &Bc);
        if (r >= nlim) Rr = scaleR * Rr;
        S.rD[r] = 1 / Rr;
        const double vel = row_value(R, r);
        S.rAref[r] = -Bc * vel - KI * (pos - margin);
      }
      __syncwarp();

      // Jaref = J * S.vec - aref for every row
      auto jaref_of = [&]() {
        basis_dot(S.Jb, S.vec, S.jbv, ncon, lane);
#pragma unroll 1
        for (int r = lane; r < nefc; r += 32) S.rJaref[r] = row_value(R, r) - S.rAref[r];
      };
      auto constraint_cost = [&]() -> double {
        double s = 0.0;
#pragma unroll 1
        for (int r = lane; r < nefc; r += 32) {
          const double jar = S.rJaref[r];
          if (jar < 0) s += 0.5 * S.rD[r] * jar * jar;
        }
        return warp_sum(s);
      };

      PH(3);
      if constexpr (INV) {
        // ---------------- mj_invConstraint: forces of the GIVEN acceleration, no solve.  Jaref = J qacc - aref, every
        // active row (pyramidal cones: plain inequalities) pushes with -D Jaref, qfrc_constraint = J' f
        __syncwarp();
        if (lane < NV) S.vec[lane] = qacc;
        __syncwarp();
        jaref_of();
        __syncwarp();
#pragma unroll 1
        for (int r = lane; r < nefc; r += 32) {
          const double jar = S.rJaref[r];
          S.rowbuf[r] = jar < 0 ? -S.rD[r] * jar : 0.0;
        }
        __syncwarp();
#pragma unroll 1
        for (int t = lane; t < ncon * 6; t += 32) {
          const int c = t / 6, kk = t - c * 6;
          const double* f = S.rowbuf + nlim + 10 * c;
          double v;
          if (kk == 0) {
            v = 0.0;
#pragma unroll
            for (int e = 0; e < 10; e++) v += f[e];
          } else {
            v = S.cfric[c][kk - 1] * (f[2 * (kk - 1)] - f[2 * (kk - 1) + 1]);
          }
          S.jbv[c][kk] = v;
        }
        __syncwarp();
        if (lane < NV) {
#pragma unroll 1
          for (int c = 0; c < ncon; c++) {
#pragma unroll
            for (int kk = 0; kk < 6; kk++) qfc += S.Jb[c][kk][lane] * S.jbv[c][kk];
          }
#pragma unroll 1
          for (int r = 0; r < nlim; r++)
            if (S.lim_dof[r] == lane) qfc += S.lim_sign[r] * S.rowbuf[r];
        }
      } else {
      // ---------------- warmstart: best of (qacc_warmstart, qacc_smooth)
      {
        __syncwarp();
        if (lane < NV) S.vec[lane] = qws;
        __syncwarp();
        jaref_of();
        double cost_ws = constraint_cost();
        const double Ma_ws = mat_vec(S.M, S.vec, lane);
        cost_ws += warp_sum(lane < NV ? 0.5 * (Ma_ws - qfs) * (qws - qas) : 0.0);
        __syncwarp();
        if (lane < NV) S.vec[lane] = qas;
        __syncwarp();
        jaref_of();
        const double cost_sm = constraint_cost();
        qacc = (cost_ws > cost_sm) ? qas : qws;
        __syncwarp();
      }
      PH(4);

      // ---------------- Newton iterations (mj_solNewton, primal)
      const double scale = 1.0 / (M.meaninertia * (double)NV);
      const double tolerance = M.tolerance;
      const int max_iter = M.iterations, ls_iterations = M.ls_iterations;
      const double ls_tolerance = M.ls_tolerance;
      double Ma, Mv = 0.0, gauss = 0.0, grad = 0.0, mgrad = 0.0, search = 0.0, cost = 0.0, oldcost = 0.0;

      __syncwarp();
      if (lane < NV) S.vec[lane] = qacc;
      __syncwarp();
      Ma = mat_vec(S.M, S.vec, lane);
      jaref_of();
      bool first = true;
#pragma unroll 1
      for (;;) {
        // ======== part A of the evaluation at the current point (mj_constraintUpdate): cost, row forces,
        // qfrc_constraint and the gradient.  The convergence test needs nothing else, so the Hessian of the point
        // that ends the solve is never assembled or factorised (one factorisation per iteration, not per evaluation).
        {
          double s = 0.0;
          __syncwarp();
#pragma unroll 1
          for (int r = lane; r < nefc; r += 32) {
            const double jar = S.rJaref[r], D = S.rD[r];
            const bool act = jar < 0;
            if (act) s += 0.5 * D * jar * jar;
            S.rowbuf[r] = act ? -D * jar : 0.0;
          }
          const double cost_c = warp_sum(s);
          gauss = warp_sum(lane < NV ? 0.5 * (Ma - qfs) * (qacc - qas) : 0.0);
          __syncwarp();
          // ---- qfrc_constraint = J' f: gather the row forces on the contact basis
#pragma unroll 1
          for (int t = lane; t < ncon * 6; t += 32) {
            const int c = t / 6, kk = t - c * 6;
            const double* f = S.rowbuf + nlim + 10 * c;
            double v;
            if (kk == 0) {
              v = 0.0;
#pragma unroll
              for (int e = 0; e < 10; e++) v += f[e];
            } else {
              v = S.cfric[c][kk - 1] * (f[2 * (kk - 1)] - f[2 * (kk - 1) + 1]);
            }
            S.jbv[c][kk] = v;
          }
          __syncwarp();
          qfc = 0.0;
          if (lane < NV) {
#pragma unroll 1
            for (int c = 0; c < ncon; c++) {
#pragma unroll
              for (int kk = 0; kk < 6; kk++) qfc += S.Jb[c][kk][lane] * S.jbv[c][kk];
            }
#pragma unroll 1
            for (int r = 0; r < nlim; r++)
              if (S.lim_dof[r] == lane) qfc += S.lim_sign[r] * S.rowbuf[r];
          }
          grad = (lane < NV) ? Ma - qfs - qfc : 0.0;
          cost = cost_c + gauss;
          PH(14);
        }
        if (!first) {
          const double improvement = scale * (oldcost - cost);
          const double gradient = scale * sqrt(warp_sum(grad * grad));
          niter++;
          if (improvement < tolerance || gradient < tolerance) break;
        }
        first = false;
        if (niter >= max_iter) break;

        // ======== part B: Hessian of the active set, factorisation, Newton direction
        {
          __syncwarp();
#pragma unroll 1
          for (int r = lane; r < nefc; r += 32) S.rowbuf[r] = (S.rJaref[r] < 0) ? S.rD[r] : 0.0;
          __syncwarp();
          // ---- per-contact weights of the active edges in the 6-row basis
#pragma unroll 1
          for (int t = lane; t < ncon * 5; t += 32) {
            const int c = t / 5, kk = 1 + (t - c * 5);
            const double* Da = S.rowbuf + nlim + 10 * c;
            const double dp = Da[2 * (kk - 1)], dm = Da[2 * (kk - 1) + 1], mu = S.cfric[c][kk - 1];
            S.W[c][kk] = mu * (dp - dm);
            S.W[c][5 + kk] = mu * mu * (dp + dm);
          }
          if (lane < ncon) {
            const double* Da = S.rowbuf + nlim + 10 * lane;
            double Wn = 0.0;
#pragma unroll
            for (int kk = 1; kk < 6; kk++) Wn += Da[2 * (kk - 1)] + Da[2 * (kk - 1) + 1];
            S.W[lane][0] = Wn;
          }
          __syncwarp();
          PH(5);
          // ---- H = M + J' Dact J from the contact basis (lower triangle), in two passes per chunk of 3 contacts:
          //   T_c = W_c Jb_c (6 x NV; W_c is the arrow-shaped 6x6 weight: Wnn, Wnk, diag Wkk), staged in the storage of H
          //   (dead until the end of the assembly: 3 x 6 x LD doubles is exactly NV x LD), then
          //   H_ij += sum_k Jb_c[k][i] T_c[k][j] with the six entries of a lane as independent accumulators.
          constexpr int NQ = (NTRI + 31) / 32;
          double acc[NQ];
          int oi[NQ], oj[NQ];
#pragma unroll
          for (int q = 0; q < NQ; q++) {
            const int t = min(lane + 32 * q, NTRI - 1);
            oi[q] = T->tri_i[t]; oj[q] = T->tri_j[t];   // L1-resident index tables (no per-env copy: 342 B of shared memory less)
            acc[q] = S.M[oi[q] * LD + oj[q]];
          }
          // diagonal of the active limit rows: lane d sums the rows of dof d (uniform loop, broadcast loads), added last
          double dd = 0.0;
#pragma unroll 4
          for (int r = 0; r < nlim; r++) {
            const int ld = S.lim_dof[r];
            const double rb = S.rowbuf[r];
            dd += (ld == lane) ? rb : 0.0;
          }
          PH(8);
          double (*Tc)[6][LD] = reinterpret_cast<double (*)[6][LD]>(S.H);
#pragma unroll 1
          for (int c0 = 0; c0 < ncon; c0 += 3) {
            const int nc = min(3, ncon - c0);
            __syncwarp();
#pragma unroll 1
            for (int e = lane; e < nc * NV; e += 32) {
              const int cc = e / NV, d = e - cc * NV;
              const double* Wc = S.W[c0 + cc];
              const double (*J)[LD] = S.Jb[c0 + cc];
              double jk[6];
#pragma unroll
              for (int kk = 0; kk < 6; kk++) jk[kk] = J[kk][d];
              double t0 = Wc[0] * jk[0];
#pragma unroll
              for (int kk = 1; kk < 6; kk++) {
                t0 += Wc[kk] * jk[kk];
                Tc[cc][kk][d] = Wc[kk] * jk[0] + Wc[5 + kk] * jk[kk];
              }
              Tc[cc][0][d] = t0;
            }
            __syncwarp();
#pragma unroll 1
            for (int cc = 0; cc < nc; cc++) {
              const double (*J)[LD] = S.Jb[c0 + cc];
#pragma unroll
              for (int kk = 0; kk < 6; kk++) {
#pragma unroll
                for (int q = 0; q < NQ; q++) acc[q] += J[kk][oi[q]] * Tc[cc][kk][oj[q]];
              }
            }
          }
          __syncwarp();
#pragma unroll
          for (int q = 0; q < NQ; q++)
            if (lane + 32 * q < NTRI) S.H[oi[q] * LD + oj[q]] = acc[q];
          __syncwarp();
          if (lane < NV) S.H[lane * LD + lane] += dd;
          PH(12);
          mgrad = chol_factor_solve(S.H, &S.chol, grad, lane);
          PH(13);
#ifdef HMJ_PHASE_PROF_CHOL2   // developer experiment: the same routine again, back to back (H holds L by now: timing only)
          { const double again = chol_factor_solve(S.H, &S.chol, grad, lane); if (again == 1.2345e300) mgrad = again; }
          PH(8);
#endif
        }
        search = -mgrad;

        // ======== exact line search along `search` (PrimalSearch)
        double alpha = 0.0;
        const double snorm = sqrt(warp_sum(lane < NV ? search * search : 0.0));
        if (snorm < HMJ_MINVAL) break;
        const double gtol = tolerance * ls_tolerance * snorm / scale;
        __syncwarp();
        if (lane < NV) S.vec[lane] = search;
        __syncwarp();
        Mv = mat_vec(S.M, S.vec, lane);
        basis_dot(S.Jb, S.vec, S.jbv, ncon, lane);
        const double qg0 = gauss;
        const double qg1 = warp_sum(lane < NV ? search * Ma - qfs * search : 0.0);
        const double qg2 = 0.5 * warp_sum(lane < NV ? search * Mv : 0.0);
#pragma unroll 1
        for (int r = lane; r < nefc; r += 32) S.rJv[r] = row_value(R, r);
        __syncwarp();
        PH(6);
        int lsiter = 0;
        auto EV = [&](LsPnt& p) { ls_eval(p, S.rD, S.rJaref, S.rJv, nefc, qg0, qg1, qg2, lane); lsiter++; };
        {
          LsPnt p0, p1, p2, pmid, p1next, p2next;
          p0.alpha = 0; EV(p0);
          p1.alpha = p0.alpha - p0.d0 / p0.d1; EV(p1);
          if (p0.cost < p1.cost) p1 = p0;
          bool found = false;
          if (fabs(p1.d0) < gtol) { alpha = p1.alpha; found = true; }
          if (!found) {
            const int dir = (p1.d0 < 0 ? +1 : -1);
            bool p2update = false;
            p2 = p1;
#pragma unroll 1
            while (p1.d0 * dir <= -gtol && lsiter < ls_iterations) {
              p2 = p1; p2update = true;
              p1.alpha -= p1.d0 / p1.d1; EV(p1);
              if (fabs(p1.d0) < gtol) { alpha = p1.alpha; found = true; break; }
            }
            if (!found && (lsiter >= ls_iterations || !p2update)) { alpha = p1.alpha; found = true; }
            if (!found) {
              p2next = p1;
              p1next.alpha = p1.alpha - p1.d0 / p1.d1; EV(p1next);
#pragma unroll 1
              while (lsiter < ls_iterations && !found) {
                pmid.alpha = 0.5 * (p1.alpha + p2.alpha); EV(pmid);
                // the three candidates stay in registers (c0..c2 are copies: p1next / p2next are re-evaluated below)
                const LsPnt c0 = p1next, c1 = p2next, c2 = pmid;
                double bestcost = 0, bestalpha = 0; bool have = false;
#define HMJ_LS_BEST(c) if (fabs(c.d0) < gtol && (!have || c.cost < bestcost)) { bestcost = c.cost; bestalpha = c.alpha; have = true; }
                HMJ_LS_BEST(c0) HMJ_LS_BEST(c1) HMJ_LS_BEST(c2)
#undef HMJ_LS_BEST
                if (have) { alpha = bestalpha; found = true; break; }
                int b1 = 0, b2 = 0;
#define HMJ_LS_BRACKET(p, b, c) if (p.d0 < 0 && c.d0 < 0 && p.d0 < c.d0) { p = c; b = 1; } else if (p.d0 > 0 && c.d0 > 0 && p.d0 > c.d0) { p = c; b = 2; }
                HMJ_LS_BRACKET(p1, b1, c0) HMJ_LS_BRACKET(p1, b1, c1) HMJ_LS_BRACKET(p1, b1, c2)
                if (b1) { p1next.alpha = p1.alpha - p1.d0 / p1.d1; EV(p1next); }
                HMJ_LS_BRACKET(p2, b2, c0) HMJ_LS_BRACKET(p2, b2, c1) HMJ_LS_BRACKET(p2, b2, c2)
#undef HMJ_LS_BRACKET
                if (b2) { p2next.alpha = p2.alpha - p2.d0 / p2.d1; EV(p2next); }
                if (!b1 && !b2) { alpha = pmid.alpha; found = true; break; }
              }
              if (!found) {
                if (p1.cost <= p2.cost && p1.cost < p0.cost) alpha = p1.alpha;
                else if (p2.cost <= p1.cost && p2.cost < p0.cost) alpha = p2.alpha;
                else alpha = 0.0;
              }
            }
          }
        }
        PH(7);
#ifdef HMJ_PHASE_PROF
        if (lane == 0) hmj_phase_env[env & 16383][15] += lsiter;
#endif
        if (alpha == 0.0) break;

        // ======== move to the new point
        if (lane < NV) { qacc += alpha * search; Ma += alpha * Mv; }
#pragma unroll 1
        for (int r = lane; r < nefc; r += 32) S.rJaref[r] += alpha * S.rJv[r];
        oldcost = cost;
      }
      }   // !INV
    }
  }

  PH(9);
  if constexpr (INV) {
    // ---------------- mj_inverse: qfrc_inverse = M qacc + qfrc_bias - qfrc_passive - qfrc_constraint
    //                                          = M qacc - qfrc_smooth + (qfrc_actuator + J' xfrc_applied) - qfrc_constraint
    double Mq = 0.0;
    if constexpr (MAXC > 0) {
      __syncwarp();
      if (lane < NV) S.vec[lane] = qacc;
      __syncwarp();
      Mq = mat_vec(S.M, S.vec, lane);
    }
    if (lane == 0) { B.ncon[env] = ncon; B.nefc[env] = nefc; B.niter[env] = 0; }
    if (lane < NV) {
      B.qfrc_constraint[(size_t)env * NV + lane] = qfc;
      B.qfrc_inverse[(size_t)env * NV + lane] = Mq - qfs + B.qfrc_ext[(size_t)env * NV + lane] - qfc;
    }
    if (status && lane == 0) atomicOr(&B.status[env], status);
    return;
  }
  // ---------------- results of mj_fwdConstraint
  if (lane == 0) { B.ncon[env] = ncon; B.nefc[env] = nefc; B.niter[env] = niter; }
  if (lane < NV) {
    B.qacc[(size_t)env * NV + lane] = qacc;
    B.qfrc_constraint[(size_t)env * NV + lane] = qfc;
    B.warmstart[(size_t)env * NV + lane] = qacc;
  }

  // ---------------- mj_checkAcc
  if (__any_sync(FULL, lane < NV && d_is_bad(qacc))) {
    status |= HMJ_STATUS_BADQACC;
    // mj_resetData: state back to qpos0 (MuJoCo then re-runs mj_forward; callers treat this as divergence)
    if (lane < HMJ_NQ) B.qpos[(size_t)env * HMJ_NQ + lane] = M.qpos0[lane];
    if (lane < NV) { B.qvel[(size_t)env * NV + lane] = 0.0; B.warmstart[(size_t)env * NV + lane] = 0.0; }
    if (lane == 0) { B.time[env] = 0.0; atomicOr(&B.status[env], status); }
    return;
  }
  if (status && lane == 0) atomicOr(&B.status[env], status);
  if (!integrate) return;

  PH(10);
  // ---------------- mj_Euler with implicit joint damping: (M + h diag(b)) a = qfrc_smooth + qfrc_constraint,
  // with the factor k_factor left in B.euler_factor
  __syncwarp();
  double qacc_e;
  if (T->tree_chol) {
    qacc_e = tree_solve_stored(B.euler_factor + (size_t)env * FAC_DOUBLES, &S.chol.col[0][0], qfs + qfc, lane);
  } else {
    const double* F = B.euler_factor + (size_t)env * FAC_DOUBLES;
    double fe[(NTRI + 31) / 32];
    const double fi = F[NTRI + (lane < NV ? lane : 0)];
#pragma unroll
    for (int k = 0; k < (NTRI + 31) / 32; k++) fe[k] = (lane + 32 * k < NTRI) ? F[lane + 32 * k] : 0.0;   // all loads in flight at once
#pragma unroll
    for (int k = 0; k < (NTRI + 31) / 32; k++) {
      const int t = lane + 32 * k;
      if (t < NTRI) S.H[T->tri_i[t] * LD + T->tri_j[t]] = fe[k];
    }
    if (lane < NV) S.chol.invd[lane] = fi;
    __syncwarp();
    qacc_e = chol_solve_only(S.H, S.chol.invd, qfs + qfc, lane);
  }
  PH(11);
  if (lane < NV) {
    qvel += h * qacc_e;
    B.qvel[(size_t)env * NV + lane] = qvel;
    S.vec[lane] = qvel;
  }
  __syncwarp();
  // mj_integratePos
  if (lane >= 6 && lane < NV) {
    const int j = M.dof_jntid[lane];
    double* qp = B.qpos + (size_t)env * HMJ_NQ + M.jnt_qposadr[j];
    *qp += h * qvel;
  } else if (lane < 3) {
    B.qpos[(size_t)env * HMJ_NQ + lane] += h * qvel;
  } else if (lane == 3) {
    double* q = B.qpos + (size_t)env * HMJ_NQ + 3;
    double quat[4] = {q[0], q[1], q[2], q[3]};
    double w[3] = {S.vec[3], S.vec[4], S.vec[5]}, qrot[4];
    const double angle = h * d_normalize3(w);
    d_axis_angle2quat(qrot, w, angle);
    d_normalize4(quat);
    d_mul_quat(quat, quat, qrot);
    q[0] = quat[0]; q[1] = quat[1]; q[2] = quat[2]; q[3] = quat[3];
  }
  if (lane == 0) B.time[env] += h;
}

template <int MAXC>
static void launch_class(const hmj_tables_t* tab, hmj_buffers_t buf, int integrate, cudaStream_t stream) {
  static bool configured[HMJ_MAX_DEVICES] = {false};   // function attributes are per device
  const size_t smem = sizeof(SolveSmem<MAXC>);
  const int dev = hmj_current_device();
  if (!configured[dev]) {
    cudaFuncSetAttribute(k_solve<MAXC, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k_solve<MAXC, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured[dev] = true;
  }
  const int self = (integrate & HMJ_SOLVE_SELFCLASS);
  if ((integrate & 3) == 2) k_solve<MAXC, true><<<buf.n, 32, smem, stream>>>(tab, buf, self);   // mj_inverse
  else k_solve<MAXC, false><<<buf.n, 32, smem, stream>>>(tab, buf, (integrate & 1) | self);
}

// Five launches: classification, then one solver launch per class.  With side streams the many-contact class and the
// unconstrained class run concurrently with the main class; otherwise the heavy class goes first.
void hmj_launch_solve(const hmj_tables_t* tab, hmj_buffers_t buf, int integrate, cudaStream_t stream,
                      const hmj_solve_ctx_t* ctx, bool zero) {
  if (ctx) cudaStreamWaitEvent(stream, ctx->fac_done, 0);   // k_factor ran on side[0] beside the collision kernels
  if (zero) cudaMemsetAsync(buf.solve_count, 0, 4 * sizeof(int32_t), stream);   // else: zeroed by this step's k_smooth
  if (buf.n <= SOLVE_SELFCLASS_MAX_N) integrate |= HMJ_SOLVE_SELFCLASS;   // the solver CTAs classify their own env
  else k_classify<<<(buf.n + 3) / 4, 128, 0, stream>>>(tab, buf, integrate == 2);
  if (!ctx) {
    launch_class<HMJ_MAXCON>(tab, buf, integrate, stream);
    launch_class<SOLVE_MID_CON>(tab, buf, integrate, stream);
    launch_class<SOLVE_SMALL_CON>(tab, buf, integrate, stream);
    launch_class<0>(tab, buf, integrate, stream);
    return;
  }
  cudaEventRecord(ctx->fork, stream);
  cudaStreamWaitEvent(ctx->side[0], ctx->fork, 0);
  cudaStreamWaitEvent(ctx->side[1], ctx->fork, 0);
  launch_class<HMJ_MAXCON>(tab, buf, integrate, ctx->side[0]);
  launch_class<SOLVE_MID_CON>(tab, buf, integrate, ctx->side[0]);
  launch_class<SOLVE_SMALL_CON>(tab, buf, integrate, stream);
  launch_class<0>(tab, buf, integrate, ctx->side[1]);
  cudaEventRecord(ctx->join[0], ctx->side[0]);
  cudaEventRecord(ctx->join[1], ctx->side[1]);
  cudaStreamWaitEvent(stream, ctx->join[0], 0);
  cudaStreamWaitEvent(stream, ctx->join[1], 0);
}

#ifdef HMJ_PHASE_PROF
extern "C" int hmj_debug_phase_cycles(unsigned long long* out, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, hmj_phase_cycles, sizeof(unsigned long long) * 16);
  if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(hmj_phase_cycles, z, sizeof(z)); }
  return 0;
}
extern "C" int hmj_debug_phase_env(unsigned int* out, int n, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, hmj_phase_env, sizeof(unsigned int) * 16 * (size_t)n);
  if (reset) { void* p = nullptr; cudaGetSymbolAddress(&p, hmj_phase_env); cudaMemset(p, 0, sizeof(unsigned int) * 16 * 16384); }
  return 0;
}
#endif
