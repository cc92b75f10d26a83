// hmj_solve.cu — kernel 3 of the step: contacts -> constraints -> Newton solver -> Euler integrator.
//
// Restates mj_fwdAcceleration's solve (mujoco.h:256), mj_makeConstraint (:358),
// mj_referenceConstraint (:367), mj_fwdConstraint with the Newton solver (:259), mj_checkAcc (:307)
// and mj_Euler with implicit joint damping (:262) for this model (MuJoCo 3.1.4 semantics, defaults of
// mjmodel.h:403-444: pyramidal cones, condim 6 -> 10 rows per contact, tolerance 1e-8, 100 iterations,
// exact line search with ls_tolerance 0.01 / 50 iterations, refsafe, warmstart).
//
// Mapping: one warp per environment, 4 warps per CTA.
//   * dense 18x18 M and Hessian factor in shared memory (row stride 19: conflict-free fp64 columns);
//     right-looking warp Cholesky, triangular solves in registers with shuffles;
//   * the 10 pyramid-edge rows of a contact are never materialised: only the 6 contact-frame basis rows
//     Jb[c][6][18] are stored (J_edge = Jn +- mu_k Jk), so J.v, J'.f and J'DJ all work on the basis;
//   * per-row scalars (D, aref, Jaref, Jv, line-search quadratics) live in the registers of the lane
//     that owns the row (row r -> lane r % 32), reductions are xor-butterfly shuffles, which give
//     bit-identical sums on all lanes, so all solver control flow is warp-uniform.
// Roofline: FP64 FMA pipe (J'DJ, Cholesky) and shuffle/issue latency (line search); per-env HBM/L2
// traffic is ~2.6 KB in (qM, cdof, slots, state) and ~0.6 KB out.
#include <cstdio>
#include "hmj_dev.cuh"

#define NV HMJ_NV
#ifndef SOLVE_MIN_CTAS
#define SOLVE_MIN_CTAS 8
#endif
#ifndef SOLVE_WARPS
#define SOLVE_WARPS 1           /* one environment per CTA: a slow Newton solve never holds idle neighbours resident */
#endif
#define LD 19                   /* padded row stride of the dense 18x18 matrices */
#define NROWREG 6               /* ceil(HMJ_MAXEFC / 32) rows per lane */
#define NLIMMAX 24
#define ST_QUAD 1

struct SolveSmem {
  double M[NV * LD];
  double H[NV * LD];
  double invd[NV + 1];
  double Jb[HMJ_MAXCON][6][LD];
  double jbv[HMJ_MAXCON][6];
  double rowbuf[HMJ_MAXEFC];
  double cpos[HMJ_MAXCON][3];
  double cframe[HMJ_MAXCON][9];
  double cfric[HMJ_MAXCON][5];
  double vec[NV + 1];           // vector being multiplied (search / qacc / qvel)
  double qfs[NV + 1];           // qfrc_smooth
  double qas[NV + 1];           // qacc_smooth
  double qfc[NV + 1];           // qfrc_constraint
  double com[4];
  int32_t cslot[HMJ_MAXCON];
  int32_t cb1[HMJ_MAXCON], cb2[HMJ_MAXCON];
  int32_t lim_dof[NLIMMAX];
  double lim_sign[NLIMMAX], lim_pos[NLIMMAX], lim_margin[NLIMMAX];
  int32_t lim_jnt[NLIMMAX];
};

// ---------------------------------------------------------------- dense helpers (warp cooperative)
// In-register right-looking Cholesky: lane i holds row i of the lower triangle (fully unrolled, static
// register indices); column j is scaled by rsqrt(d_j) computed redundantly on every lane and the rank-1
// update pulls l_kj from lane k with one shuffle per (j, k) pair: 153 shuffles + 153 FMAs, no shared
// memory traffic or barriers inside the factorisation.  Entries above the diagonal carry garbage.
__device__ __noinline__ void chol_factor(double* H, double* invd, const hmj_tables_t*, int lane) {
  const unsigned FULL = 0xffffffffu;
  const int row = lane < NV ? lane : NV - 1;
  double a[NV];
  __syncwarp();
#pragma unroll
  for (int k = 0; k < NV; k++) a[k] = H[row * LD + k];
#pragma unroll
  for (int j = 0; j < NV; j++) {
    double dj = __shfl_sync(FULL, a[j], j);
    if (dj < HMJ_MINVAL) dj = HMJ_MINVAL;
    const double inv = rsqrt(dj);
    const double lij = (lane == j) ? dj * inv : a[j] * inv;
    a[j] = lij;
    if (lane == j) invd[j] = inv;
#pragma unroll
    for (int k = j + 1; k < NV; k++) {
      const double lkj = __shfl_sync(FULL, lij, k);
      a[k] -= lij * lkj;
    }
  }
  if (lane < NV) {
#pragma unroll
    for (int k = 0; k < NV; k++)
      if (k <= lane) H[lane * LD + k] = a[k];
  }
  __syncwarp();
}

// x (one entry per lane, lanes >= NV ignored) <- H^-1 x with H = L L'
__device__ __noinline__ double chol_solve(const double* H, const double* invd, double x, int lane) {
  const unsigned FULL = 0xffffffffu;
  for (int k = 0; k < NV; k++) {
    const double xk = __shfl_sync(FULL, x, k) * invd[k];
    if (lane == k) x = xk;
    else if (lane > k && lane < NV) x -= H[lane * LD + k] * xk;
  }
  for (int k = NV - 1; k >= 0; k--) {
    const double xk = __shfl_sync(FULL, x, k) * invd[k];
    if (lane == k) x = xk;
    else if (lane < k) x -= H[k * LD + lane] * xk;
  }
  return x;
}

__device__ __forceinline__ double mat_vec(const double* A, const double* v, int lane) {
  double s = 0.0;
  if (lane < NV) {
#pragma unroll
    for (int j = 0; j < NV; j++) s += A[lane * LD + j] * v[j];
  }
  return s;
}

__device__ __forceinline__ void get_impedance(const double* solimp, double pos, double margin, double* imp) {
  if (solimp[0] == solimp[1] || solimp[2] <= HMJ_MINVAL) { *imp = 0.5 * (solimp[0] + solimp[1]); return; }
  double x = (pos - margin) / solimp[2];
  if (x < 0) x = -x;
  if (x >= 1 || x <= 0) { *imp = (x >= 1 ? solimp[1] : solimp[0]); return; }
  double y;
  if (solimp[4] == 1) {
    y = x;
  } else if (x <= solimp[3]) {
    double a = 1 / pow(solimp[3], solimp[4] - 1);
    y = a * pow(x, solimp[4]);
  } else {
    double b = 1 / pow(1 - solimp[3], solimp[4] - 1);
    y = 1 - b * pow(1 - x, solimp[4]);
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
#define PH(i) do { const long long t_ = clock64(); if (lane == 0) atomicAdd(&hmj_phase_cycles[i], (unsigned long long)(t_ - ph_t0)); ph_t0 = clock64(); } while (0)
#else
#define PH(i) do { } while (0)
#endif

__global__ void __launch_bounds__(32 * SOLVE_WARPS, SOLVE_MIN_CTAS)
k_solve(const hmj_tables_t* __restrict__ T, hmj_buffers_t B, int integrate) {
  extern __shared__ double smem_raw[];
  const int lane = threadIdx.x & 31;
  const int wib = threadIdx.x >> 5;
  const int env = blockIdx.x * SOLVE_WARPS + wib;
  if (env >= B.n) return;
  SolveSmem& S = reinterpret_cast<SolveSmem*>(smem_raw)[wib];
  const hmj_model_t& M = T->m;
  const hmj_params_t* P = B.params + env;
  const unsigned FULL = 0xffffffffu;
  const double h = P->timestep;
#ifdef HMJ_PHASE_PROF
  long long ph_t0 = clock64();
#endif

  // ---------------- load: dense M, qfrc_smooth, state
  for (int t = lane; t < NV * LD; t += 32) S.M[t] = 0.0;
  __syncwarp();
  for (int e = lane; e < HMJ_NM; e += 32) {
    const double v = B.qM[(size_t)env * HMJ_NM + e];
    const int i = T->M_row[e], j = T->M_col[e];
    S.M[i * LD + j] = v;
    S.M[j * LD + i] = v;
  }
  double qfs = 0, qvel = 0, qws = 0;
  if (lane < NV) {
    qfs = B.qfrc_smooth[(size_t)env * NV + lane];
    qvel = B.qvel[(size_t)env * NV + lane];
    qws = B.warmstart[(size_t)env * NV + lane];
    S.qfs[lane] = qfs;
  }
  if (lane < 3) S.com[lane] = B.com[(size_t)env * 3 + lane];
  __syncwarp();

  // ---------------- qacc_smooth = M^-1 qfrc_smooth (dense Cholesky instead of the sparse L'DL: same result to rounding)
  for (int t = lane; t < NV * LD; t += 32) S.H[t] = S.M[t];
  __syncwarp();
  chol_factor(S.H, S.invd, T, lane);
  const double qas = chol_solve(S.H, S.invd, qfs, lane);
  if (lane < NV) { S.qas[lane] = qas; B.qacc_smooth[(size_t)env * NV + lane] = qas; }

  PH(1);
  // ---------------- contacts: ordered compaction of the narrowphase slots (slot order == MuJoCo contact order)
  int ncon = 0;
  int status = 0;
  {
    const double* slots = B.slots + (size_t)env * HMJ_NSLOT * HMJ_SLOT_DOUBLES;
    for (int base = 0; base < HMJ_NSLOT; base += 32) {
      const int s = base + lane;
      const bool valid = (s < HMJ_NSLOT) && (slots[(size_t)s * HMJ_SLOT_DOUBLES + 7] != 0.0);
      const unsigned bal = __ballot_sync(FULL, valid);
      const int idx = ncon + __popc(bal & ((1u << lane) - 1u));
      if (valid && idx < HMJ_MAXCON) S.cslot[idx] = s;
      ncon += __popc(bal);
    }
    if (ncon > HMJ_MAXCON) { ncon = HMJ_MAXCON; status |= HMJ_STATUS_CONTACTFULL; }
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
      S.lim_margin[before] = margin; S.lim_jnt[before] = j;
    }
    if (act1) {
      const int r = before + (act0 ? 1 : 0);
      S.lim_dof[r] = M.jnt_dofadr[j]; S.lim_sign[r] = -1.0; S.lim_pos[r] = dist1;
      S.lim_margin[r] = margin; S.lim_jnt[r] = j;
    }
    nlim = __popc(b0) + __popc(b1);
  }
  const int nefc = nlim + 10 * ncon;
  __syncwarp();
  if (lane == 0) { B.ncon[env] = ncon; B.nefc[env] = nefc; }

  double qacc = qas;  // lanes < NV
  double qfc = 0.0;
  int niter = 0;

  if (nefc > 0) {
    PH(2);
    // ---------------- contact-frame Jacobian basis Jb[c][k][d] (mj_jacDifPair rotated into the contact frame)
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

    // ---------------- per-row constants in the owning lane's registers
    double rD[NROWREG], rAref[NROWREG], rMu[NROWREG], rJaref[NROWREG], rJv[NROWREG];
    double rq0[NROWREG], rq1[NROWREG], rq2[NROWREG];
    int rInfo[NROWREG];  // limit: dof | 0x100 ; contact: c*8 + k
#pragma unroll
    for (int k = 0; k < NROWREG; k++) { rD[k] = 0; rAref[k] = 0; rMu[k] = 0; rJaref[k] = 0; rJv[k] = 0; rq0[k] = rq1[k] = rq2[k] = 0; rInfo[k] = 0; }

    // basis products with a vector held in S.vec
    auto basis_dot = [&]() {
      for (int t = lane; t < ncon * 6; t += 32) {
        const int c = t / 6, kk = t - c * 6;
        double s = 0.0;
#pragma unroll
        for (int d = 0; d < NV; d++) s += S.Jb[c][kk][d] * S.vec[d];
        S.jbv[c][kk] = s;
      }
      __syncwarp();
    };
    auto row_value = [&](int k) -> double {  // (J v)_r for the row in register slot k, after basis_dot()
      const int info = rInfo[k];
      if (info & 0x100) return rMu[k] * S.vec[info & 0xff];
      const int c = info >> 3, kk = info & 7;
      return S.jbv[c][0] + rMu[k] * S.jbv[c][kk];
    };

    // efc_vel = J qvel  -> aref ; D
    if (lane < NV) S.vec[lane] = qvel;
    __syncwarp();
    basis_dot();
#pragma unroll
    for (int k = 0; k < NROWREG; k++) {
      const int r = lane + 32 * k;
      if (r < nefc) {
        double R, KI, Bc, pos, margin;
        if (r < nlim) {
          const int j = S.lim_jnt[r];
          rInfo[k] = 0x100 | S.lim_dof[r];
          rMu[k] = S.lim_sign[r];
          pos = S.lim_pos[r]; margin = S.lim_margin[r];
          row_impedance(M.jnt_solref[j], M.jnt_solimp[j], pos, margin, M.dof_invweight0[S.lim_dof[r]], h, &R, &KI, &Bc);
        } else {
          const int c = (r - nlim) / 10, e = (r - nlim) - 10 * c;
          const int kk = 1 + (e >> 1);
          rInfo[k] = (c << 3) | kk;
          rMu[k] = (e & 1) ? -S.cfric[c][kk - 1] : S.cfric[c][kk - 1];
          const hmj_contact_t* C = B.contact + (size_t)env * HMJ_MAXCON + c;
          pos = C->dist; margin = 0.0;
          const int b1 = S.cb1[c], b2 = S.cb2[c];
          const double tran = M.body_invweight0[b1][0] + M.body_invweight0[b2][0];
          const double mu0 = S.cfric[c][0];
          // all edges share Rpy = 2 mu0^2 R(first edge); first edge: diagApprox = tran + mu0^2 tran
          double R0;
          row_impedance(C->solref, C->solimp, pos, margin, tran + mu0 * mu0 * tran, h, &R0, &KI, &Bc);
          R = 2 * mu0 * mu0 * R0;
        }
        rD[k] = 1 / R;
        const double vel = row_value(k);
        rAref[k] = -Bc * vel - KI * (pos - margin);
#ifdef HMJ_TRACE
        if (env == HMJ_TRACE) printf("GPU row %d D %.17g aref %.17g vel %.17g pos %.17g\n", r, rD[k], rAref[k], vel, pos);
#endif
      }
    }
    __syncwarp();

    // helpers that need the row registers
    auto jaref_of = [&](double* out) {  // out[k] = (J * S.vec)_r - aref_r
      basis_dot();
#pragma unroll
      for (int k = 0; k < NROWREG; k++) {
        const int r = lane + 32 * k;
        out[k] = (r < nefc) ? row_value(k) - rAref[k] : 0.0;
      }
    };
    auto constraint_cost = [&](const double* jar) -> double {
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < NROWREG; k++) {
        const int r = lane + 32 * k;
        if (r < nefc && jar[k] < 0) s += 0.5 * rD[k] * jar[k] * jar[k];
      }
      return warp_sum(s);
    };

    PH(3);
    // ---------------- warmstart: best of (qacc_warmstart, qacc_smooth)
    {
      double jar[NROWREG];
      if (lane < NV) S.vec[lane] = qws;
      __syncwarp();
      jaref_of(jar);
      double cost_ws = constraint_cost(jar);
      const double Ma_ws = mat_vec(S.M, S.vec, lane);
      cost_ws += warp_sum(lane < NV ? 0.5 * (Ma_ws - qfs) * (qws - qas) : 0.0);
      __syncwarp();
      if (lane < NV) S.vec[lane] = qas;
      __syncwarp();
      jaref_of(jar);
      const double cost_sm = constraint_cost(jar);
      qacc = (cost_ws > cost_sm) ? qas : qws;
      __syncwarp();
    }

    PH(4);
    // ---------------- Newton iterations (mj_solNewton, primal)
    const double scale = 1.0 / (M.meaninertia * (double)NV);
    const double tolerance = M.tolerance;
    double Ma, gauss = 0.0, grad = 0.0, mgrad = 0.0;

    // Everything that depends on the current point: active set, cost, Hessian factor,
    // qfrc_constraint, gradient and Newton direction.  Returns the total cost.
    auto evaluate = [&]() -> double {
      // ---- mj_constraintUpdate: cost, active set, forces
      double s = 0.0;
      double rF[NROWREG];
#pragma unroll
      for (int k = 0; k < NROWREG; k++) {
        const int r = lane + 32 * k;
        rF[k] = 0.0;
        if (r < nefc) {
          const double jar = rJaref[k];
          const bool act = jar < 0;
          if (act) { s += 0.5 * rD[k] * jar * jar; rF[k] = -rD[k] * jar; }
          S.rowbuf[r] = act ? rD[k] : 0.0;
        }
      }
      const double cost_c = warp_sum(s);
      gauss = warp_sum(lane < NV ? 0.5 * (Ma - qfs) * (qacc - qas) : 0.0);
      __syncwarp();

      // ---- H = M + J' Dact J from the contact basis (lower triangle), then Cholesky
      for (int t = lane; t < NV * (NV + 1) / 2; t += 32) {
        const int i = T->tri_i[t], j = T->tri_j[t];
        double acc = S.M[i * LD + j];
        for (int c = 0; c < ncon; c++) {
          const double* Da = S.rowbuf + nlim + 10 * c;
          const double* fr = S.cfric[c];
          const double Jni = S.Jb[c][0][i], Jnj = S.Jb[c][0][j];
          double Wn = 0.0, tj = 0.0, ti = 0.0;
#pragma unroll
          for (int kk = 1; kk < 6; kk++) {
            const double dp = Da[2 * (kk - 1)], dm = Da[2 * (kk - 1) + 1];
            const double mu = fr[kk - 1];
            const double wk = mu * (dp - dm), wkk = mu * mu * (dp + dm);
            const double Jki = S.Jb[c][kk][i], Jkj = S.Jb[c][kk][j];
            Wn += dp + dm;
            tj += wk * Jkj;
            ti += Jki * (wk * Jnj + wkk * Jkj);
          }
          acc += Jni * (Wn * Jnj + tj) + ti;
        }
        if (i == j) {
          for (int r = 0; r < nlim; r++)
            if (S.lim_dof[r] == i) acc += S.rowbuf[r];
        }
        S.H[i * LD + j] = acc;
      }
      __syncwarp();
      PH(12);
      chol_factor(S.H, S.invd, T, lane);
      PH(13);

      // ---- qfrc_constraint = J' f: publish row forces, gather on the contact basis
#pragma unroll
      for (int k = 0; k < NROWREG; k++) {
        const int r = lane + 32 * k;
        if (r < nefc) S.rowbuf[r] = rF[k];
      }
      __syncwarp();
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
        for (int c = 0; c < ncon; c++) {
#pragma unroll
          for (int kk = 0; kk < 6; kk++) qfc += S.Jb[c][kk][lane] * S.jbv[c][kk];
        }
        for (int r = 0; r < nlim; r++)
          if (S.lim_dof[r] == lane) qfc += S.lim_sign[r] * S.rowbuf[r];
      }
      __syncwarp();

      // ---- gradient and Newton direction
      grad = (lane < NV) ? Ma - qfs - qfc : 0.0;
      PH(14);
      mgrad = chol_solve(S.H, S.invd, grad, lane);
      return cost_c + gauss;
    };

    if (lane < NV) S.vec[lane] = qacc;
    __syncwarp();
    Ma = mat_vec(S.M, S.vec, lane);
    jaref_of(rJaref);
    double cost = evaluate();
    PH(5);
    double search = -mgrad;

    while (niter < M.iterations) {
      // ---- exact line search along `search` (PrimalSearch)
      double alpha = 0.0;
      const double snorm = sqrt(warp_sum(lane < NV ? search * search : 0.0));
      if (snorm < HMJ_MINVAL) break;
      const double gtol = tolerance * M.ls_tolerance * snorm / scale;
      __syncwarp();
      if (lane < NV) S.vec[lane] = search;
      __syncwarp();
      const double Mv = mat_vec(S.M, S.vec, lane);
      basis_dot();
      const double qg0 = gauss;
      const double qg1 = warp_sum(lane < NV ? search * Ma - qfs * search : 0.0);
      const double qg2 = 0.5 * warp_sum(lane < NV ? search * Mv : 0.0);
#pragma unroll
      for (int k = 0; k < NROWREG; k++) {
        const int r = lane + 32 * k;
        if (r < nefc) {
          rJv[k] = row_value(k);
          const double DJ0 = rD[k] * rJaref[k];
          rq0[k] = 0.5 * (rJaref[k] * DJ0);
          rq1[k] = rJv[k] * DJ0;
          rq2[k] = 0.5 * (rJv[k] * rD[k] * rJv[k]);
        }
      }
      PH(6);
      int lsiter = 0;
      auto ls_eval = [&](LsPnt& p) {
        double a0 = 0, a1 = 0, a2 = 0;
#pragma unroll
        for (int k = 0; k < NROWREG; k++) {
          const int r = lane + 32 * k;
          if (r < nefc && (rJaref[k] + p.alpha * rJv[k]) < 0) { a0 += rq0[k]; a1 += rq1[k]; a2 += rq2[k]; }
        }
        const double q0 = qg0 + warp_sum(a0), q1 = qg1 + warp_sum(a1), q2 = qg2 + warp_sum(a2);
        p.cost = p.alpha * p.alpha * q2 + p.alpha * q1 + q0;
        p.d0 = 2 * p.alpha * q2 + q1;
        p.d1 = 2 * q2;
        if (p.d1 <= 0) p.d1 = HMJ_MINVAL;
        lsiter++;
      };
      {
        const int ls_iterations = M.ls_iterations;
        LsPnt p0, p1, p2, pmid, p1next, p2next;
        p0.alpha = 0; ls_eval(p0);
        p1.alpha = p0.alpha - p0.d0 / p0.d1; ls_eval(p1);
        if (p0.cost < p1.cost) p1 = p0;
        bool found = false;
        if (fabs(p1.d0) < gtol) { alpha = p1.alpha; found = true; }
        if (!found) {
          const int dir = (p1.d0 < 0 ? +1 : -1);
          bool p2update = false;
          p2 = p1;
          while (p1.d0 * dir <= -gtol && lsiter < ls_iterations) {
            p2 = p1; p2update = true;
            p1.alpha -= p1.d0 / p1.d1; ls_eval(p1);
            if (fabs(p1.d0) < gtol) { alpha = p1.alpha; found = true; break; }
          }
          if (!found && (lsiter >= ls_iterations || !p2update)) { alpha = p1.alpha; found = true; }
          if (!found) {
            p2next = p1;
            p1next.alpha = p1.alpha - p1.d0 / p1.d1; ls_eval(p1next);
            while (lsiter < ls_iterations && !found) {
              pmid.alpha = 0.5 * (p1.alpha + p2.alpha); ls_eval(pmid);
              const LsPnt cand[3] = {p1next, p2next, pmid};
              double bestcost = 0; int bestind = -1;
              for (int i = 0; i < 3; i++)
                if (fabs(cand[i].d0) < gtol && (bestind == -1 || cand[i].cost < bestcost)) { bestcost = cand[i].cost; bestind = i; }
              if (bestind >= 0) { alpha = cand[bestind].alpha; found = true; break; }
              int b1 = 0, b2 = 0;
              for (int i = 0; i < 3; i++) {
                if (p1.d0 < 0 && cand[i].d0 < 0 && p1.d0 < cand[i].d0) { p1 = cand[i]; b1 = 1; }
                else if (p1.d0 > 0 && cand[i].d0 > 0 && p1.d0 > cand[i].d0) { p1 = cand[i]; b1 = 2; }
              }
              if (b1) { p1next.alpha = p1.alpha - p1.d0 / p1.d1; ls_eval(p1next); }
              for (int i = 0; i < 3; i++) {
                if (p2.d0 < 0 && cand[i].d0 < 0 && p2.d0 < cand[i].d0) { p2 = cand[i]; b2 = 1; }
                else if (p2.d0 > 0 && cand[i].d0 > 0 && p2.d0 > cand[i].d0) { p2 = cand[i]; b2 = 2; }
              }
              if (b2) { p2next.alpha = p2.alpha - p2.d0 / p2.d1; ls_eval(p2next); }
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
#ifdef HMJ_TRACE
      if (env == HMJ_TRACE && lane == 0) printf("GPU it %d alpha %.17g lsiter %d snorm %.17g gtol %.6g cost %.17g\n", niter, alpha, lsiter, snorm, gtol, cost);
#endif
      PH(7);
      if (alpha == 0.0) break;

      // ---- move to the new point, re-evaluate, test convergence
      if (lane < NV) { qacc += alpha * search; Ma += alpha * Mv; }
#pragma unroll
      for (int k = 0; k < NROWREG; k++) rJaref[k] += alpha * rJv[k];
      const double oldcost = cost;
      cost = evaluate();
      PH(8);
      const double improvement = scale * (oldcost - cost);
      const double gradient = scale * sqrt(warp_sum(grad * grad));
      niter++;
#ifdef HMJ_TRACE
      if (env == HMJ_TRACE && lane == 0) printf("GPU    -> cost %.17g improvement %.6g gradient %.6g\n", cost, improvement, gradient);
#endif
      if (improvement < tolerance || gradient < tolerance) break;
      search = -mgrad;
    }
  }

  PH(9);
  // ---------------- results of mj_fwdConstraint
  if (lane < NV) {
    B.qacc[(size_t)env * NV + lane] = qacc;
    B.qfrc_constraint[(size_t)env * NV + lane] = qfc;
    B.warmstart[(size_t)env * NV + lane] = qacc;
  }
  if (lane == 0) B.niter[env] = niter;

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
  // ---------------- mj_Euler with implicit joint damping: (M + h diag(b)) a = qfrc_smooth + qfrc_constraint
  __syncwarp();
  for (int t = lane; t < NV * LD; t += 32) S.H[t] = S.M[t];
  __syncwarp();
  if (lane < NV) S.H[lane * LD + lane] += h * M.dof_damping[lane];
  __syncwarp();
  chol_factor(S.H, S.invd, T, lane);
  const double qacc_e = chol_solve(S.H, S.invd, qfs + qfc, lane);
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

void hmj_launch_solve(const hmj_tables_t* tab, hmj_buffers_t buf, int integrate, cudaStream_t stream) {
  static bool configured = false;
  const size_t smem = sizeof(SolveSmem) * SOLVE_WARPS;
  if (!configured) {
    cudaFuncSetAttribute(k_solve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = true;
  }
  const int grid = (buf.n + SOLVE_WARPS - 1) / SOLVE_WARPS;
  k_solve<<<grid, 32 * SOLVE_WARPS, smem, stream>>>(tab, buf, integrate);
}

#ifdef HMJ_PHASE_PROF
extern "C" int hmj_debug_phase_cycles(unsigned long long* out, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, hmj_phase_cycles, sizeof(unsigned long long) * 16);
  if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(hmj_phase_cycles, z, sizeof(z)); }
  return 0;
}
#endif
