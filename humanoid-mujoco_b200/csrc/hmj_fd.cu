// hmj_fd.cu — finite-differenced transition matrices for all knots of a trajectory in ONE batched step.
//
// mjpc's ModelDerivatives::Compute (mjpc/planners/model_derivatives.cc:51-96) schedules one mjd_transitionFD per knot on
// its thread pool; each of those is 1 + 48 (forward) or 1 + 96 (centred) serial mj_steps of the same state with one
// coordinate nudged.  Here every nudged copy is an environment of the batch: T knots x 97 slots are set up by k_fd_setup,
// advanced by ONE pass of the ordinary step kernels (k_kin ... k_solve, planner sensors on), and differenced by k_fd_diff.
//   slot 0: the knot itself; slot 1 + 2 c + s: column c (0..17 position tangent, 18..35 velocity, 36..47 control),
//   s = 0 nudged by +eps, s = 1 by -eps.  Positions are nudged with mj_integratePos (quaternion exponential for the
//   root), controls forward where the control range allows and backward otherwise (both when centred), and every copy
//   starts from the knot's warm start and time — mjd_stepFD's rules.
// Outputs (row-major, per knot): A [36][36] = dy/dx, B [36][12] = dy/du with y the next state in tangent coordinates
// (mj_differentiatePos), C [28][36] = ds/dx, D [28][12] = ds/du with s the 28 task residuals of that step's forward pass.
// oracle/hmj_oracle_fd.c is the CPU restatement (checker).  Streaming kernels, thread per slot / per column.
#include "hmj_dev.cuh"

#define FD_SLOTS 97
#define FD_NDX (2 * HMJ_NV)

__global__ void k_fd_setup(const hmj_tables_t* __restrict__ T, hmj_buffers_t B, int nknot, const double* __restrict__ x,
                           const double* __restrict__ u, const double* __restrict__ time, const double* __restrict__ ws,
                           double eps, int centered, int32_t* __restrict__ valid) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nknot * FD_SLOTS) return;
  const hmj_model_t& M = T->m;
  const int t = e / FD_SLOTS, s = e - t * FD_SLOTS;
  double q[HMJ_NQ], v[HMJ_NV], c[HMJ_NU];
  for (int k = 0; k < HMJ_NQ; k++) q[k] = x[(size_t)t * HMJ_NSTATE + k];
  for (int k = 0; k < HMJ_NV; k++) v[k] = x[(size_t)t * HMJ_NSTATE + HMJ_NQ + k];
  for (int k = 0; k < HMJ_NU; k++) c[k] = u[(size_t)t * HMJ_NU + k];
  int ok = 1;
  if (s > 0) {
    const int col = (s - 1) >> 1;
    const bool minus = (s - 1) & 1;
    const double d = minus ? -eps : eps;
    if (col < HMJ_NV) {   // mj_integratePos along tangent coordinate `col`
      if (col < 3) {
        q[col] += d;
      } else if (col < 6) {
        double axis[3] = {0, 0, 0}, qrot[4], quat[4] = {q[3], q[4], q[5], q[6]};
        axis[col - 3] = d < 0 ? -1.0 : 1.0;
        d_axis_angle2quat(qrot, axis, fabs(d));
        d_normalize4(quat);
        d_mul_quat(quat, quat, qrot);
        q[3] = quat[0]; q[4] = quat[1]; q[5] = quat[2]; q[6] = quat[3];
      } else {
        q[M.jnt_qposadr[M.dof_jntid[col]]] += d;
      }
      ok = minus ? (centered != 0) : 1;
    } else if (col < FD_NDX) {
      v[col - HMJ_NV] += d;
      ok = minus ? (centered != 0) : 1;
    } else {
      const int i = col - FD_NDX;
      const double lo = M.act_ctrlrange[i][0], hi = M.act_ctrlrange[i][1];
      const bool fwd = (c[i] + eps <= hi) && (c[i] + eps >= lo);
      const bool back = (centered || !fwd) && (c[i] - eps >= lo) && (c[i] - eps <= hi);
      ok = minus ? back : fwd;
      c[i] += d;
    }
  }
  for (int k = 0; k < HMJ_NQ; k++) B.qpos[(size_t)e * HMJ_NQ + k] = q[k];
  for (int k = 0; k < HMJ_NV; k++) {
    B.qvel[(size_t)e * HMJ_NV + k] = v[k];
    B.warmstart[(size_t)e * HMJ_NV + k] = ws ? ws[(size_t)t * HMJ_NV + k] : 0.0;
  }
  for (int k = 0; k < HMJ_NU; k++) B.ctrl[(size_t)e * HMJ_NU + k] = c[k];
  for (int k = 0; k < 6; k++) B.xfrc[(size_t)e * 6 + k] = 0.0;
  B.xfrc_body[e] = 0;
  B.time[e] = time[t];
  B.status[e] = 0;
  B.cost_sum[e] = 0.0;
  valid[e] = ok;
}

// (q2 (-) q1) / h in tangent coordinates (mj_differentiatePos)
__device__ __forceinline__ void fd_diff_pos(const hmj_model_t& M, double* dq, double h, const double* q1, const double* q2) {
  for (int k = 0; k < 3; k++) dq[k] = (q2[k] - q1[k]) / h;
  {
    const double qneg[4] = {q1[3], -q1[4], -q1[5], -q1[6]};
    double qdif[4];
    d_mul_quat(qdif, qneg, q2 + 3);
    double axis[3] = {qdif[1], qdif[2], qdif[3]};
    const double sn = d_normalize3(axis);
    double speed = 2 * atan2(sn, qdif[0]);
    if (speed > HMJ_PI) speed -= 2 * HMJ_PI;
    speed /= h;
    for (int k = 0; k < 3; k++) dq[3 + k] = axis[k] * speed;
  }
  for (int d = 6; d < HMJ_NV; d++) {
    const int a = M.jnt_qposadr[M.dof_jntid[d]];
    dq[d] = (q2[a] - q1[a]) / h;
  }
}

// thread per (knot, column): clampedStateDiff / clampedDiff of mjd_stepFD, written transposed into A/B/C/D
__global__ void k_fd_diff(const hmj_tables_t* __restrict__ T, hmj_buffers_t B, int nknot, const int32_t* __restrict__ valid,
                          double eps, double* __restrict__ A, double* __restrict__ Bm, double* __restrict__ C,
                          double* __restrict__ D) {
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= nknot * 48) return;
  const hmj_model_t& M = T->m;
  const int t = id / 48, col = id - t * 48;
  const int e0 = t * FD_SLOTS, ep = e0 + 1 + 2 * col, em = ep + 1;
  const bool hp = valid[ep] != 0, hm = valid[em] != 0;
  const int ea = hp ? ep : e0, eb = hm ? em : e0;
  const double h = (hp && hm) ? 2 * eps : eps;
  double dy[FD_NDX];
  if (!hp && !hm) {
    for (int r = 0; r < FD_NDX; r++) dy[r] = 0.0;
  } else {
    fd_diff_pos(M, dy, h, B.qpos + (size_t)eb * HMJ_NQ, B.qpos + (size_t)ea * HMJ_NQ);
    for (int r = 0; r < HMJ_NV; r++) dy[HMJ_NV + r] = (B.qvel[(size_t)ea * HMJ_NV + r] - B.qvel[(size_t)eb * HMJ_NV + r]) / h;
  }
  const bool isu = col >= FD_NDX;
  double* Y = isu ? Bm + (size_t)t * FD_NDX * HMJ_NU : A + (size_t)t * FD_NDX * FD_NDX;
  double* S = isu ? D + (size_t)t * HMJ_NRESIDUAL * HMJ_NU : C + (size_t)t * HMJ_NRESIDUAL * FD_NDX;
  const int ld = isu ? HMJ_NU : FD_NDX, cc = isu ? col - FD_NDX : col;
  for (int r = 0; r < FD_NDX; r++) Y[r * ld + cc] = dy[r];
  for (int r = 0; r < HMJ_NRESIDUAL; r++) {
    const double a = B.residual[(size_t)ea * HMJ_NRESIDUAL + r], b = B.residual[(size_t)eb * HMJ_NRESIDUAL + r];
    S[r * ld + cc] = (hp || hm) ? (a - b) / h : 0.0;
  }
}

void hmj_launch_fd_setup(const hmj_tables_t* tab, hmj_buffers_t view, int nknot, const double* x_dev, const double* u_dev,
                         const double* time_dev, const double* ws_dev, double eps, int centered, int32_t* valid_dev,
                         cudaStream_t stream) {
  const int n = nknot * FD_SLOTS;
  k_fd_setup<<<(n + 127) / 128, 128, 0, stream>>>(tab, view, nknot, x_dev, u_dev, time_dev, ws_dev, eps, centered, valid_dev);
}
void hmj_launch_fd_diff(const hmj_tables_t* tab, hmj_buffers_t view, int nknot, const int32_t* valid_dev, double eps, double* A,
                        double* B, double* C, double* D, cudaStream_t stream) {
  const int n = nknot * 48;
  k_fd_diff<<<(n + 127) / 128, 128, 0, stream>>>(tab, view, nknot, valid_dev, eps, A, B, C, D);
}
