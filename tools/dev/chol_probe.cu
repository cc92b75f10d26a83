// Developer probe: cycles of chol_factor_solve (both variants) alone and under SM-level contention.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I humanoid-mujoco_b200/csrc -I include -o tools/dev/chol_probe tools/dev/chol_probe.cu
#include "hmj_solve.cu"
template <int VAR>
__global__ void __launch_bounds__(32) k_probe(long long* cyc, double* out, int reps) {
  extern __shared__ double sm[];
  double* H = sm;
  CholScratch* C = reinterpret_cast<CholScratch*>(sm + NV * LD + 8);
  const int lane = threadIdx.x;
  double x = 1.0 + lane;
  long long t = 0;
  for (int r = 0; r < reps; r++) {
    for (int e = lane; e < NV * LD; e += 32) { const int i = e / LD, j = e % LD; H[e] = (i == j) ? 20.0 + i : 1.0 / (1 + i + j); }
    __syncwarp();
    const long long t0 = clock64();
    x = (VAR == 0 ? chol_factor_solve(H, C, x, lane) : chol_factor_solve_shfl(H, C, x, lane)) + 1.0;
    t += clock64() - t0;
  }
  if (lane == 0) cyc[blockIdx.x] = t / reps;
  out[blockIdx.x * 32 + lane] = x;
}
int main() {
  long long* cyc; double* out; cudaMalloc(&cyc, 8 * 65536); cudaMalloc(&out, 8 * 32 * 65536);
  const size_t smem = (NV * LD + 8) * 8 + sizeof(CholScratch);
  for (int var = 0; var < 2; var++)
    for (int grid : {1, 148, 148 * 4, 148 * 8, 148 * 16, 148 * 24}) {
      if (var == 0) k_probe<0><<<grid, 32, smem>>>(cyc, out, 20); else k_probe<1><<<grid, 32, smem>>>(cyc, out, 20);
      cudaDeviceSynchronize();
      static long long h[65536]; cudaMemcpy(h, cyc, 8 * grid, cudaMemcpyDeviceToHost);
      double s = 0; long long mx = 0; for (int i = 0; i < grid; i++) { s += h[i]; if (h[i] > mx) mx = h[i]; }
      printf("variant %s  warps/SM %3d : mean %8.0f max %8lld cycles per factor+solve\n", var ? "shfl-unrolled" : "rotating-rolled", grid / 148, s / grid, mx);
    }
  // agreement of the two variants on one solve
  static double r0[32], r1[32];
  k_probe<0><<<1, 32, smem>>>(cyc, out, 1); cudaMemcpy(r0, out, sizeof(r0), cudaMemcpyDeviceToHost);
  k_probe<1><<<1, 32, smem>>>(cyc, out, 1); cudaMemcpy(r1, out, sizeof(r1), cudaMemcpyDeviceToHost);
  double worst = 0; for (int i = 0; i < NV; i++) { double d = r0[i] - r1[i]; if (d < 0) d = -d; if (d > worst) worst = d; }
  printf("max |x_rolled - x_unrolled| = %.3e   x[0] = %.17g\n", worst, r0[0]);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
