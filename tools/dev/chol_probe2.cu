// Developer probe 2: chol_factor_solve with (a) the instruction caches thrashed between calls, (b) a wide-range matrix.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I humanoid-mujoco_b200/csrc -I include -o tools/dev/chol_probe2 tools/dev/chol_probe2.cu
#include "hmj_solve.cu"
__device__ __noinline__ double junk(double y) {
#pragma unroll
  for (int i = 0; i < 6000; i++) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(y) : "d"(1.0000001), "d"(1e-9));
  return y;
}
template <int MODE>
__global__ void __launch_bounds__(32) k_probe(long long* cyc, double* out, int reps) {
  extern __shared__ double sm[];
  double* H = sm;
  CholScratch* C = reinterpret_cast<CholScratch*>(sm + NV * LD + 8);
  const int lane = threadIdx.x;
  double x = 1.0 + lane;
  long long t = 0;
  for (int r = 0; r < reps; r++) {
    for (int e = lane; e < NV * LD; e += 32) {
      const int i = e / LD, j = e % LD;
      if (MODE & 2) H[e] = (i == j) ? 3.0e4 * (1 + i) + 1e-3 : 100.0 / (1 + i + j);
      else H[e] = (i == j) ? 20.0 + i : 1.0 / (1 + i + j);
    }
    if (MODE & 1) x = junk(x) * 1e-30 + 1.0 + lane;
    __syncwarp();
    const long long t0 = clock64();
    x = chol_factor_solve(H, C, x, lane) + 1.0;
    t += clock64() - t0;
  }
  if (lane == 0) cyc[blockIdx.x] = t / reps;
  out[blockIdx.x * 32 + lane] = x;
}
int main() {
  long long* cyc; double* out; cudaMalloc(&cyc, 8 * 65536); cudaMalloc(&out, 8 * 32 * 65536);
  const size_t smem = (NV * LD + 8) * 8 + sizeof(CholScratch);
  for (int mode = 0; mode < 4; mode++)
    for (int grid : {1, 148, 148 * 4}) {
      if (mode == 0) k_probe<0><<<grid, 32, smem>>>(cyc, out, 20);
      if (mode == 1) k_probe<1><<<grid, 32, smem>>>(cyc, out, 20);
      if (mode == 2) k_probe<2><<<grid, 32, smem>>>(cyc, out, 20);
      if (mode == 3) k_probe<3><<<grid, 32, smem>>>(cyc, out, 20);
      cudaDeviceSynchronize();
      static long long h[65536]; cudaMemcpy(h, cyc, 8 * grid, cudaMemcpyDeviceToHost);
      double s = 0; long long mx = 0; for (int i = 0; i < grid; i++) { s += h[i]; if (h[i] > mx) mx = h[i]; }
      printf("mode %d (icache thrash %d, wide matrix %d)  warps/SM %3d : mean %8.0f max %8lld cycles per factor+solve\n", mode, mode & 1, (mode >> 1) & 1, grid / 148, s / grid, mx);
    }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
