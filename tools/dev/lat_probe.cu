// Developer micro-probe: dependent-issue latencies (cycles) of the operations k_solve's chains are made of, one warp.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lat_probe lat_probe.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>
#define N 512
__global__ void probe(double* out, long long* cyc, double seed, int nwarps_busy) {
  __shared__ double sm[64];
  const int lane = threadIdx.x & 31;
  double x = seed + lane * 1e-3, y = 1.0000001, z = 1e-9;
  long long t0, t1;
  int k = 0;
  sm[lane] = x; sm[lane + 32] = y;
  __syncwarp();
#define RUN(body) t0 = clock64(); _Pragma("unroll 1") for (int i = 0; i < N / 8; i++) { body body body body body body body body } t1 = clock64(); if (threadIdx.x == 0) cyc[k] = t1 - t0; k++;
  RUN(x = fma(x, y, z);)
  RUN(x = x * y;)
  RUN(x = x + z;)
  RUN(x = rsqrt(x) + 1.5;)
  RUN(x = 1.0 / x + 1.5;)
  RUN(x = sqrt(x) + 1.5;)
  RUN(x = __shfl_sync(0xffffffffu, x, (lane + 1) & 31);)
  RUN(x = __shfl_xor_sync(0xffffffffu, x, 1) + z;)
  RUN(sm[lane] = x; __syncwarp(); x = sm[(lane + 1) & 31]; __syncwarp();)
  RUN(x = sm[((int)x) & 31] + 1.0;)
  { float f = (float)x; RUN(f = fmaf(f, 1.0001f, 1e-7f);) x += f; }
  { int q = lane; RUN(q = q * 3 + 1;) x += q; }
  out[threadIdx.x] = x;
}
int main() {
  double* out; long long* cyc; cudaMalloc(&out, 1024 * 8); cudaMalloc(&cyc, 16 * 8);
  const char* names[] = {"DFMA", "DMUL", "DADD", "rsqrt(f64)+add", "1/x(f64)+add", "sqrt(f64)+add", "SHFL.IDX f64", "SHFL.BFLY f64 + DADD",
                         "STS+sync+LDS+sync", "LDS dependent + DADD", "FFMA", "IMAD"};
  for (int threads : {32, 128, 512}) {
    probe<<<1, threads>>>(out, cyc, 1.25, 0);
    cudaDeviceSynchronize();
    long long h[16]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("---- %d threads in the CTA (warps per SMSP: %d)\n", threads, threads / 128 ? threads / 128 : 1);
    for (int i = 0; i < 12; i++) printf("%-24s %7.1f cycles/op\n", names[i], (double)h[i] / N);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
