#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double* out, long long* cyc, int l1, int l2, double m) {
  double x = 1.0 + threadIdx.x;
  long long t0, t1;
  // 64-bit shuffle chain with a run-time source lane
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 128; i++) x = __shfl_sync(0xffffffffu, x, (l1 + i) & 31);
  t1 = clock64();
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
  // shuffle + DMUL + DFMA + DMUL per step (the elimination chain)
  double y = x * m, z = x + m;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; i++) {
    const double p = __shfl_sync(0xffffffffu, y, (l1 + i) & 31);
    const double l = __shfl_sync(0xffffffffu, y, (l2 + i) & 31);
    y = fma(z, p, -(y * l)) * m;
  }
  t1 = clock64();
  if (threadIdx.x == 0) cyc[1] = t1 - t0;
  // 32-bit shuffle chain
  int q = threadIdx.x;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 128; i++) q = __shfl_sync(0xffffffffu, q, (l1 + q) & 31);
  t1 = clock64();
  if (threadIdx.x == 0) cyc[2] = t1 - t0;
  // 8 independent 64-bit shuffles then a use (throughput of a burst)
  double a[8];
  t0 = clock64();
#pragma unroll
  for (int r = 0; r < 16; r++) {
#pragma unroll
    for (int c = 0; c < 8; c++) a[c] = __shfl_sync(0xffffffffu, y + c, (l1 + c + r) & 31);
    y = ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
  }
  t1 = clock64();
  if (threadIdx.x == 0) cyc[3] = t1 - t0;
  out[threadIdx.x] = x + y + q;
}
int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, 1024 * 8); cudaMalloc(&cyc, 16 * 8);
  for (int it = 0; it < 2; it++) k<<<1, 32>>>(out, cyc, 3, 5, 1.0000001);
  cudaDeviceSynchronize();
  long long h[16];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  printf("SHFL64 chain %.1f  elimination step (2 SHFL64 + DMUL + DFMA + DMUL) %.1f  SHFL32 chain %.1f  burst of 8 SHFL64 + adds %.1f\n", h[0] / 128.0, h[1] / 64.0, h[2] / 128.0, h[3] / 16.0);
  return 0;
}
