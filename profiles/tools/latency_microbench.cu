#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double* out, long long* cyc, double x0, double m) {
  double x = x0 + threadIdx.x;
  long long t0, t1;
  // DFMA chain
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) x = fma(x, m, 1e-9);
  t1 = clock64();
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
  // DMUL chain
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) x = x * m;
  t1 = clock64();
  if (threadIdx.x == 0) cyc[1] = t1 - t0;
  // DADD chain
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) x = x + m;
  t1 = clock64();
  if (threadIdx.x == 0) cyc[2] = t1 - t0;
  // SHFL (64-bit) chain
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) x = __shfl_sync(0xffffffffu, x, (i * 7 + 1) & 31);
  t1 = clock64();
  if (threadIdx.x == 0) cyc[3] = t1 - t0;
  // SHFL + DFMA alternating
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 128; i++) { x = __shfl_sync(0xffffffffu, x, (i * 7 + 1) & 31); x = fma(x, m, 1e-9); }
  t1 = clock64();
  if (threadIdx.x == 0) cyc[4] = t1 - t0;
  // rsqrt chain (library)
  double y = fabs(x) + 1.5;
  t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 64; i++) y = rsqrt(y) + 1.5;
  t1 = clock64();
  if (threadIdx.x == 0) cyc[5] = t1 - t0;
  // FFMA chain for reference
  float f = (float)x;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) f = fmaf(f, 1.0001f, 1e-9f);
  t1 = clock64();
  if (threadIdx.x == 0) cyc[6] = t1 - t0;
  // LDS chain
  __shared__ int sidx[64];
  if (threadIdx.x < 64) sidx[threadIdx.x] = (threadIdx.x * 5 + 3) & 63;
  __syncthreads();
  int j = threadIdx.x;
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 256; i++) j = sidx[j];
  t1 = clock64();
  if (threadIdx.x == 0) cyc[7] = t1 - t0;
  // barrier cost
  t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 64; i++) __syncthreads();
  t1 = clock64();
  if (threadIdx.x == 0) cyc[8] = t1 - t0;
  out[threadIdx.x] = x + y + f + j;
}
int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, 1024 * 8); cudaMalloc(&cyc, 16 * 8);
  for (int nt : {32, 128, 512}) {
    k<<<1, nt>>>(out, cyc, 1.0, 1.0000001);
    cudaDeviceSynchronize();
    k<<<1, nt>>>(out, cyc, 1.0, 1.0000001);
    long long h[16];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("threads %d: DFMA %.1f DMUL %.1f DADD %.1f SHFL64 %.1f SHFL64+DFMA %.1f rsqrt+DADD %.1f FFMA %.1f LDS %.1f BAR %.1f\n", nt,
           h[0] / 256.0, h[1] / 256.0, h[2] / 256.0, h[3] / 256.0, h[4] / 128.0, h[5] / 64.0, h[6] / 256.0, h[7] / 256.0, h[8] / 64.0);
  }
  return 0;
}
