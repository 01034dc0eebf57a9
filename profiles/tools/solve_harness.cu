// standalone timing + correctness harness for k_solve_cam (development only)
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <random>
#ifndef KH
#define KH "../../multicam_calibration_b200/csrc/kernels.cuh"   // build from profiles/tools: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo [-DMCC_SOLVE_TIMING] -o solve_harness solve_harness.cu
#endif
#include KH
using namespace mcc;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)
int main(int argc, char** argv) {
  const int C = argc > 1 ? atoi(argv[1]) : 106;
  const int reps = 50;
  std::mt19937_64 rng(7);
  std::normal_distribution<double> nd(0.0, 1.0);
  // SPD matrix with unit-ish diagonal
  std::vector<double> B((size_t)C * (C + 8)), A((size_t)C * C), g(C);
  for (auto& v : B) v = nd(rng);
  for (int i = 0; i < C; i++) for (int j = 0; j < C; j++) { double s = 0; for (int k = 0; k < C + 8; k++) s += B[(size_t)i * (C + 8) + k] * B[(size_t)j * (C + 8) + k]; A[(size_t)i * C + j] = s; }
  std::vector<double> d(C); for (int i = 0; i < C; i++) d[i] = 1.0 / std::sqrt(A[(size_t)i * C + i]);
  for (int i = 0; i < C; i++) for (int j = 0; j < C; j++) A[(size_t)i * C + j] *= d[i] * d[j];
  for (int i = 0; i < C; i++) g[i] = nd(rng);
  const double radius = 1e4, mind = 1e-6, maxd = 1e32;
  // CPU reference: (A + diag(clamp(Aii)/radius)) y = g
  std::vector<double> M(A), y(g);
  for (int i = 0; i < C; i++) M[(size_t)i * C + i] += std::fmin(std::fmax(A[(size_t)i * C + i], mind), maxd) / radius;
  { // Cholesky
    std::vector<double>& L = M;
    for (int j = 0; j < C; j++) { double s = L[(size_t)j * C + j]; for (int k = 0; k < j; k++) s -= L[(size_t)j * C + k] * L[(size_t)j * C + k]; L[(size_t)j * C + j] = std::sqrt(s);
      for (int i = j + 1; i < C; i++) { double t = L[(size_t)i * C + j]; for (int k = 0; k < j; k++) t -= L[(size_t)i * C + k] * L[(size_t)j * C + k]; L[(size_t)i * C + j] = t / L[(size_t)j * C + j]; } }
    for (int i = 0; i < C; i++) { double t = y[i]; for (int k = 0; k < i; k++) t -= L[(size_t)i * C + k] * y[k]; y[i] = t / L[(size_t)i * C + i]; }
    for (int i = C - 1; i >= 0; i--) { double t = y[i]; for (int k = i + 1; k < C; k++) t -= L[(size_t)k * C + i] * y[k]; y[i] = t / L[(size_t)i * C + i]; }
  }
  const int NT = (C + 1 + SY_T - 1) / SY_T, npairs = NT * (NT + 1) / 2;
  double *dH, *dg, *dss, *dsc, *dcamp, *dcand, *dyc, *dysc, *dM, *dscal, *dcamrec; int *dcol, *dfail;
  CK(cudaMalloc(&dH, sizeof(double) * C * C)); CK(cudaMalloc(&dg, sizeof(double) * C)); CK(cudaMalloc(&dss, sizeof(double) * npairs * 64)); CK(cudaMalloc(&dsc, sizeof(double) * C));
  CK(cudaMalloc(&dcamp, sizeof(double) * C)); CK(cudaMalloc(&dcand, sizeof(double) * C)); CK(cudaMalloc(&dyc, sizeof(double) * C)); CK(cudaMalloc(&dysc, sizeof(double) * C));
  CK(cudaMalloc(&dM, sizeof(double) * (C + 1) * C)); CK(cudaMalloc(&dscal, sizeof(double) * 8)); CK(cudaMalloc(&dcamrec, sizeof(double) * 64)); CK(cudaMalloc(&dcol, sizeof(int) * C)); CK(cudaMalloc(&dfail, sizeof(int)));
  std::vector<double> ones(C, 1.0), zeros(C, 0.0); std::vector<int> col(C); for (int i = 0; i < C; i++) col[i] = i;
  CK(cudaMemcpy(dH, A.data(), sizeof(double) * C * C, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dg, g.data(), sizeof(double) * C, cudaMemcpyHostToDevice));
  CK(cudaMemset(dss, 0, sizeof(double) * npairs * 64)); CK(cudaMemcpy(dsc, ones.data(), sizeof(double) * C, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dcamp, zeros.data(), sizeof(double) * C, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dcol, col.data(), sizeof(int) * C, cudaMemcpyHostToDevice)); CK(cudaMemset(dfail, 0, sizeof(int)));
  const size_t need = ((size_t)C * (C + 1) / 2 + C + 1) * sizeof(double);
  const bool packed = need <= 184 * 1024;
  if (packed) CK(cudaFuncSetAttribute(k_solve_cam<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto launch = [&]() {
    if (packed) k_solve_cam<true><<<1, CH_THREADS, need>>>(C, radius, mind, maxd, dH, dg, dss, dsc, dcol, C, dcamp, dcand, dyc, dysc, dM, dscal, dfail, nullptr, 0, dcamrec, nullptr);
    else k_solve_cam<false><<<1, CH_THREADS, 0>>>(C, radius, mind, maxd, dH, dg, dss, dsc, dcol, C, dcamp, dcand, dyc, dysc, dM, dscal, dfail, nullptr, 0, dcamrec, nullptr);
  };
  for (int i = 0; i < 5; i++) launch();
  CK(cudaDeviceSynchronize());
  cudaEventRecord(e0);
  for (int i = 0; i < reps; i++) launch();
  cudaEventRecord(e1); CK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  std::vector<double> yg(C); int fl;
  CK(cudaMemcpy(yg.data(), dyc, sizeof(double) * C, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(&fl, dfail, sizeof(int), cudaMemcpyDeviceToHost));
  double err = 0, nrm = 0; for (int i = 0; i < C; i++) { err = std::fmax(err, std::fabs(yg[i] - y[i])); nrm = std::fmax(nrm, std::fabs(y[i])); }
  printf("C %d %s: %.2f us per launch (back to back), max |y - y_ref| / max|y| = %.3e, fail flags %d\n", C, packed ? "packed" : "global", 1e3 * ms / reps, err / nrm, fl);
  return 0;
}
