// init_pose.cpp -- initial board pose of one camera view from a homography (host C++, part of libmcc.so).
//
// The step right before the solve path.  Restates what the reference does with OpenCV in
// /root/reference/src/get_init_pose.cpp:81-140 (get_init_pose) and :46-75 (decomposeHomography):
//   world (x, y) and image points are narrowed to float (cv::Point2f, :98-113); the image points are undistorted
//   to normalised coordinates with the camera's model (cv::fisheye::undistortPoints / cv::undistortPoints, :114-118,
//   again to float); a homography world plane -> normalised image is fitted to ALL points (cv::findHomography with
//   its default method, :126: normalised DLT followed by a few damped Gauss-Newton refinement steps on the
//   reprojection error); the homography is sign-fixed, scaled by its SECOND singular value, its first two columns
//   orthonormalised into r1, r2, r3 = r1 x r2, t = third column (:46-75).
// OpenCV is not available to link here, so the numerical kernels below (iterative undistortion, DLT through a
// Jacobi eigen-solver, refinement, 3x3 singular values) are this repo's own; agreement with cv2 is therefore close
// but not bitwise (tests/test_init_pose.py states the tolerance).  The LM solve converges from either.
#include <cmath>
#include <cstring>
#include <vector>
#include <thread>
#include <atomic>
#include <algorithm>

#include "../../include/mcc.h"

namespace {

// cyclic Jacobi for a symmetric n x n matrix (n <= 9): eigenvalues in d, eigenvectors in the columns of V
template <int N>
void jacobi_eig(double A[N][N], double d[N], double V[N][N]) {
  for (int i = 0; i < N; i++) { for (int j = 0; j < N; j++) V[i][j] = i == j ? 1.0 : 0.0; }
  for (int sweep = 0; sweep < 60; sweep++) {
    double off = 0.0;
    for (int i = 0; i < N; i++) for (int j = i + 1; j < N; j++) off += A[i][j] * A[i][j];
    if (off < 1e-300) break;
    for (int p = 0; p < N; p++)
      for (int q = p + 1; q < N; q++) {
        if (std::fabs(A[p][q]) < 1e-300) continue;
        const double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
        const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
        const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < N; k++) { const double akp = A[k][p], akq = A[k][q]; A[k][p] = c * akp - s * akq; A[k][q] = s * akp + c * akq; }
        for (int k = 0; k < N; k++) { const double apk = A[p][k], aqk = A[q][k]; A[p][k] = c * apk - s * aqk; A[q][k] = s * apk + c * aqk; }
        for (int k = 0; k < N; k++) { const double vkp = V[k][p], vkq = V[k][q]; V[k][p] = c * vkp - s * vkq; V[k][q] = s * vkp + c * vkq; }
      }
  }
  for (int i = 0; i < N; i++) d[i] = A[i][i];
}

// equidistant model: theta_d = theta (1 + k1 th^2 + k2 th^4 + k3 th^6 + k4 th^8)  ->  normalised point
void undistort_equidistant(double u, double v, const double* K4, const double* D, int nd, double* xo, double* yo) {
  const double px = (u - K4[2]) / K4[0], py = (v - K4[3]) / K4[1];
  double thd = std::sqrt(px * px + py * py);
  thd = std::fmin(std::fmax(-M_PI / 2.0, thd), M_PI / 2.0);
  double scale = 0.0;
  if (thd > 1e-8) {
    double th = thd;
    for (int it = 0; it < 20; it++) {
      double f = 1.0, fp = 1.0, p = th * th, th2 = p;
      for (int i = 0; i < nd && i < 4; i++) { f += D[i] * p; fp += (2 * i + 3) * D[i] * p; p *= th2; }
      const double fix = (th * f - thd) / fp;
      th -= fix;
      if (std::fabs(fix) < 1e-12) break;
    }
    scale = std::tan(th) / thd;
  } else {
    scale = 1.0;
  }
  *xo = px * scale; *yo = py * scale;
}

// radtan model (k1,k2,p1,p2,k3,k4,k5,k6): fixed-point iteration on the normalised point
void undistort_radtan(double u, double v, const double* K4, const double* Din, int nd, double* xo, double* yo) {
  double D[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int i = 0; i < nd && i < 8; i++) D[i] = Din[i];
  const double x0 = (u - K4[2]) / K4[0], y0 = (v - K4[3]) / K4[1];
  double x = x0, y = y0;
  for (int it = 0; it < 20; it++) {
    const double r2 = x * x + y * y;
    const double icd = (1.0 + ((D[7] * r2 + D[6]) * r2 + D[5]) * r2) / (1.0 + ((D[4] * r2 + D[1]) * r2 + D[0]) * r2);
    const double dx = 2.0 * D[2] * x * y + D[3] * (r2 + 2.0 * x * x);
    const double dy = D[2] * (r2 + 2.0 * y * y) + 2.0 * D[3] * x * y;
    const double xn = (x0 - dx) * icd, yn = (y0 - dy) * icd;
    const double ch = std::fabs(xn - x) + std::fabs(yn - y);
    x = xn; y = yn;
    if (ch < 1e-14) break;
  }
  *xo = x; *yo = y;
}

// H (row-major 3x3, H[8] = 1) mapping (X, Y, 1) -> (x, y, 1): normalised DLT + damped Gauss-Newton refinement
bool fit_homography(int n, const float* W, const float* P, double* H) {
  if (n < 4) return false;
  // isotropic normalisation of both point sets
  double cw[2] = {0, 0}, cp[2] = {0, 0};
  for (int i = 0; i < n; i++) { cw[0] += W[2 * i]; cw[1] += W[2 * i + 1]; cp[0] += P[2 * i]; cp[1] += P[2 * i + 1]; }
  for (int k = 0; k < 2; k++) { cw[k] /= n; cp[k] /= n; }
  double dw = 0, dp = 0;
  for (int i = 0; i < n; i++) {
    dw += std::hypot(W[2 * i] - cw[0], W[2 * i + 1] - cw[1]);
    dp += std::hypot(P[2 * i] - cp[0], P[2 * i + 1] - cp[1]);
  }
  if (!(dw > 0) || !(dp > 0)) return false;
  const double sw = std::sqrt(2.0) * n / dw, sp = std::sqrt(2.0) * n / dp;
  double A[9][9];
  std::memset(A, 0, sizeof(A));
  for (int i = 0; i < n; i++) {
    const double X = (W[2 * i] - cw[0]) * sw, Y = (W[2 * i + 1] - cw[1]) * sw;
    const double x = (P[2 * i] - cp[0]) * sp, y = (P[2 * i + 1] - cp[1]) * sp;
    const double r1[9] = {X, Y, 1, 0, 0, 0, -x * X, -x * Y, -x};
    const double r2[9] = {0, 0, 0, X, Y, 1, -y * X, -y * Y, -y};
    for (int a = 0; a < 9; a++) for (int b = a; b < 9; b++) A[a][b] += r1[a] * r1[b] + r2[a] * r2[b];
  }
  for (int a = 0; a < 9; a++) for (int b = 0; b < a; b++) A[a][b] = A[b][a];
  double d[9], V[9][9];
  jacobi_eig<9>(A, d, V);
  int m = 0;
  for (int i = 1; i < 9; i++) if (d[i] < d[m]) m = i;
  double Hn[9];
  for (int i = 0; i < 9; i++) Hn[i] = V[i][m];
  // denormalise: H = Tp^-1 Hn Tw
  const double Tw[9] = {sw, 0, -sw * cw[0], 0, sw, -sw * cw[1], 0, 0, 1};
  const double Tpi[9] = {1 / sp, 0, cp[0], 0, 1 / sp, cp[1], 0, 0, 1};
  double T1[9], T2[9];
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { double s = 0; for (int k = 0; k < 3; k++) s += Hn[i * 3 + k] * Tw[k * 3 + j]; T1[i * 3 + j] = s; }
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { double s = 0; for (int k = 0; k < 3; k++) s += Tpi[i * 3 + k] * T1[k * 3 + j]; T2[i * 3 + j] = s; }
  if (std::fabs(T2[8]) < 1e-300) return false;
  for (int i = 0; i < 9; i++) H[i] = T2[i] / T2[8];
  // refinement of the 8 free entries on the reprojection error (10 damped Gauss-Newton steps)
  double lambda = 1e-3;
  auto cost_of = [&](const double* h) {
    double c = 0;
    for (int i = 0; i < n; i++) {
      const double X = W[2 * i], Y = W[2 * i + 1];
      const double w = h[6] * X + h[7] * Y + 1.0;
      const double ex = (h[0] * X + h[1] * Y + h[2]) / w - P[2 * i], ey = (h[3] * X + h[4] * Y + h[5]) / w - P[2 * i + 1];
      c += ex * ex + ey * ey;
    }
    return c;
  };
  double cost = cost_of(H);
  for (int it = 0; it < 10; it++) {
    double JtJ[8][8], Jtr[8];
    std::memset(JtJ, 0, sizeof(JtJ));
    std::memset(Jtr, 0, sizeof(Jtr));
    for (int i = 0; i < n; i++) {
      const double X = W[2 * i], Y = W[2 * i + 1];
      const double w = H[6] * X + H[7] * Y + 1.0, iw = 1.0 / w;
      const double px = (H[0] * X + H[1] * Y + H[2]) * iw, py = (H[3] * X + H[4] * Y + H[5]) * iw;
      const double jx[8] = {X * iw, Y * iw, iw, 0, 0, 0, -X * px * iw, -Y * px * iw};
      const double jy[8] = {0, 0, 0, X * iw, Y * iw, iw, -X * py * iw, -Y * py * iw};
      const double ex = px - P[2 * i], ey = py - P[2 * i + 1];
      for (int a = 0; a < 8; a++) {
        Jtr[a] += jx[a] * ex + jy[a] * ey;
        for (int b = a; b < 8; b++) JtJ[a][b] += jx[a] * jx[b] + jy[a] * jy[b];
      }
    }
    bool improved = false;
    for (int attempt = 0; attempt < 5 && !improved; attempt++) {
      double L[8][8], y[8];
      for (int a = 0; a < 8; a++) for (int b = 0; b < 8; b++) L[a][b] = a <= b ? JtJ[a][b] : JtJ[b][a];
      for (int a = 0; a < 8; a++) L[a][a] *= 1.0 + lambda;
      bool ok = true;
      for (int j = 0; j < 8 && ok; j++) {   // Cholesky, lower triangle in place
        double s = L[j][j];
        for (int k = 0; k < j; k++) s -= L[j][k] * L[j][k];
        if (!(s > 0)) { ok = false; break; }
        L[j][j] = std::sqrt(s);
        for (int i = j + 1; i < 8; i++) { double v = L[i][j]; for (int k = 0; k < j; k++) v -= L[i][k] * L[j][k]; L[i][j] = v / L[j][j]; }
      }
      if (!ok) { lambda *= 10; continue; }
      for (int i = 0; i < 8; i++) { double v = Jtr[i]; for (int k = 0; k < i; k++) v -= L[i][k] * y[k]; y[i] = v / L[i][i]; }
      for (int i = 7; i >= 0; i--) { double v = y[i]; for (int k = i + 1; k < 8; k++) v -= L[k][i] * y[k]; y[i] = v / L[i][i]; }
      double Hc[9];
      for (int i = 0; i < 8; i++) Hc[i] = H[i] - y[i];
      Hc[8] = 1.0;
      const double c2 = cost_of(Hc);
      if (c2 <= cost) { std::memcpy(H, Hc, sizeof(Hc)); improved = true; lambda = std::fmax(lambda * 0.1, 1e-12); if (cost - c2 <= 1e-14 * cost) { cost = c2; it = 10; } cost = c2; }
      else lambda *= 10;
    }
    if (!improved) break;
  }
  return true;
}

}  // namespace

extern "C" mcc_status mcc_init_pose(int model, const double* intr, int n_dist, const double* dist, int npts, const double* world_xyz,
                                    const double* image_uv, double* T_cam_world) {
  if (!intr || !T_cam_world || npts < 0 || (npts && (!world_xyz || !image_uv)) || n_dist < 0 || n_dist > MCC_MAX_DIST || (n_dist && !dist))
    return MCC_ERR_BAD_ARG;
  if (npts < 4) return MCC_ERR_NO_DATA;
  std::vector<float> W(2 * (size_t)npts), P(2 * (size_t)npts);
  for (int i = 0; i < npts; i++) {
    W[2 * i] = (float)world_xyz[3 * i]; W[2 * i + 1] = (float)world_xyz[3 * i + 1];      // cv::Point2f(w.x, w.y)
    const float u = (float)image_uv[2 * i], v = (float)image_uv[2 * i + 1];                // cv::Point2f(i.x, i.y)
    double x, y;
    if (model == MCC_MODEL_EQUIDISTANT) undistort_equidistant(u, v, intr, dist, n_dist, &x, &y);
    else if (model == MCC_MODEL_RADTAN) undistort_radtan(u, v, intr, dist, n_dist, &x, &y);
    else { x = u; y = v; }                                                                  // "unknown distortion model": ipu = ip
    P[2 * i] = (float)x; P[2 * i + 1] = (float)y;
  }
  double H[9];
  if (!fit_homography(npts, W.data(), P.data(), H)) return MCC_ERR_NUMERICAL;
  // decomposeHomography (get_init_pose.cpp:46-75)
  if (H[8] < 0) for (int i = 0; i < 9; i++) H[i] = -H[i];
  double HtH[3][3], d[3], V[3][3];
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { double s = 0; for (int k = 0; k < 3; k++) s += H[k * 3 + i] * H[k * 3 + j]; HtH[i][j] = s; }
  jacobi_eig<3>(HtH, d, V);
  double sv[3] = {d[0], d[1], d[2]};
  for (int i = 0; i < 3; i++) for (int j = i + 1; j < 3; j++) if (sv[j] > sv[i]) { const double t = sv[i]; sv[i] = sv[j]; sv[j] = t; }
  const double s2 = std::sqrt(std::fmax(sv[1], 0.0));      // the second singular value
  if (!(s2 > 0)) return MCC_ERR_NUMERICAL;
  for (int i = 0; i < 9; i++) H[i] /= s2;
  double r1[3] = {H[0], H[3], H[6]}, r2[3] = {H[1], H[4], H[7]};
  const double n1 = std::sqrt(r1[0] * r1[0] + r1[1] * r1[1] + r1[2] * r1[2]);
  for (int i = 0; i < 3; i++) r1[i] /= n1;
  const double dp = r2[0] * r1[0] + r2[1] * r1[1] + r2[2] * r1[2];
  for (int i = 0; i < 3; i++) r2[i] -= dp * r1[i];
  const double n2 = std::sqrt(r2[0] * r2[0] + r2[1] * r2[1] + r2[2] * r2[2]);
  for (int i = 0; i < 3; i++) r2[i] /= n2;
  const double r3[3] = {r1[1] * r2[2] - r1[2] * r2[1], r1[2] * r2[0] - r1[0] * r2[2], r1[0] * r2[1] - r1[1] * r2[0]};
  for (int i = 0; i < 3; i++) {
    T_cam_world[i * 4 + 0] = r1[i]; T_cam_world[i * 4 + 1] = r2[i]; T_cam_world[i * 4 + 2] = r3[i];
    T_cam_world[i * 4 + 3] = H[i * 3 + 2];
  }
  T_cam_world[12] = T_cam_world[13] = T_cam_world[14] = 0.0; T_cam_world[15] = 1.0;
  return MCC_OK;
}

// many views at once: independent, so they are dealt to host threads by a shared counter
extern "C" mcc_status mcc_init_pose_batch(int n_views, const mcc_view* views, double* T_cam_world, mcc_status* status, int n_threads) {
  if (n_views < 0 || (n_views && (!views || !T_cam_world))) return MCC_ERR_BAD_ARG;
  if (n_threads <= 0) n_threads = (int)std::max(1u, std::thread::hardware_concurrency());
  n_threads = std::min(n_threads, std::max(1, n_views / 16));      // not worth a thread for a handful of views
  std::atomic<int> next(0);
  auto work = [&]() {
    for (;;) {
      const int v0 = next.fetch_add(16);
      if (v0 >= n_views) break;
      for (int v = v0; v < std::min(n_views, v0 + 16); v++) {
        const mcc_view& w = views[v];
        double* T = T_cam_world + (size_t)v * 16;
        const mcc_status st = mcc_init_pose(w.model, w.intr, w.n_dist, w.dist, w.npts, w.world_xyz, w.image_uv, T);
        if (st != MCC_OK) std::memset(T, 0, 16 * sizeof(double));
        if (status) status[v] = st;
      }
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < n_threads; t++) pool.emplace_back(work);
  work();
  for (auto& th : pool) th.join();
  return MCC_OK;
}
