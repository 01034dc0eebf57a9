// stereo_rectify.h -- rectification_matrix / projection_matrix of the first two cameras for the output YAML.
//
// The reference gets them from OpenCV in compute_RPQ (/root/reference/src/calibration_nodelet.cpp:373-410):
// cv::fisheye::stereoRectify for "equidistant" pairs (rotation vector and translation narrowed to float on the way,
// get_init_pose.cpp:33-44) and cv::stereoRectify for "radtan"/"plumb_bob" pairs, both with CALIB_ZERO_DISPARITY, and
// only when both cameras share distortion model and resolution.  OpenCV cannot be linked here, so the published
// algorithm of those two functions is restated:
//   1. both cameras are rotated by half of the relative rotation (r_r = Rodrigues(-om/2)), then by the smallest
//      rotation that brings the baseline onto the x (or y) axis:  R1 = wR r_r^T,  R2 = wR r_r,  t' = R2 t;
//   2. fisheye: new camera matrices from the undistorted, rotated mid-points of the four image edges
//      (estimateNewCameraMatrixForUndistortRectify, balance 0), f = min of the two fy; pinhole: f = mean of the two
//      fy, principal point from the undistorted, rotated image corners (float32 points, as in cvStereoRectify);
//   3. principal points averaged (CALIB_ZERO_DISPARITY), P2(0,3) = t'_x f.
// Golden vectors made with Python cv2 pin this restatement (tests/golden/rectify_golden.json).
#ifndef MULTICAM_CALIBRATION_B200_STEREO_RECTIFY_H
#define MULTICAM_CALIBRATION_B200_STEREO_RECTIFY_H
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <string>
#include <vector>

#include "multicam_calibration/calibration_data.h"

namespace multicam_calibration {
namespace rectify_detail {

struct M3 { double m[9]; };

inline M3 mul(const M3 &A, const M3 &B) {
  M3 C;
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) C.m[i * 3 + j] = A.m[i * 3] * B.m[j] + A.m[i * 3 + 1] * B.m[3 + j] + A.m[i * 3 + 2] * B.m[6 + j];
  return C;
}
inline M3 transpose(const M3 &A) {
  M3 T;
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) T.m[i * 3 + j] = A.m[j * 3 + i];
  return T;
}
inline void mulv(const M3 &A, const double *v, double *o) {
  for (int i = 0; i < 3; i++) o[i] = A.m[i * 3] * v[0] + A.m[i * 3 + 1] * v[1] + A.m[i * 3 + 2] * v[2];
}
// rotation vector -> matrix
inline M3 rodrigues(const double *r) {
  const double th = std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  M3 R = {{1, 0, 0, 0, 1, 0, 0, 0, 1}};
  if (th < DBL_EPSILON) return R;
  const double c = std::cos(th), s = std::sin(th), c1 = 1.0 - c, x = r[0] / th, y = r[1] / th, z = r[2] / th;
  const double M[9] = {c + c1 * x * x, c1 * x * y - s * z, c1 * x * z + s * y, c1 * x * y + s * z, c + c1 * y * y, c1 * y * z - s * x,
                       c1 * x * z - s * y, c1 * y * z + s * x, c + c1 * z * z};
  for (int i = 0; i < 9; i++) R.m[i] = M[i];
  return R;
}
// rotation matrix -> rotation vector (angle in [0, pi])
inline void rodrigues_inv(const M3 &R, double *r) {
  const double rx = R.m[7] - R.m[5], ry = R.m[2] - R.m[6], rz = R.m[3] - R.m[1];
  const double s = std::sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
  double c = (R.m[0] + R.m[4] + R.m[8] - 1.0) * 0.5;
  c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
  const double th = std::acos(c);
  if (s < 1e-5) {
    if (c > 0) { r[0] = r[1] = r[2] = 0.0; return; }
    // angle ~ pi: axis from the symmetric part
    double x = std::sqrt(std::max((R.m[0] + 1) * 0.5, 0.0)), y = std::sqrt(std::max((R.m[4] + 1) * 0.5, 0.0)) * (R.m[1] < 0 ? -1.0 : 1.0);
    double z = std::sqrt(std::max((R.m[8] + 1) * 0.5, 0.0)) * (R.m[2] < 0 ? -1.0 : 1.0);
    if (std::fabs(x) < std::fabs(y) && std::fabs(x) < std::fabs(z) && (R.m[5] > 0) != (y * z > 0)) z = -z;
    const double n = th / std::sqrt(x * x + y * y + z * z);
    r[0] = x * n; r[1] = y * n; r[2] = z * n;
    return;
  }
  const double k = 0.5 * th / s;
  r[0] = rx * k; r[1] = ry * k; r[2] = rz * k;
}

// step 1 of both OpenCV functions
inline void rectifying_rotations(const double *om_in, const double *t_in, M3 *R1, M3 *R2, double *t_new, int *idx_out) {
  const double om[3] = {-0.5 * om_in[0], -0.5 * om_in[1], -0.5 * om_in[2]};
  const M3 r_r = rodrigues(om);
  double t[3];
  mulv(r_r, t_in, t);
  const int idx = std::fabs(t[0]) > std::fabs(t[1]) ? 0 : 1;
  const double c = t[idx], nt = std::sqrt(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]);
  double uu[3] = {0, 0, 0};
  uu[idx] = c > 0 ? 1.0 : -1.0;
  double ww[3] = {t[1] * uu[2] - t[2] * uu[1], t[2] * uu[0] - t[0] * uu[2], t[0] * uu[1] - t[1] * uu[0]};
  const double nw = std::sqrt(ww[0] * ww[0] + ww[1] * ww[1] + ww[2] * ww[2]);
  if (nw > 0.0) { const double k = std::acos(std::fabs(c) / nt) / nw; ww[0] *= k; ww[1] *= k; ww[2] *= k; }
  const M3 wR = rodrigues(ww);
  *R1 = mul(wR, transpose(r_r));
  *R2 = mul(wR, r_r);
  mulv(*R2, t_in, t_new);
  *idx_out = idx;
}

// fisheye: pixel -> normalised, undistorted, rotated by R
inline void fisheye_undistort_rotate(double u, double v, const std::vector<double> &K4, const std::vector<double> &D, const M3 &R, double *xo, double *yo) {
  const double px = (u - K4[2]) / K4[0], py = (v - K4[3]) / K4[1];
  double thd = std::sqrt(px * px + py * py);
  thd = std::min(std::max(-M_PI / 2.0, thd), M_PI / 2.0);
  double scale = 1.0;
  if (thd > 1e-8) {
    double th = thd;
    for (int it = 0; it < 20; it++) {
      const double t2 = th * th, t4 = t2 * t2, t6 = t4 * t2, t8 = t6 * t2;
      const double k0 = D.size() > 0 ? D[0] : 0, k1 = D.size() > 1 ? D[1] : 0, k2 = D.size() > 2 ? D[2] : 0, k3 = D.size() > 3 ? D[3] : 0;
      const double fix = (th * (1 + k0 * t2 + k1 * t4 + k2 * t6 + k3 * t8) - thd) / (1 + 3 * k0 * t2 + 5 * k1 * t4 + 7 * k2 * t6 + 9 * k3 * t8);
      th -= fix;
      if (std::fabs(fix) < 1e-12) break;
    }
    scale = std::tan(th) / thd;
  }
  const double p[3] = {px * scale, py * scale, 1.0};
  double pr[3];
  mulv(R, p, pr);
  *xo = pr[0] / pr[2];
  *yo = pr[1] / pr[2];
}

// cv::fisheye::estimateNewCameraMatrixForUndistortRectify with balance = 0, fov_scale = 1, same image size
inline void fisheye_new_camera(const std::vector<double> &K4, const std::vector<double> &D, int w, int h, const M3 &R, double *fx, double *fy,
                               double *cx, double *cy) {
  double x[4], y[4], mx = 0, my = 0;
  // mid-points of the four image edges (w / 2 and h / 2 are integer divisions in OpenCV)
  const double ptsi[4][2] = {{(double)(w / 2), 0.0}, {(double)w, (double)(h / 2)}, {(double)(w / 2), (double)h}, {0.0, (double)(h / 2)}};
  for (int i = 0; i < 4; i++) { fisheye_undistort_rotate(ptsi[i][0], ptsi[i][1], K4, D, R, &x[i], &y[i]); mx += x[i]; my += y[i]; }
  mx /= 4; my /= 4;
  const double aspect = K4[0] / K4[1];
  my *= aspect;
  double minx = DBL_MAX, miny = DBL_MAX, maxx = -DBL_MAX, maxy = -DBL_MAX;
  for (int i = 0; i < 4; i++) {
    y[i] *= aspect;
    minx = std::min(minx, x[i]); maxx = std::max(maxx, x[i]); miny = std::min(miny, y[i]); maxy = std::max(maxy, y[i]);
  }
  const double f1 = w * 0.5 / (mx - minx), f2 = w * 0.5 / (maxx - mx), f3 = h * 0.5 * aspect / (my - miny), f4 = h * 0.5 * aspect / (maxy - my);
  const double f = std::max(f1, std::max(f2, std::max(f3, f4)));   // balance 0 -> fmax
  *fx = f; *fy = f / aspect;
  *cx = -mx * f + w * 0.5;
  *cy = (-my * f + h * aspect * 0.5) / aspect;
}

// pinhole: pixel -> normalised, undistorted (OpenCV's 5 fixed-point sweeps), result narrowed to float like cvUndistortPoints on CV_32F
inline void radtan_undistort(float u, float v, const std::vector<double> &K4, const std::vector<double> &Din, float *xo, float *yo) {
  double D[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (size_t i = 0; i < Din.size() && i < 8; i++) D[i] = Din[i];
  const double x0 = (u - K4[2]) / K4[0], y0 = (v - K4[3]) / K4[1];
  double x = x0, y = y0;
  for (int it = 0; it < 5; it++) {
    const double r2 = x * x + y * y;
    const double icd = (1.0 + ((D[7] * r2 + D[6]) * r2 + D[5]) * r2) / (1.0 + ((D[4] * r2 + D[1]) * r2 + D[0]) * r2);
    const double dx = 2.0 * D[2] * x * y + D[3] * (r2 + 2.0 * x * x), dy = D[2] * (r2 + 2.0 * y * y) + 2.0 * D[3] * x * y;
    x = (x0 - dx) * icd; y = (y0 - dy) * icd;
  }
  *xo = (float)x; *yo = (float)y;
}

}  // namespace rectify_detail

struct Rectification {
  double R[2][9];    // rectification_matrix of camera 0 and 1 (row-major 3x3)
  double P[2][12];   // projection_matrix (row-major 3x4)
};

// compute_RPQ (calibration_nodelet.cpp:373-410).  false: not a pair the reference would rectify.
inline bool compute_RPQ(unsigned int cam1_idx, unsigned int cam2_idx, const CalibDataVec &calib, Rectification *out) {
  using namespace rectify_detail;
  if (cam1_idx >= calib.size() || cam2_idx >= calib.size()) return false;
  const CameraIntrinsics &ci1 = calib[cam1_idx].intrinsics, &ci2 = calib[cam2_idx].intrinsics;
  const std::string &distm = ci1.distortion_model;
  if (distm != ci2.distortion_model || ci1.resolution[0] != ci2.resolution[0] || ci1.resolution[1] != ci2.resolution[1]) return false;
  const CameraExtrinsics &T = calib[cam2_idx].T_cn_cnm1;
  M3 Rm;
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) Rm.m[i * 3 + j] = T(i, j);
  double om[3], t[3] = {T(0, 3), T(1, 3), T(2, 3)};
  rodrigues_inv(Rm, om);
  const int w = ci1.resolution[0], h = ci1.resolution[1];
  M3 R1, R2;
  double tn[3], f, cc[2][2];
  int idx = 0;
  if (distm == "equidistant") {
    for (int i = 0; i < 3; i++) { om[i] = (double)(float)om[i]; t[i] = (double)(float)t[i]; }   // cv::Affine3f::Vec3
    rectifying_rotations(om, t, &R1, &R2, tn, &idx);
    double fx1, fy1, fx2, fy2;
    fisheye_new_camera(ci1.intrinsics, ci1.distortion_coeffs, w, h, R1, &fx1, &fy1, &cc[0][0], &cc[0][1]);
    fisheye_new_camera(ci2.intrinsics, ci2.distortion_coeffs, w, h, R2, &fx2, &fy2, &cc[1][0], &cc[1][1]);
    f = std::min(fy1, fy2);
    idx = 0;   // the fisheye version always writes t'_x f into P2(0,3)
  } else if (distm == "radtan" || distm == "plumb_bob") {
    rectifying_rotations(om, t, &R1, &R2, tn, &idx);
    f = (ci1.intrinsics[idx == 0 ? 1 : 0] + ci2.intrinsics[idx == 0 ? 1 : 0]) * 0.5;
    for (int k = 0; k < 2; k++) {
      const CameraIntrinsics &ci = k == 0 ? ci1 : ci2;
      const M3 &Rk = k == 0 ? R1 : R2;
      double ax = 0, ay = 0;
      for (int i = 0; i < 4; i++) {
        const float u = (float)((i % 2) * (w - 1)), v = (float)((i < 2 ? 0 : 1) * (h - 1));
        float xn, yn;
        radtan_undistort(u, v, ci.intrinsics, ci.distortion_coeffs, &xn, &yn);
        const double p[3] = {xn, yn, 1.0};
        double pr[3];
        mulv(Rk, p, pr);
        ax += (double)(float)(f * pr[0] / pr[2]);
        ay += (double)(float)(f * pr[1] / pr[2]);
      }
      cc[k][0] = (w - 1) / 2.0 - ax / 4;
      cc[k][1] = (h - 1) / 2.0 - ay / 4;
    }
  } else {
    return false;
  }
  // CALIB_ZERO_DISPARITY: one common principal point
  const double cxm = (cc[0][0] + cc[1][0]) * 0.5, cym = (cc[0][1] + cc[1][1]) * 0.5;
  for (int k = 0; k < 2; k++) {
    const M3 &Rk = k == 0 ? R1 : R2;
    for (int i = 0; i < 9; i++) out->R[k][i] = Rk.m[i];
    const double P[12] = {f, 0, cxm, 0, 0, f, cym, 0, 0, 0, 1, 0};
    for (int i = 0; i < 12; i++) out->P[k][i] = P[i];
  }
  out->P[1][idx == 0 ? 3 : 7] = tn[idx] * f;
  return true;
}

}  // namespace multicam_calibration
#endif
