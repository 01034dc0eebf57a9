// camera_extrinsics.h -- 4x4 rigid transform.
// The reference aliases Eigen::Matrix<double,4,4> (/root/reference/include/multicam_calibration/
// camera_extrinsics.h:13-17, src/camera_extrinsics.cpp:11-23); this is a dependency-free stand-in with the
// operations the Calibrator API and the YAML I/O use: (i,j) access, product, inverse, Identity/Zero.
#ifndef MULTICAM_CALIBRATION_B200_CAMERA_EXTRINSICS_H
#define MULTICAM_CALIBRATION_B200_CAMERA_EXTRINSICS_H
#include <cmath>
#include <cstring>
#include <ostream>
#include <utility>
#include <vector>

namespace multicam_calibration {

class CameraExtrinsics {
 public:
  CameraExtrinsics() { std::memset(m_, 0, sizeof(m_)); }
  static CameraExtrinsics Zero() { return CameraExtrinsics(); }
  static CameraExtrinsics Identity() {
    CameraExtrinsics T;
    for (int i = 0; i < 4; i++) T(i, i) = 1.0;
    return T;
  }
  double &operator()(int i, int j) { return m_[4 * i + j]; }
  double operator()(int i, int j) const { return m_[4 * i + j]; }
  const double *data() const { return m_; }   // row-major, what the C ABI expects
  double *data() { return m_; }
  unsigned int rows() const { return 4; }
  unsigned int cols() const { return 4; }
  CameraExtrinsics operator*(const CameraExtrinsics &B) const {
    CameraExtrinsics C;
    for (int i = 0; i < 4; i++)
      for (int j = 0; j < 4; j++) {
        double s = 0;
        for (int k = 0; k < 4; k++) s += (*this)(i, k) * B(k, j);
        C(i, j) = s;
      }
    return C;
  }
  bool operator==(const CameraExtrinsics &B) const { return std::memcmp(m_, B.m_, sizeof(m_)) == 0 || equal_values(B); }
  bool operator!=(const CameraExtrinsics &B) const { return !(*this == B); }
  // general inverse (Gauss-Jordan with partial pivoting)
  CameraExtrinsics inverse() const {
    double a[4][8];
    for (int i = 0; i < 4; i++)
      for (int j = 0; j < 4; j++) { a[i][j] = (*this)(i, j); a[i][4 + j] = i == j ? 1.0 : 0.0; }
    for (int c = 0; c < 4; c++) {
      int p = c;
      for (int r = c + 1; r < 4; r++) if (std::fabs(a[r][c]) > std::fabs(a[p][c])) p = r;
      if (p != c) for (int j = 0; j < 8; j++) std::swap(a[c][j], a[p][j]);
      const double d = 1.0 / a[c][c];
      for (int j = 0; j < 8; j++) a[c][j] *= d;
      for (int r = 0; r < 4; r++)
        if (r != c && a[r][c] != 0.0) { const double f = a[r][c]; for (int j = 0; j < 8; j++) a[r][j] -= f * a[c][j]; }
    }
    CameraExtrinsics R;
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) R(i, j) = a[i][4 + j];
    return R;
  }

 private:
  bool equal_values(const CameraExtrinsics &B) const {
    for (int i = 0; i < 16; i++) if (m_[i] != B.m_[i]) return false;
    return true;
  }
  double m_[16];
};

inline CameraExtrinsics zeros() { return CameraExtrinsics::Zero(); }
inline CameraExtrinsics identity() { return CameraExtrinsics::Identity(); }
// camera_extrinsics.cpp:17-22: "non-zero" means the sum of all 16 entries is not 0
inline bool isNonZero(const CameraExtrinsics &T) {
  double s = 0;
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) s += T(i, j);
  return s != 0;
}
inline std::ostream &operator<<(std::ostream &os, const CameraExtrinsics &T) {
  for (int i = 0; i < 4; i++) os << T(i, 0) << " " << T(i, 1) << " " << T(i, 2) << " " << T(i, 3) << (i < 3 ? "\n" : "");
  return os;
}
typedef std::vector<CameraExtrinsics> CameraExtrinsicsVec;

}  // namespace multicam_calibration
#endif
