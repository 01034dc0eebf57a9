// camera_intrinsics.h -- same fields and defaults as the reference's CameraIntrinsics
// (/root/reference/include/multicam_calibration/camera_intrinsics.h:14-25).
#ifndef MULTICAM_CALIBRATION_B200_CAMERA_INTRINSICS_H
#define MULTICAM_CALIBRATION_B200_CAMERA_INTRINSICS_H
#include <ostream>
#include <string>
#include <vector>

namespace multicam_calibration {
struct CameraIntrinsics {
  CameraIntrinsics() : distortion_coeffs(4, 0.0), intrinsics(4, 0.0), resolution(2, 0) {}
  std::vector<double> distortion_coeffs;
  std::vector<double> intrinsics;   // fx, fy, cx, cy
  std::vector<int> resolution;
  std::string distortion_model;     // "equidistant" | "radtan"
  std::string camera_model;         // "pinhole"
};
inline std::ostream &operator<<(std::ostream &os, const CameraIntrinsics &ci) {
  auto vec = [&](const char *name, const auto &v) {
    os << name << "[";
    for (const auto &e : v) os << " " << e;
    os << "]" << std::endl;
  };
  os << "camera mod: " << ci.camera_model << std::endl;
  vec("intrinsics: ", ci.intrinsics);
  os << "dist model: " << ci.distortion_model << std::endl;
  vec("distortion: ", ci.distortion_coeffs);
  vec("resolution: ", ci.resolution);
  return os;
}
}  // namespace multicam_calibration
#endif
