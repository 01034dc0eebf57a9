// get_init_pose.h -- initial board pose of a camera view and the cam0 pose guess of a frame.
// Same entry point as /root/reference/include/multicam_calibration/get_init_pose.h:24-28 (get_init_pose) plus the
// nodelet's guessCameraPose (/root/reference/src/calibration_nodelet.cpp:510-543), on top of mcc_init_pose
// (include/mcc.h), which restates the OpenCV-based computation without OpenCV.
#ifndef MULTICAM_CALIBRATION_B200_GET_INIT_POSE_H
#define MULTICAM_CALIBRATION_B200_GET_INIT_POSE_H
#include <iostream>
#include <string>
#include <vector>

#include "mcc.h"
#include "multicam_calibration/calibration_data.h"

namespace multicam_calibration {
namespace get_init_pose {

// get_init_pose.cpp:81-140; returns the zero matrix if no pose could be computed (fewer than 4 points, degenerate)
inline CameraExtrinsics get_init_pose(const std::vector<Point3<double>> &world_points, const std::vector<Point2<double>> &image_points,
                                      const std::vector<double> &intrinsics, const std::string &distModel,
                                      const std::vector<double> &distcoeff = std::vector<double>()) {
  int model = -1;
  if (distModel == "equidistant") model = MCC_MODEL_EQUIDISTANT;
  else if (distModel == "radtan") model = MCC_MODEL_RADTAN;
  else std::cout << "WARNING: unknown distortion model: " << distModel << std::endl;
  CameraExtrinsics T;
  if (world_points.size() != image_points.size() || world_points.empty() || intrinsics.size() < 4) return T;
  const mcc_status st = mcc_init_pose(model, intrinsics.data(), (int)distcoeff.size(), distcoeff.empty() ? nullptr : distcoeff.data(),
                                      (int)world_points.size(), &world_points[0].x, &image_points[0].x, T.data());
  if (st != MCC_OK) return CameraExtrinsics::Zero();
  return T;
}

}  // namespace get_init_pose

// calibration_nodelet.cpp:510-543: pose of cam0 from the first camera that saw the board, chained back through
// T_cn_cnm1; later cameras are only checked for consistency with it (warning when 1.5 - tr(R_err)/2 > 0.25).
inline bool guessCameraPose(const CalibDataVec &cameras, const CamWorldPoints &wp, const CamImagePoints &ip, CameraExtrinsics *T_0_w,
                            int frameNum) {
  CameraExtrinsics T_n_0 = identity();
  bool poseFound = false;
  for (size_t k = 0; k < cameras.size(); k++) {
    const CalibrationData &cd = cameras[k];
    T_n_0 = cd.T_cn_cnm1 * T_n_0;
    if (wp[k].empty()) continue;
    const CameraIntrinsics &ci = cd.intrinsics;
    const CameraExtrinsics T_n_w = get_init_pose::get_init_pose(wp[k], ip[k], ci.intrinsics, ci.distortion_model, ci.distortion_coeffs);
    if (!isNonZero(T_n_w)) continue;
    if (!poseFound) {
      *T_0_w = T_n_0.inverse() * T_n_w;
      poseFound = true;
    } else {
      const CameraExtrinsics T_err = (T_n_0.inverse() * T_n_w).inverse() * (*T_0_w);
      const double rot_err = 1.5 - 0.5 * (T_err(0, 0) + T_err(1, 1) + T_err(2, 2));
      if (rot_err > 0.25) {
        std::cerr << "init tf for camera " << k << " frame " << frameNum << " is off from your initial .yaml file!" << std::endl;
        std::cerr << "expected T_n_0 from " << k << " to cam 0 is roughly:" << std::endl << T_n_w * T_0_w->inverse() << std::endl;
      }
    }
  }
  return poseFound;
}

}  // namespace multicam_calibration
#endif
