// get_init_pose.h -- initial board pose of a camera view and the cam0 pose guess of a frame.
// Same entry point as /root/reference/include/multicam_calibration/get_init_pose.h:24-28 (get_init_pose) plus the
// nodelet's guessCameraPose (/root/reference/src/calibration_nodelet.cpp:510-543), on top of mcc_init_pose
// (include/mcc.h), which restates the OpenCV-based computation without OpenCV.
#ifndef MULTICAM_CALIBRATION_B200_GET_INIT_POSE_H
#define MULTICAM_CALIBRATION_B200_GET_INIT_POSE_H
#include <algorithm>
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

namespace detail {
inline int model_id(const std::string &distModel) {
  if (distModel == "equidistant") return MCC_MODEL_EQUIDISTANT;
  if (distModel == "radtan") return MCC_MODEL_RADTAN;
  return -1;
}
// calibration_nodelet.cpp:510-543 without the pose computation itself: pose of cam0 from the first camera that saw the
// board, chained back through T_cn_cnm1; later cameras are only checked for consistency with it (warning when
// 1.5 - tr(R_err)/2 > 0.25).  has_points(k): camera k has detections in this frame; view_pose(k): its T_n_w (zero = none).
template <class HasPoints, class ViewPose>
inline bool chain_camera_poses(const CalibDataVec &cameras, HasPoints has_points, ViewPose view_pose, CameraExtrinsics *T_0_w, int frameNum) {
  CameraExtrinsics T_n_0 = identity();
  bool poseFound = false;
  for (size_t k = 0; k < cameras.size(); k++) {
    T_n_0 = cameras[k].T_cn_cnm1 * T_n_0;
    if (!has_points(k)) continue;
    const CameraExtrinsics T_n_w = view_pose(k);
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
}  // namespace detail

// calibration_nodelet.cpp:510-543 for one frame
inline bool guessCameraPose(const CalibDataVec &cameras, const CamWorldPoints &wp, const CamImagePoints &ip, CameraExtrinsics *T_0_w,
                            int frameNum) {
  return detail::chain_camera_poses(
      cameras, [&](size_t k) { return !wp[k].empty(); },
      [&](size_t k) {
        const CameraIntrinsics &ci = cameras[k].intrinsics;
        return get_init_pose::get_init_pose(wp[k], ip[k], ci.intrinsics, ci.distortion_model, ci.distortion_coeffs);
      },
      T_0_w, frameNum);
}

// The same for every frame of a data set (worldPoints[camera][frame], as read_corners_csv fills them): all views go through
// mcc_init_pose_batch, i.e. over the host's threads (they are independent; a large corners file has tens of thousands), then
// the frames are chained one by one exactly as above -- the same poses and the same messages as a loop over guessCameraPose.
// Returns, per frame, whether a pose was found.
inline std::vector<bool> guessCameraPoses(const CalibDataVec &cameras, const std::vector<CamWorldPoints> &worldPoints,
                                          const std::vector<CamImagePoints> &imagePoints, std::vector<CameraExtrinsics> *T_0_w, int n_threads = 0) {
  const size_t ncam = cameras.size(), nframes = ncam ? worldPoints[0].size() : 0;
  std::vector<mcc_view> views;
  std::vector<long> view_of(ncam * nframes, -1);
  for (size_t f = 0; f < nframes; f++)
    for (size_t k = 0; k < ncam; k++) {
      const auto &w = worldPoints[k][f];
      const auto &im = imagePoints[k][f];
      const CameraIntrinsics &ci = cameras[k].intrinsics;
      if (w.empty() || w.size() != im.size() || ci.intrinsics.size() < 4) continue;      // get_init_pose returns the zero matrix
      mcc_view v;
      v.model = detail::model_id(ci.distortion_model); v.n_dist = (int)ci.distortion_coeffs.size(); v.npts = (int)w.size();
      v.intr = ci.intrinsics.data(); v.dist = ci.distortion_coeffs.empty() ? nullptr : ci.distortion_coeffs.data();
      v.world_xyz = &w[0].x; v.image_uv = &im[0].x;
      view_of[f * ncam + k] = (long)views.size();
      views.push_back(v);
    }
  std::vector<double> T(views.size() * 16);
  if (mcc_init_pose_batch((int)views.size(), views.data(), T.data(), nullptr, n_threads) != MCC_OK) std::fill(T.begin(), T.end(), 0.0);
  T_0_w->assign(nframes, CameraExtrinsics());
  std::vector<bool> have(nframes, false);
  for (size_t f = 0; f < nframes; f++)
    have[f] = detail::chain_camera_poses(
        cameras, [&](size_t k) { return !worldPoints[k][f].empty(); },
        [&](size_t k) {
          if (detail::model_id(cameras[k].intrinsics.distortion_model) < 0)
            std::cout << "WARNING: unknown distortion model: " << cameras[k].intrinsics.distortion_model << std::endl;
          CameraExtrinsics Tv;
          const long v = view_of[f * ncam + k];
          if (v < 0) return Tv;
          for (int i = 0; i < 16; i++) Tv.data()[i] = T[(size_t)v * 16 + i];
          return Tv;
        },
        &(*T_0_w)[f], (int)f);
  return have;
}

}  // namespace multicam_calibration
#endif
