// calibration_data.h -- per-camera calibration record, same fields and defaults as the reference's
// CalibrationData (/root/reference/include/multicam_calibration/calibration_data.h:20-36).  The reference fills it
// from the ROS parameter server (parse_cameras, src/calibration_data.cpp:57-87); here it is filled from a Kalibr
// YAML file (kalibr_yaml.h), since ROS is out of scope.
#ifndef MULTICAM_CALIBRATION_B200_CALIBRATION_DATA_H
#define MULTICAM_CALIBRATION_B200_CALIBRATION_DATA_H
#include <ostream>
#include <string>
#include <vector>

#include "multicam_calibration/camera_extrinsics.h"
#include "multicam_calibration/camera_intrinsics.h"
#include "multicam_calibration/types.h"

namespace multicam_calibration {
struct CalibrationData {
  typedef std::vector<CalibrationData> CalibDataVec;
  std::string name;
  CameraIntrinsics intrinsics;
  CameraExtrinsics T_cam_imu = zeros();
  CameraExtrinsics T_cn_cnm1 = identity();
  std::string rostopic;
  int tagCount{0};
  int rotateCode{-1};
  int flipCode{2};
  bool fixIntrinsics{false};
  bool fixExtrinsics{false};
  bool active{true};
  // Kalibr YAML file -> cameras cam0, cam1, ... until one is missing (calibration_nodelet.cpp:143-194)
  static CalibDataVec parse_cameras(const std::string &yaml_file);
};
using CalibDataVec = CalibrationData::CalibDataVec;
inline std::ostream &operator<<(std::ostream &os, const CalibrationData &cd) { return os << cd.intrinsics; }
}  // namespace multicam_calibration
#endif
