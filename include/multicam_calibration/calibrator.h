// calibrator.h -- C++ facade with the reference's Calibrator interface on top of the C ABI (include/mcc.h).
//
// Same public methods as /root/reference/include/multicam_calibration/calibrator.h:17-49:
// setFixAllIntrinsics, showCameraStatus, setCameras, addPoints, runCalibration, getCalibrationResults,
// testCalibration, removeOutliers.  Everything below the class (problem assembly, the Ceres solve) runs on the GPU
// in libmcc.so; this header only marshals the reference's container types into flat arrays.
// Extensions (not in the reference): summary(), lastStatus(), setDevice(), setDevices(), setVerbose().
#ifndef MULTICAM_CALIBRATION_B200_CALIBRATOR_H
#define MULTICAM_CALIBRATION_B200_CALIBRATOR_H
#include <cstdio>
#include <iostream>
#include <string>
#include <vector>

#include "mcc.h"
#include "multicam_calibration/calibration_data.h"

namespace multicam_calibration {

class Calibrator {
 public:
  // the reference uses Eigen::Matrix<double,2,1>; (i) access is all its callers need
  struct Vec2d {
    double v[2];
    Vec2d(double a = 0, double b = 0) : v{a, b} {}
    double operator()(int i) const { return v[i]; }
    double &operator()(int i) { return v[i]; }
  };
  using Residuals = std::vector<std::vector<std::vector<Vec2d>>>;   // [frame][cam][point]

  Calibrator() { mcc_create(&h_); }
  ~Calibrator() { mcc_destroy(h_); }
  Calibrator(const Calibrator &) = delete;
  Calibrator &operator=(const Calibrator &) = delete;

  void setFixAllIntrinsics(bool f) { mcc_set_fix_all_intrinsics(h_, f ? 1 : 0); }

  // calibrator.cpp:217-223
  void showCameraStatus() const {
    for (const auto &cam : calibrationData_)
      std::printf("%-10s: fix intr: %d, fix extr: %d, active: %d\n", cam.name.c_str(), (int)cam.fixIntrinsics, (int)cam.fixExtrinsics, (int)cam.active);
  }

  // calibrator.cpp:225-227
  void setCameras(const CalibDataVec &calibData) {
    calibrationData_ = calibData;
    std::vector<mcc_camera> cams(calibData.size());
    for (size_t k = 0; k < calibData.size(); k++) {
      const CameraIntrinsics &ci = calibData[k].intrinsics;
      mcc_camera &c = cams[k];
      c.model = ci.distortion_model == "equidistant" ? MCC_MODEL_EQUIDISTANT : (ci.distortion_model == "radtan" ? MCC_MODEL_RADTAN : -1);
      c.n_intr = (int)ci.intrinsics.size();
      c.n_dist = (int)ci.distortion_coeffs.size();
      c.intr = ci.intrinsics.data();
      c.dist = ci.distortion_coeffs.data();
      for (int i = 0; i < 16; i++) c.T_cn_cnm1[i] = calibData[k].T_cn_cnm1.data()[i];
      c.fix_intr = calibData[k].fixIntrinsics;
      c.fix_extr = calibData[k].fixExtrinsics;
      c.active = calibData[k].active;
    }
    status_ = mcc_set_cameras(h_, (int)cams.size(), cams.data());
    if (status_ != MCC_OK) std::cout << "ERROR: " << mcc_last_error(h_) << std::endl;   // e.g. unknown distortion model (calibrator.cpp:131-133)
  }

  // calibrator.cpp:229-249; wp / ip are indexed by camera.  frameNum is unused, as in the reference.
  void addPoints(int frameNum, const CamWorldPoints &wp, const CamImagePoints &ip, const CameraExtrinsics &cam0PoseGuess) {
    (void)frameNum;
    if (wp.size() != calibrationData_.size() || ip.size() != calibrationData_.size()) {
      std::cerr << "number of cams != width of array!" << std::endl;
      status_ = MCC_ERR_BAD_ARG;
      return;
    }
    const size_t ncam = wp.size();
    std::vector<int> n(ncam);
    std::vector<const double *> w(ncam), im(ncam);
    for (size_t k = 0; k < ncam; k++) {
      n[k] = (int)wp[k].size();
      if (ip[k].size() != wp[k].size()) { std::cerr << "world / image point count mismatch for cam " << k << std::endl; status_ = MCC_ERR_BAD_ARG; return; }
      w[k] = wp[k].empty() ? nullptr : &wp[k][0].x;
      im[k] = ip[k].empty() ? nullptr : &ip[k][0].x;
    }
    status_ = mcc_add_frame(h_, (int)ncam, n.data(), w.data(), im.data(), cam0PoseGuess.data());
    if (status_ == MCC_OK) npts_.push_back(n);
  }

  // calibrator.cpp:285-313
  void runCalibration() {
    if (calibrationData_.empty() || npts_.empty()) {
      std::fprintf(stderr, "no data to run on!\n");
      status_ = MCC_ERR_NO_DATA;
      return;
    }
    std::printf("camera status at start:\n");
    showCameraStatus();
    mcc_options o;
    mcc_default_options(&o);
    o.verbose = verbose_ ? 1 : 0;
    status_ = mcc_run(h_, &o, &summary_);
    if (status_ != MCC_OK) { std::fprintf(stderr, "calibration failed: %s\n", mcc_last_error(h_)); return; }
    if (verbose_) printReport();
  }

  // calibrator.cpp:415-463
  CalibDataVec getCalibrationResults() const {
    CalibDataVec res;
    if (calibrationData_.empty() || npts_.empty()) return res;
    const CameraExtrinsics &T_cam0_imu = calibrationData_[0].T_cam_imu;
    CameraExtrinsics T_cn_0 = CameraExtrinsics::Identity();
    for (size_t k = 0; k < calibrationData_.size(); k++) {
      CalibrationData cd(calibrationData_[k]);
      double intr[4], dist[MCC_MAX_DIST], T[16];
      if (mcc_get_camera(h_, (int)k, intr, dist, T) != MCC_OK) return CalibDataVec();
      for (size_t i = 0; i < cd.intrinsics.intrinsics.size() && i < 4; i++) cd.intrinsics.intrinsics[i] = intr[i];
      for (size_t i = 0; i < cd.intrinsics.distortion_coeffs.size(); i++) cd.intrinsics.distortion_coeffs[i] = dist[i];
      if (k > 0) {
        for (int i = 0; i < 16; i++) cd.T_cn_cnm1.data()[i] = T[i];
        T_cn_0 = cd.T_cn_cnm1 * T_cn_0;
        if (isNonZero(T_cam0_imu)) cd.T_cam_imu = T_cn_0 * T_cam0_imu;
      }
      res.push_back(cd);
    }
    return res;
  }

  // calibrator.cpp:316-351: residuals[frame][cam][point]; frames without residual block stay empty
  void testCalibration(Residuals *resptr) {
    Residuals &res = *resptr;
    const long long n = mcc_num_residuals(h_);
    std::vector<double> flat((size_t)(n > 0 ? n : 1));
    status_ = mcc_eval_residuals(h_, flat.data(), (size_t)n);
    if (status_ != MCC_OK) { std::fprintf(stderr, "testCalibration failed: %s\n", mcc_last_error(h_)); return; }
    std::vector<uint8_t> mask((size_t)mcc_num_observations(h_) + 1);
    mcc_get_mask(h_, mask.data(), (size_t)mcc_num_observations(h_));
    size_t o = 0, r = 0;
    for (size_t f = 0; f < npts_.size(); f++) {
      res.push_back(std::vector<std::vector<Vec2d>>());
      size_t nvalid = 0, oo = o;
      for (size_t k = 0; k < npts_[f].size(); k++) {
        if (calibrationData_[k].active) for (int i = 0; i < npts_[f][k]; i++) nvalid += mask[oo + i] != 0;
        oo += npts_[f][k];
      }
      if (nvalid > 0) {
        for (size_t k = 0; k < npts_[f].size(); k++) {
          res.back().push_back(std::vector<Vec2d>());
          if (calibrationData_[k].active)
            for (int i = 0; i < npts_[f][k]; i++)
              if (mask[o + i]) { res.back().back().push_back(Vec2d(flat[r], flat[r + 1])); r += 2; }
          o += npts_[f][k];
        }
      } else {
        o = oo;
      }
    }
  }

  // calibrator.cpp:372-413
  void removeOutliers(double maxErr) {
    std::cout << "removing outliers with pixel error more than " << maxErr << std::endl;
    long long nd = 0;
    status_ = mcc_remove_outliers(h_, maxErr, &nd);
    if (status_ == MCC_OK) std::cout << "dropped " << nd << " bad points" << std::endl;
  }

  // ---- extensions
  const mcc_summary &summary() const { return summary_; }
  mcc_status lastStatus() const { return status_; }
  const char *lastError() const { return mcc_last_error(h_); }
  void setDevice(int d) { mcc_set_device(h_, d); }
  // several GPUs behind this one Calibrator (before the first addPoints); the call sequence of the reference's nodelet stays as it is
  void setDevices(const std::vector<int> &ids) { status_ = mcc_set_devices(h_, (int)ids.size(), ids.data()); }
  void setVerbose(bool v) { verbose_ = v; }
  mcc_handle *handle() { return h_; }
  std::vector<double> params() const {
    std::vector<double> p((size_t)mcc_num_params(h_));
    if (mcc_get_params(h_, p.data(), p.size()) != MCC_OK) p.clear();
    return p;
  }

 private:
  void printReport() const {
    static const char *why[] = {"Gradient tolerance reached.", "Parameter tolerance reached.", "Function tolerance reached.", "Maximum number of iterations reached.",
                                "Number of successive invalid steps more than Solver::Options::max_num_consecutive_invalid_steps.",
                                "Minimum trust region radius reached.", "Linear solver failure.", "No data."};
    static const char *term[] = {"CONVERGENCE", "NO_CONVERGENCE", "FAILURE"};
    const mcc_summary &s = summary_;
    std::printf("\nSolver Summary (multicam_calibration_b200: %s)\n\n", mcc_version());
    std::printf("Parameters                          %8d   (reduced %d)\nResidual                            %8lld\n\n", s.num_params, s.num_reduced_params, s.num_residuals);
    std::printf("Cost:\nInitial                          %e\nFinal                            %e\nChange                           %e\n\n", s.initial_cost, s.final_cost,
                s.initial_cost - s.final_cost);
    std::printf("Minimizer iterations                 %8d\nSuccessful steps                     %8d\nUnsuccessful steps                   %8d\n\n", s.num_iterations,
                s.num_successful, s.num_unsuccessful);
    std::printf("Time (in seconds):\n  Upload / structure                 %.4f\n  Jacobian evaluation + assembly     %.4f\n  Linear solver + candidate cost     %.4f\nTotal                                %.4f\n\n",
                s.t_upload, s.t_jacobian, s.t_linear, s.t_total);
    if (s.why == 1) std::printf("Termination:                      %s (%s Relative step_norm: %e <= 1.000000e-12.)\n\n", term[s.termination], why[s.why], s.why_value);
    else if (s.why == 2) std::printf("Termination:                      %s (%s |cost_change|/cost: %e <= 1.000000e-12)\n\n", term[s.termination], why[s.why], s.why_value);
    else std::printf("Termination:                      %s (%s)\n\n", term[s.termination], why[s.why]);
  }

  mcc_handle *h_ = nullptr;
  CalibDataVec calibrationData_;
  std::vector<std::vector<int>> npts_;   // [frame][cam]
  mcc_summary summary_{};
  mcc_status status_ = MCC_OK;
  bool verbose_ = true;
};

}  // namespace multicam_calibration
#endif
