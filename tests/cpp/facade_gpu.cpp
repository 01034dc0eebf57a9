// GPU test driver for the C++ facade (include/multicam_calibration/calibrator.h): reads a rig from a text file written by
// tests/test_cpp_facade.py, feeds it through the reference's own call sequence (setCameras, addPoints per frame,
// runCalibration, testCalibration, getCalibrationResults -- calibration_nodelet.cpp:196-208, 294, 642) and prints what
// came back so that the Python side can compare it with the oracle.
//
//   rig file:  ncam nframes
//              per camera: model n_dist fx fy cx cy d0..d{n-1}  T(16)  fix_intr fix_extr active
//              per frame:  T_c0_w(16)  then per camera: n  followed by n lines "x y z u v"
//   usage:     facade_gpu <rig file> [<devices: - or e.g. 0,1> [run]]
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <string>

#include "multicam_calibration/calibrator.h"

using namespace multicam_calibration;

int main(int argc, char **argv) {
  if (argc < 2) return 2;
  std::ifstream in(argv[1]);
  int ncam = 0, nframes = 0;
  in >> ncam >> nframes;
  CalibDataVec cams(ncam);
  for (int k = 0; k < ncam; k++) {
    int model, nd;
    in >> model >> nd;
    CalibrationData &c = cams[k];
    c.name = "cam" + std::to_string(k);
    c.intrinsics.distortion_model = model == 0 ? "equidistant" : "radtan";
    c.intrinsics.intrinsics.resize(4);
    c.intrinsics.distortion_coeffs.resize(nd);
    for (int i = 0; i < 4; i++) in >> c.intrinsics.intrinsics[i];
    for (int i = 0; i < nd; i++) in >> c.intrinsics.distortion_coeffs[i];
    for (int i = 0; i < 16; i++) in >> c.T_cn_cnm1.data()[i];
    int fi, fe, ac;
    in >> fi >> fe >> ac;
    c.fixIntrinsics = fi; c.fixExtrinsics = fe; c.active = ac;
  }
  Calibrator cal;
  cal.setVerbose(false);
  const std::string devs = argc > 2 ? argv[2] : "-";
  if (devs != "-") {
    std::vector<int> ids;
    size_t p = 0;
    while (p < devs.size()) { ids.push_back(std::atoi(devs.c_str() + p)); p = devs.find(',', p); if (p == std::string::npos) break; p++; }
    cal.setDevices(ids);
    if (cal.lastStatus() != MCC_OK) { std::printf("SETDEVICES_FAILED %s\n", cal.lastError()); return 1; }
  }
  cal.setCameras(cams);
  for (int f = 0; f < nframes; f++) {
    CameraExtrinsics T;
    for (int i = 0; i < 16; i++) in >> T.data()[i];
    CamWorldPoints wp(ncam);
    CamImagePoints ip(ncam);
    for (int k = 0; k < ncam; k++) {
      int n = 0;
      in >> n;
      for (int i = 0; i < n; i++) {
        double x, y, z, u, v;
        in >> x >> y >> z >> u >> v;
        wp[k].emplace_back(x, y, z);
        ip[k].emplace_back(u, v);
      }
    }
    cal.addPoints(f, wp, ip, T);
    if (cal.lastStatus() != MCC_OK) { std::printf("ADDPOINTS_FAILED %s\n", cal.lastError()); return 1; }
  }
  const bool run = argc > 3 && std::string(argv[3]) == "run";
  if (run) {
    cal.runCalibration();
    if (cal.lastStatus() != MCC_OK) { std::printf("RUN_FAILED %s\n", cal.lastError()); return 1; }
    const mcc_summary &s = cal.summary();
    std::printf("SUMMARY %.17g %.17g %d %d %d %d\n", s.initial_cost, s.final_cost, s.num_iterations, s.num_successful, s.num_unsuccessful, s.why);
    std::vector<double> p = cal.params();
    std::printf("PARAMS %zu", p.size());
    for (double v : p) std::printf(" %.17g", v);
    std::printf("\n");
  } else {
    if (mcc_init_params(cal.handle()) != MCC_OK) { std::printf("INIT_FAILED\n"); return 1; }
  }
  Calibrator::Residuals res;
  cal.testCalibration(&res);
  if (cal.lastStatus() != MCC_OK) { std::printf("TEST_FAILED %s\n", cal.lastError()); return 1; }
  for (size_t f = 0; f < res.size(); f++)
    for (size_t k = 0; k < res[f].size(); k++) {
      std::printf("RES %zu %zu %zu", f, k, res[f][k].size());
      for (const auto &r : res[f][k]) std::printf(" %.17g %.17g", r(0), r(1));
      std::printf("\n");
    }
  std::printf("FACADE_GPU_OK\n");
  return 0;
}
