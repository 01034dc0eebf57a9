// CPU test of the C++ facade (include/multicam_calibration/*.h): Kalibr YAML in/out, corners.csv reader with the
// reference's float parse, CameraExtrinsics algebra, Calibrator host-side behaviour.  No GPU needed: compute calls
// are expected to fail loudly (MCC_ERR_NO_DEVICE) when none is present.
#include <cassert>
#include <cmath>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <sstream>

#include "multicam_calibration/calibrator.h"
#include "multicam_calibration/corners_csv.h"
#include "multicam_calibration/get_init_pose.h"
#include "multicam_calibration/kalibr_yaml.h"

using namespace multicam_calibration;

#define CHECK(c)                                                              \
  do {                                                                        \
    if (!(c)) { std::printf("CHECK FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } \
  } while (0)

int main(int argc, char **argv) {
  const std::string dir = argc > 1 ? argv[1] : "/tmp";
  // ---- YAML in (README.md:35-51 style, plus flags and a comment)
  const std::string y = dir + "/in.yaml";
  {
    std::ofstream o(y);
    o << "cam0:\n  camera_model: pinhole\n  intrinsics: [605.054, 604.66, 641.791, 508.728]\n  distortion_model: equidistant\n"
         "  distortion_coeffs: [-0.0146915, 0.000370396, -0.00425216, 0.0015107]\n  resolution: [1280, 1024]\n  rostopic: /rig/left/image_mono\n"
         "  T_cn_cnm1:\n  - [0.5, 0.0, 0.0, 0.0]\n  - [0.0, 1.0, 0.0, 0.0]\n  - [0.0, 0.0, 1.0, 0.0]\n  - [0.0, 0.0, 0.0, 1.0]\n"
         "cam1:\n  T_cn_cnm1:\n  - [ 0.99999965648, -0.00013331925,  0.00081808159, -0.19946344647]\n  - [ 0.00013388601,  0.99999975107, -0.00069277676, -0.00005674605]\n"
         "  - [-0.00081798903,  0.00069288605,  0.99999942540,  0.00010022941]\n  - [ 0.00000000000,  0.00000000000,  0.00000000000,  1.00000000000]\n"
         "  camera_model: pinhole\n  intrinsics: [605.097, 604.321, 698.772, 573.558]\n  distortion_model: radtan # comment\n"
         "  distortion_coeffs: [-0.0146155, -0.00291494, -0.000681981, 0.000221855, 0.1]\n  resolution: [1280, 1024]\n  rostopic: /rig/right/image_mono\n"
         "  fix_intrinsics: true\n  active: false\n"
         "cam2:\n  camera_model: pinhole\n  intrinsics: [600.0, 600.0, 640.0, 512.0]\n  distortion_model: equidistant\n  distortion_coeffs: [0.0, 0.0, 0.0, 0.0]\n"
         "  resolution: [1280, 1024]\n  rostopic: /c2\n  T_cn_cnm1:\n  - [1.0, 0.0, 0.0, 0.1]\n  - [0.0, 1.0, 0.0, 0.0]\n  - [0.0, 0.0, 1.0, 0.0]\n  - [0.0, 0.0, 0.0, 1.0]\n"
         "cam4:\n  camera_model: never_reached\n"
         "T_imu_body:\n  - [1.0, 0.0, 0.0, 0.0]\n  - [0.0, 1.0, 0.0, 0.0]\n  - [0.0, 0.0, 1.0, 0.5]\n  - [0.0, 0.0, 0.0, 1.0]\n";
  }
  CameraExtrinsics Tib;
  CalibDataVec cams = read_kalibr_yaml(y, &Tib);
  CHECK(cams.size() == 3);                               // cam3 is missing: parsing stops there
  CHECK(cams[0].T_cn_cnm1 == identity());                // cam0 forced to identity
  CHECK(cams[1].intrinsics.distortion_model == "radtan" && cams[1].intrinsics.distortion_coeffs.size() == 5);
  CHECK(cams[1].fixIntrinsics && !cams[1].active && !cams[1].fixExtrinsics && cams[0].active);
  CHECK(std::fabs(cams[1].T_cn_cnm1(0, 3) + 0.19946344647) < 1e-15);
  CHECK(cams[2].T_cn_cnm1(1, 1) == std::cos(1e-5) && cams[2].T_cn_cnm1(2, 1) == std::sin(1e-5));   // identity rotation perturbed
  CHECK(cams[2].T_cn_cnm1(0, 3) == 0.1);
  CHECK(Tib(2, 3) == 0.5 && isNonZero(Tib) && !isNonZero(cams[0].T_cam_imu));
  CHECK(cams[0].intrinsics.resolution[0] == 1280 && cams[0].rostopic == "/rig/left/image_mono");
  // integer entries in a transform are not "doubles" for the reference: default is used
  {
    std::ofstream o(dir + "/int.yaml");
    o << "cam0:\n  camera_model: pinhole\n  intrinsics: [1.0, 1.0, 1.0, 1.0]\n  distortion_model: radtan\n  distortion_coeffs: [0.0]\n  resolution: [2, 2]\n  rostopic: /a\n"
         "cam1:\n  camera_model: pinhole\n  intrinsics: [1.0, 1.0, 1.0, 1.0]\n  distortion_model: radtan\n  distortion_coeffs: [0.0]\n  resolution: [2, 2]\n  rostopic: /b\n"
         "  T_cn_cnm1:\n  - [0, 1, 0, 0]\n  - [1, 0, 0, 0]\n  - [0, 0, 1, 0]\n  - [0, 0, 0, 1]\n";
  }
  CalibDataVec ci = read_kalibr_yaml(dir + "/int.yaml");
  CHECK(ci.size() == 2 && ci[1].T_cn_cnm1(0, 0) == 1.0 && ci[1].T_cn_cnm1(1, 1) == std::cos(1e-5));
  // missing required field
  {
    std::ofstream o(dir + "/bad.yaml");
    o << "cam0:\n  camera_model: pinhole\n  intrinsics: [1.0, 1.0, 1.0, 1.0]\n";
  }
  bool threw = false;
  try { read_kalibr_yaml(dir + "/bad.yaml"); } catch (const std::exception &e) { threw = std::string(e.what()).find("cannot find distortion_model for cam cam0") != std::string::npos; }
  CHECK(threw);

  // ---- YAML out: format of writeCalibration and round trip
  std::stringstream ss;
  write_calibration(ss, cams, Tib);
  const std::string out = ss.str();
  CHECK(out.find("cam1:\n  T_cn_cnm1:\n  - [ 0.99999965648, -0.00013331925,  0.00081808159, -0.19946344647] \n") != std::string::npos);
  CHECK(out.find("  intrinsics: [605.054, 604.66, 641.791, 508.728]\n") != std::string::npos);
  CHECK(out.find("  distortion_coeffs: [-0.0146915, 0.000370396, -0.00425216, 0.0015107]\n") != std::string::npos);
  CHECK(out.find("  resolution: [1280, 1024]\n  rostopic: /rig/left/image_mono\n") != std::string::npos);
  CHECK(out.find("cam0:\n  camera_model: pinhole\n") != std::string::npos);   // no T_cn_cnm1 / T_cam_imu for cam0
  CHECK(out.find("T_imu_body:\n  - [ 1.00000000000,  0.00000000000,  0.00000000000,  0.00000000000] \n") != std::string::npos);
  { std::ofstream o(dir + "/out.yaml"); o << out; }
  CalibDataVec back = read_kalibr_yaml(dir + "/out.yaml");
  CHECK(back.size() == 3);
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) CHECK(std::fabs(back[1].T_cn_cnm1(i, j) - cams[1].T_cn_cnm1(i, j)) < 1e-11);

  // ---- stereo pair of the same model: rectification entries for cameras 0 and 1 (calibration_nodelet.cpp:428-439);
  // values are cv2.fisheye.stereoRectify's for the README rig (tests/golden/rectify_golden.json, case 0)
  {
    CalibDataVec st(cams.begin(), cams.begin() + 2);
    st[1].intrinsics.distortion_model = "equidistant";
    st[1].intrinsics.distortion_coeffs = {-0.0146155, -0.00291494, -0.000681981, 0.000221855};
    std::stringstream s2;
    write_calibration(s2, st);
    const std::string o2 = s2.str();
    CHECK(o2.find("cam0:\n  rectification_matrix:\n  - [ 0.99999993889,  0.00015082") != std::string::npos);
    CHECK(o2.find("  projection_matrix:\n  - [439.5640") != std::string::npos);
    CHECK(o2.find(",  0.00000000000, 712.866") != std::string::npos);
    CHECK(o2.find("-87.67696") != std::string::npos);
    CHECK(o2.find("cam1:\n  T_cn_cnm1:\n") != std::string::npos && o2.find("cam1:\n  T_cn_cnm1:\n") < o2.rfind("  rectification_matrix:"));
  }

  // ---- CameraExtrinsics
  CameraExtrinsics P = cams[1].T_cn_cnm1 * cams[2].T_cn_cnm1, Pi = P.inverse(), I = P * Pi;
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) CHECK(std::fabs(I(i, j) - (i == j)) < 1e-14);

  // ---- corners.csv: float parse, foreign lines skipped, frame ranking over the union, empty (cam, frame) cells
  const std::string c = dir + "/corners.csv";
  {
    std::ofstream o(c);
    o << "10, intrinsics, 0, 0.100000, 0.200000, 100.123456, 50.500000\n"
         "10, intrinsics, 0, 0.300000, 0.200000, 101.000001, 51.250000\n"
         "10, extrinsics, 0, 1, 2, 3\n"
         "12, intrinsics, 1, -0.050000, 0.000000, 7.700000, 8.800000\n"
         "garbage line\n"
         "15, intrinsics, 0, 0.000000, 0.000000, 1.000000, 2.000000\n"
         "15, intrinsics, 1, 0.000000, 0.050000, 3.000000, 4.000000\n";
  }
  std::vector<CamWorldPoints> wp;
  std::vector<CamImagePoints> ip;
  std::vector<int> fn;
  CHECK(read_corners_csv(c, 2, &wp, &ip, &fn));
  CHECK(fn.size() == 3 && fn[0] == 10 && fn[1] == 12 && fn[2] == 15);
  CHECK(wp[0][0].size() == 2 && wp[0][1].empty() && wp[0][2].size() == 1 && wp[1][0].empty() && wp[1][1].size() == 1 && wp[1][2].size() == 1);
  CHECK(ip[0][0][0].x == (double)100.123456f && ip[0][0][0].x != 100.123456);   // float32-exact, as the reference reads it
  CHECK(wp[0][0][1].x == (double)0.3f && wp[0][0][1].z == 0.0);
  CHECK(!read_corners_csv(dir + "/does_not_exist.csv", 2, &wp, &ip));

  // ---- Calibrator host side
  Calibrator cal;
  cal.setVerbose(false);
  cal.runCalibration();
  CHECK(cal.lastStatus() == MCC_ERR_NO_DATA);            // "no data to run on!"
  CHECK(cal.getCalibrationResults().empty());
  cal.setCameras(cams);
  CHECK(cal.lastStatus() == MCC_OK);
  CamWorldPoints w1(1);
  CamImagePoints i1(1);
  cal.addPoints(0, w1, i1, identity());
  CHECK(cal.lastStatus() == MCC_ERR_BAD_ARG);            // "number of cams != width of array!"
  CamWorldPoints w3(3);
  CamImagePoints i3(3);
  for (int k = 0; k < 3; k++)
    for (int p = 0; p < 5; p++) { w3[k].emplace_back(0.01 * p, 0.02 * k, 0.0); i3[k].emplace_back(600.0 + p, 500.0 + k); }
  CameraExtrinsics T = identity();
  T(2, 3) = 0.5;
  cal.addPoints(0, w3, i3, T);
  cal.addPoints(1, w3, i3, T);
  CHECK(cal.lastStatus() == MCC_OK);
  CHECK(mcc_num_frames(cal.handle()) == 2 && mcc_num_observations(cal.handle()) == 30);
  CHECK(mcc_init_params(cal.handle()) == MCC_OK);
  std::vector<double> p = cal.params();
  CHECK(p.size() == 8 + 9 + 8 + 12 + 12);
  CHECK(p[0] == 605.054 && p[8 + 4 + 4] == 0.1 /* cam1 5th coeff */ && std::fabs(p[25 + 3] + 0.19946344647) < 1e-15);
  CHECK(p[37 + 5] == 0.5 && p[37 + 0] == 0.0);           // frame pose: identity rotation, t_z = 0.5
  CalibDataVec res = cal.getCalibrationResults();
  CHECK(res.size() == 3);
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) CHECK(std::fabs(res[2].T_cn_cnm1(i, j) - cams[2].T_cn_cnm1(i, j)) < 1e-12);
  CalibrationData bad = cams[0];
  bad.intrinsics.distortion_model = "fisheye624";
  CalibDataVec bv(1, bad);
  Calibrator cal2;
  cal2.setCameras(bv);
  CHECK(cal2.lastStatus() == MCC_ERR_BAD_MODEL);
  {
    // guessCameraPose (calibration_nodelet.cpp:510-543) frame by frame against guessCameraPoses over all frames at once:
    // two undistorted pinhole cameras 0.2 m apart looking at a 7 x 5 board; camera 0 misses the board in every third frame
    CalibDataVec two(2);
    for (int k = 0; k < 2; k++) {
      two[k].name = "cam" + std::to_string(k);
      two[k].intrinsics.distortion_model = k ? "radtan" : "equidistant";
      two[k].intrinsics.intrinsics = {600.0, 600.0, 640.0, 512.0};
      two[k].intrinsics.distortion_coeffs = {0.0, 0.0, 0.0, 0.0};
      two[k].T_cn_cnm1 = identity();
    }
    two[1].T_cn_cnm1(0, 3) = -0.2;
    const int NF = 40;
    std::vector<CamWorldPoints> W(2, CamWorldPoints(NF));
    std::vector<CamImagePoints> I(2, CamImagePoints(NF));
    for (int f = 0; f < NF; f++) {
      const double a = 0.02 * f - 0.3, tz = 1.0 + 0.01 * f;          // board rotated about y, in front of the cameras
      for (int k = 0; k < 2; k++) {
        if (k == 0 && f % 3 == 2) continue;
        for (int i = 0; i < 35; i++) {
          const double X = 0.05 * (i % 7) - 0.15, Y = 0.05 * (i / 7) - 0.1;
          const double xc = std::cos(a) * X + (k ? -0.2 : 0.0), yc = Y, zc = -std::sin(a) * X + tz;
          double u = xc / zc, v = yc / zc;
          if (k == 0) { const double r = std::sqrt(u * u + v * v), th = std::atan(r); const double sc = r > 0 ? th / r : 1.0; u *= sc; v *= sc; }
          W[k][f].emplace_back(X, Y, 0.0);
          I[k][f].emplace_back(600.0 * u + 640.0, 600.0 * v + 512.0);
        }
      }
    }
    std::vector<CameraExtrinsics> batch;
    const std::vector<bool> have = guessCameraPoses(two, W, I, &batch, 3);
    CHECK(have.size() == (size_t)NF && batch.size() == (size_t)NF);
    for (int f = 0; f < NF; f++) {
      CamWorldPoints wp = {W[0][f], W[1][f]};
      CamImagePoints ip = {I[0][f], I[1][f]};
      CameraExtrinsics one;
      const bool ok = guessCameraPose(two, wp, ip, &one, f);
      CHECK(ok && have[f]);
      for (int i = 0; i < 16; i++) CHECK(one.data()[i] == batch[f].data()[i]);      // the same arithmetic, bit for bit
      const double a = 0.02 * f - 0.3;
      CHECK(std::fabs(one(0, 0) - std::cos(a)) < 1e-4 && std::fabs(one(2, 0) + std::sin(a)) < 1e-4 && std::fabs(one(2, 3) - (1.0 + 0.01 * f)) < 1e-3 && std::fabs(one(0, 3)) < 1e-3);
    }
  }
  std::printf("FACADE_OK\n");
  return 0;
}
