// cli.cpp -- mcc_calibrate: the offline calibration flow of the reference's nodelet without ROS.
//
// Reproduces CalibrationNodelet::onInit with `corners_file` + `run_calib_on_init`
// (/root/reference/src/calibration_nodelet.cpp:53-126) and calibrate() (:196-246): parse the Kalibr YAML, read
// corners.csv, add every frame with its initial cam0 pose, run the calibration (optionally remove outliers and run
// again, :202-207), write the result YAML to a file and to stdout (:413-453), feed the results back as the
// cameras (updateCameras, :455-469) and print the reprojection report with per-camera residual dumps (:292-345).
//
// Initial poses: computed per frame like the reference's guessCameraPose (:510-543, homography of the first camera
// that saw the board, get_init_pose.h), or -- with --init-poses -- read from a file: one line per frame rank,
// 16 numbers (row-major 4x4 T_c0_w); frames without a pose line or with "nan" are skipped like frames where the
// reference finds no detections (:638-641).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include "multicam_calibration/calibrator.h"
#include "multicam_calibration/corners_csv.h"
#include "multicam_calibration/get_init_pose.h"
#include "multicam_calibration/kalibr_yaml.h"

using namespace multicam_calibration;

static void usage() {
  std::fprintf(stderr,
               "usage: mcc_calibrate --calib initial.yaml --corners corners.csv [--init-poses poses.txt] [--output out.yaml]\n"
               "                     [--outlier-pixel-threshold px] [--fix-intrinsics] [--device n | --devices 0,1,...] [--residuals-dir dir]\n"
               "                     [--latest-link name.yaml] [--quiet] [--init-poses-out poses.txt] [--pose-threads n] [--dry-run]\n"
               "                     [--stage 'cam0:fix_intrinsics=0,fix_extrinsics=1,active=1;cam1:active=0'] ...\n"
               "  every --stage sets camera flags (the nodelet's set_parameter service) and runs the calibrate() sequence once;\n"
               "  the result of a stage is the starting point of the next (example_calib_manager.py); no --stage: one run\n"
               "  --init-poses-out writes the initial cam0 poses (the format --init-poses reads); --dry-run stops after them\n");
}

// calibration_nodelet.cpp:292-345
static std::string reprojection_report(Calibrator &cal, const CalibDataVec &calib, const std::vector<CamWorldPoints> &worldPoints,
                                       const std::vector<CamImagePoints> &imagePoints, const std::vector<size_t> &frame_of_added,
                                       const std::string &residuals_dir) {
  Calibrator::Residuals res;
  cal.testCalibration(&res);
  if (res.empty()) { std::printf("no residuals found for test!\n"); return "calibration failed: no data found!"; }
  double tot = 0; size_t ntot = 0; double maxErr = 0; size_t maxFrame = 0, maxCam = 0;
  std::vector<double> camSum(calib.size(), 0.0);
  std::vector<size_t> camCnt(calib.size(), 0);
  std::vector<std::ofstream> files;
  if (!residuals_dir.empty())
    for (size_t k = 0; k < calib.size(); k++) files.emplace_back(residuals_dir + "/residuals_cam" + std::to_string(k) + ".txt");
  for (size_t f = 0; f < res.size(); f++)
    for (size_t k = 0; k < res[f].size(); k++)
      for (size_t i = 0; i < res[f][k].size(); i++) {
        const Calibrator::Vec2d v = res[f][k][i];
        const double err = v(0) * v(0) + v(1) * v(1);
        tot += err; ntot++; camSum[k] += err; camCnt[k]++;
        if (!files.empty()) {
          // the reference indexes the points by residual index (valid only while no point is masked)
          const auto &wp = worldPoints[k][frame_of_added[f]][i];
          const auto &ip = imagePoints[k][frame_of_added[f]][i];
          const double dx = ip.x - calib[k].intrinsics.intrinsics[2], dy = ip.y - calib[k].intrinsics.intrinsics[3];
          files[k] << wp.x << " " << wp.y << " " << ip.x << " " << ip.y << " " << dx << " " << dy << " " << std::sqrt(dx * dx + dy * dy) << " "
                   << std::sqrt(err) << " " << v(0) << " " << v(1) << std::endl;
        }
        if (err > maxErr) { maxErr = err; maxFrame = f; maxCam = k; }
      }
  std::cout << "----------------- reprojection errors: ---------------" << std::endl;
  std::cout << "total error:     " << std::sqrt(ntot ? tot / ntot : 0.0) << " px" << std::endl;
  std::stringstream ss;
  ss << "calib proj err: ";
  for (size_t k = 0; k < calib.size(); k++) {
    const double px = std::sqrt(camCnt[k] ? camSum[k] / camCnt[k] : 0.0);
    std::cout << "avg error cam " << k << ": " << px << " px" << std::endl;
    ss << "cam" << k << ": " << px << "px ";
  }
  std::cout << "max error: " << std::sqrt(maxErr) << " px at frame: " << maxFrame << " for cam: " << maxCam << std::endl;
  return ss.str();
}

int main(int argc, char **argv) {
  std::string calib, corners, poses_file, poses_out, output, residuals_dir, latest_link;
  bool dry_run = false;
  int pose_threads = 0;
  std::vector<std::string> stages;
  double outlier = -1.0;
  bool fix_intr = false, quiet = false;
  int device = -1;
  std::vector<int> devices;      // --devices 0,1,...: several GPUs behind the one Calibrator
  for (int i = 1; i < argc; i++) {
    const std::string a = argv[i];
    auto next = [&]() -> std::string { if (i + 1 >= argc) { usage(); std::exit(2); } return argv[++i]; };
    if (a == "--calib") calib = next();
    else if (a == "--corners") corners = next();
    else if (a == "--init-poses") poses_file = next();
    else if (a == "--init-poses-out") poses_out = next();
    else if (a == "--pose-threads") pose_threads = std::atoi(next().c_str());
    else if (a == "--dry-run") dry_run = true;
    else if (a == "--output") output = next();
    else if (a == "--outlier-pixel-threshold") outlier = std::atof(next().c_str());
    else if (a == "--fix-intrinsics") fix_intr = true;
    else if (a == "--device") device = std::atoi(next().c_str());
    else if (a == "--devices") {
      const std::string l = next();
      size_t p = 0;
      while (p < l.size()) { devices.push_back(std::atoi(l.c_str() + p)); p = l.find(',', p); if (p == std::string::npos) break; p++; }
    }
    else if (a == "--residuals-dir") residuals_dir = next();
    else if (a == "--latest-link") latest_link = next();
    else if (a == "--quiet") quiet = true;
    else if (a == "--stage") stages.push_back(next());
    else { usage(); return 2; }
  }
  if (calib.empty() || corners.empty()) { usage(); return 2; }

  CameraExtrinsics T_imu_body;
  CalibDataVec cameras;
  try {
    cameras = read_kalibr_yaml(calib, &T_imu_body);
  } catch (const std::exception &e) {
    std::fprintf(stderr, "error parsing cameras: %s\n", e.what());   // calibration_nodelet.cpp:57-62
    return 1;
  }
  std::printf("Found %zu cameras!\n", cameras.size());
  if (cameras.empty()) return 1;

  Calibrator cal;
  if (!devices.empty()) {
    cal.setDevices(devices);
    if (cal.lastStatus() != MCC_OK) { std::fprintf(stderr, "--devices: %s\n", cal.lastError()); return 1; }
  } else if (device >= 0) cal.setDevice(device);
  cal.setVerbose(!quiet);
  cal.setFixAllIntrinsics(fix_intr);
  cal.setCameras(cameras);
  if (cal.lastStatus() != MCC_OK) return 1;

  std::vector<CamWorldPoints> worldPoints;
  std::vector<CamImagePoints> imagePoints;
  if (!read_corners_csv(corners, (unsigned)cameras.size(), &worldPoints, &imagePoints)) {
    std::fprintf(stderr, "file not found or bad: %s\n", corners.c_str());
    return 1;
  }
  const size_t nframes = worldPoints[0].size();
  // initial poses
  std::vector<CameraExtrinsics> poses(nframes);
  std::vector<bool> have(nframes, false);
  if (!poses_file.empty()) {
    std::ifstream in(poses_file);
    if (!in.is_open()) { std::fprintf(stderr, "cannot open %s\n", poses_file.c_str()); return 1; }
    std::string line;
    size_t f = 0;
    while (std::getline(in, line) && f < nframes) {
      std::stringstream ss(line);
      double v[16];
      int n = 0;
      while (n < 16 && (ss >> v[n])) n++;
      if (n == 16 && std::isfinite(v[0])) { for (int i = 0; i < 16; i++) poses[f].data()[i] = v[i]; have[f] = true; }
      f++;
    }
  }
  // no file: the reference's guessCameraPose for every frame, all views at once over the host's threads (get_init_pose.h)
  if (poses_file.empty()) have = guessCameraPoses(cameras, worldPoints, imagePoints, &poses, pose_threads);
  if (!poses_out.empty()) {
    // what --init-poses reads back: one line per frame, 16 numbers row-major, "nan" for a frame without a pose
    std::ofstream po(poses_out);
    po.precision(17);
    for (size_t f = 0; f < nframes; f++) {
      if (!have[f]) { po << "nan\n"; continue; }
      for (int i = 0; i < 16; i++) po << poses[f].data()[i] << (i == 15 ? "\n" : " ");
    }
  }
  if (dry_run) { std::printf("dry run: initial poses of %zu frames, no solve\n", nframes); return 0; }
  std::vector<size_t> frame_of_added;
  for (size_t f = 0; f < nframes; f++) {
    CamWorldPoints wp; CamImagePoints ip;
    bool any = false;
    for (size_t k = 0; k < cameras.size(); k++) { wp.push_back(worldPoints[k][f]); ip.push_back(imagePoints[k][f]); any = any || !wp.back().empty(); }
    if (!any || !have[f]) { std::printf("no detections found, skipping frame %zu\n", f); continue; }
    cal.addPoints((int)f, wp, ip, poses[f]);
    frame_of_added.push_back(f);
  }
  std::printf("read %zu frames from %s\n", frame_of_added.size(), corners.c_str());

  if (stages.empty()) stages.push_back("");
  CalibDataVec results;
  for (size_t st = 0; st < stages.size(); st++) {
    // set_parameter service (calibration_nodelet.cpp:257-285): "<camera>:<flag>=<0|1>,...;<camera>:..."
    if (!stages[st].empty()) {
      std::stringstream all(stages[st]);
      std::string one;
      while (std::getline(all, one, ';')) {
        const size_t colon = one.find(':');
        if (colon == std::string::npos) { std::fprintf(stderr, "bad --stage entry: %s\n", one.c_str()); return 2; }
        const std::string cam = one.substr(0, colon);
        size_t idx = cameras.size();
        for (size_t k = 0; k < cameras.size(); k++) if (cameras[k].name == cam) idx = k;
        if (idx == cameras.size()) { std::fprintf(stderr, "invalid camera name: %s\n", cam.c_str()); return 2; }   // getCameraIndex (:248-255)
        std::stringstream fl(one.substr(colon + 1));
        std::string kv;
        while (std::getline(fl, kv, ',')) {
          const size_t eq = kv.find('=');
          if (eq == std::string::npos) { std::fprintf(stderr, "bad --stage flag: %s\n", kv.c_str()); return 2; }
          const std::string key = kv.substr(0, eq);
          const bool val = std::atoi(kv.substr(eq + 1).c_str()) != 0;
          if (key == "fix_intrinsics") cameras[idx].fixIntrinsics = val;
          else if (key == "fix_extrinsics") cameras[idx].fixExtrinsics = val;
          else if (key == "active") cameras[idx].active = val;
          else { std::fprintf(stderr, "invalid parameter: %s\n", key.c_str()); return 2; }
        }
      }
      cal.setCameras(cameras);
      if (stages.size() > 1) std::printf("--------- calibration stage %zu of %zu ---------\n", st + 1, stages.size());
    }
    // calibrate()  (calibration_nodelet.cpp:196-246)
    cal.showCameraStatus();
    cal.runCalibration();
    if (cal.lastStatus() != MCC_OK) return 1;
    if (outlier > 0) {
      cal.removeOutliers(outlier);
      std::printf("removed outliers, rerunning calibration!\n");
      cal.runCalibration();
      if (cal.lastStatus() != MCC_OK) return 1;
    }
    results = cal.getCalibrationResults();
    if (results.empty()) { std::fprintf(stderr, "empty calibration results, no file written!\n"); return 1; }
    std::string out_name = output;
    if (out_name.empty()) {
      char buf[64];
      std::time_t t = std::time(nullptr);
      std::strftime(buf, sizeof(buf), "%Y-%m-%d-%H-%M-%S", std::localtime(&t));   // make_filename (:19-26)
      out_name = std::string("calibration_output-") + buf + ".yaml";
    }
    std::printf("writing calibration to %s\n", out_name.c_str());
    {
      std::ofstream of(out_name);
      write_calibration(of, results, T_imu_body);
    }
    // updateCameras (:455-469): results become the next run's initial values; flags are kept
    for (size_t k = 0; k < cameras.size(); k++) {
      cameras[k].intrinsics = results[k].intrinsics;
      cameras[k].T_cn_cnm1 = results[k].T_cn_cnm1;
      cameras[k].T_cam_imu = results[k].T_cam_imu;
    }
    cal.setCameras(cameras);   // (:227) params_ of the finished run stay in place for the test below
    write_calibration(std::cout, results, T_imu_body);
    if (!latest_link.empty()) {
      // "ln -sf <file> <calib_dir>/<latest_link_name>" (calibration_nodelet.cpp:232-235)
      const std::string cmd = "ln -sf " + out_name + " " + latest_link;
      if (std::system(nullptr) && std::system(cmd.c_str())) std::fprintf(stderr, "link command failed: %s\n", cmd.c_str());
    }
  }
  const std::string msg = reprojection_report(cal, results, worldPoints, imagePoints, frame_of_added, residuals_dir);
  std::printf("%s\n", msg.c_str());
  return 0;
}
