// corners_csv.h -- reader for the detector's corners.csv checkpoint file.
// Restates get_intrinsics_csv_data / readPointsFromFile of /root/reference/src/calibration_nodelet.cpp:545-645:
// lines "<frame>, intrinsics, <cam>, <wx>, <wy>, <u>, <v>"; the four coordinates are parsed with %f into FLOAT and
// widened (so they are float32-exact); other lines are skipped; world z = 0; a run of lines with the same frame
// number forms one frame of that camera; frames are then ranked over the sorted union of frame numbers.
#ifndef MULTICAM_CALIBRATION_B200_CORNERS_CSV_H
#define MULTICAM_CALIBRATION_B200_CORNERS_CSV_H
#include <cstdio>
#include <fstream>
#include <iterator>
#include <set>
#include <string>
#include <vector>

#include "multicam_calibration/types.h"

namespace multicam_calibration {

// one camera's frames in file order
inline bool get_intrinsics_csv_data(const std::string &filename, unsigned int camera_id, std::vector<int> *frame_nums,
                                    CamWorldPoints &world_points, CamImagePoints &img_pts, int skip_factor = 0) {
  std::ifstream in(filename);
  if (!in.is_open()) return false;
  std::string line;
  int prev = -1;
  frame_nums->clear();
  while (std::getline(in, line)) {
    int frame_num;
    unsigned int cam_idx;
    float wx, wy, cx, cy;
    if (std::sscanf(line.c_str(), "%d, intrinsics, %u, %f, %f, %f, %f", &frame_num, &cam_idx, &wx, &wy, &cx, &cy) != 6) continue;
    if (cam_idx != camera_id) continue;
    if (frame_num % (skip_factor + 1) != 0) continue;
    if (frame_num != prev) {
      prev = frame_num;
      world_points.push_back(FrameWorldPoints());
      img_pts.push_back(FrameImagePoints());
      frame_nums->push_back(frame_num);
    }
    world_points.back().emplace_back(wx, wy, 0);
    img_pts.back().emplace_back(cx, cy);
  }
  std::printf("read %u frames from corners file for cam %u\n", (unsigned)frame_nums->size(), camera_id);
  return true;
}

// worldPoints[cam][frame_rank], imagePoints[cam][frame_rank]; frame ranks over the union of frame numbers
inline bool read_corners_csv(const std::string &fname, unsigned int ncam, std::vector<CamWorldPoints> *worldPoints,
                             std::vector<CamImagePoints> *imagePoints, std::vector<int> *frame_numbers = nullptr) {
  std::set<int> fnum_set;
  std::vector<std::vector<int>> fnums(ncam);
  std::vector<CamWorldPoints> wp(ncam);
  std::vector<CamImagePoints> ip(ncam);
  for (unsigned int cam = 0; cam < ncam; cam++) {
    if (!get_intrinsics_csv_data(fname, cam, &fnums[cam], wp[cam], ip[cam])) return false;
    fnum_set.insert(fnums[cam].begin(), fnums[cam].end());
  }
  std::printf("total number of frames: %zu\n", fnum_set.size());
  worldPoints->assign(ncam, CamWorldPoints(fnum_set.size()));
  imagePoints->assign(ncam, CamImagePoints(fnum_set.size()));
  for (unsigned int cam = 0; cam < ncam; cam++)
    for (size_t i = 0; i < fnums[cam].size(); i++) {
      const size_t rank = (size_t)std::distance(fnum_set.begin(), fnum_set.find(fnums[cam][i]));
      (*worldPoints)[cam][rank] = wp[cam][i];
      (*imagePoints)[cam][rank] = ip[cam][i];
    }
  if (frame_numbers) frame_numbers->assign(fnum_set.begin(), fnum_set.end());
  return true;
}

}  // namespace multicam_calibration
#endif
