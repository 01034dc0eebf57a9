// Test driver for include/multicam_calibration/stereo_rectify.h: reads camera pairs from stdin, one per line
//   <model> <w> <h> <intr1 x4> <intr2 x4> <nd> <dist1 x nd> <dist2 x nd> <T_cn_cnm1 x16>
// and prints R1 (9) R2 (9) P1 (12) P2 (12) per line with 17 significant digits.
#include <cstdio>
#include <iostream>
#include <string>

#include "multicam_calibration/stereo_rectify.h"

using namespace multicam_calibration;

int main() {
  std::string model;
  while (std::cin >> model) {
    CalibDataVec cams(2);
    int w, h, nd;
    std::cin >> w >> h;
    for (int k = 0; k < 2; k++) {
      cams[k].intrinsics.distortion_model = model;
      cams[k].intrinsics.resolution = {w, h};
      for (int i = 0; i < 4; i++) std::cin >> cams[k].intrinsics.intrinsics[i];
    }
    std::cin >> nd;
    for (int k = 0; k < 2; k++) {
      cams[k].intrinsics.distortion_coeffs.assign(nd, 0.0);
      for (int i = 0; i < nd; i++) std::cin >> cams[k].intrinsics.distortion_coeffs[i];
    }
    for (int i = 0; i < 16; i++) std::cin >> cams[1].T_cn_cnm1.data()[i];
    Rectification r;
    if (!compute_RPQ(0, 1, cams, &r)) { std::printf("none\n"); continue; }
    for (int k = 0; k < 2; k++) for (int i = 0; i < 9; i++) std::printf("%.17g ", r.R[k][i]);
    for (int k = 0; k < 2; k++) for (int i = 0; i < 12; i++) std::printf("%.17g ", r.P[k][i]);
    std::printf("\n");
  }
  return 0;
}
