// types.h -- point and container types of the Calibrator API.
// Same names and meaning as /root/reference/include/multicam_calibration/types.h:17-55, without Eigen
// (the reference only needs Eigen's aligned allocator there because its points can hold ceres::Jet).
#ifndef MULTICAM_CALIBRATION_B200_TYPES_H
#define MULTICAM_CALIBRATION_B200_TYPES_H
#include <ostream>
#include <utility>
#include <vector>

namespace multicam_calibration {

template <typename T>
struct Point2 {
  T x, y;
  Point2() : x(), y() {}
  Point2(T x_, T y_) : x(x_), y(y_) {}
};
template <typename T>
std::ostream &operator<<(std::ostream &os, const Point2<T> &p) { return os << "(" << p.x << ", " << p.y << ")"; }

template <typename T>
struct Point3 {
  T x, y, z;
  Point3() : x(), y(), z() {}
  Point3(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
};
template <typename T>
std::ostream &operator<<(std::ostream &os, const Point3<T> &p) { return os << "(" << p.x << ", " << p.y << ", " << p.z << ")"; }

// the C ABI takes these arrays as flat doubles
static_assert(sizeof(Point2<double>) == 2 * sizeof(double), "Point2<double> must be two packed doubles");
static_assert(sizeof(Point3<double>) == 3 * sizeof(double), "Point3<double> must be three packed doubles");

using FrameWorldPoints = std::vector<Point3<double>>;   // one camera, one frame
using FrameImagePoints = std::vector<Point2<double>>;
using CamWorldPoints = std::vector<FrameWorldPoints>;   // indexed by camera (addPoints) or by frame (storage)
using CamImagePoints = std::vector<FrameImagePoints>;
using Pose = std::pair<Point3<double>, Point3<double>>;

}  // namespace multicam_calibration
#endif
