// kalibr_yaml.h -- Kalibr-style calibration YAML in and out.
//
// In : what the reference loads onto the ROS parameter server and reads in parseCameras
//      (/root/reference/src/calibration_nodelet.cpp:143-194): cam0, cam1, ... until one is missing, each with
//      camera_model, intrinsics, distortion_model, distortion_coeffs, resolution, rostopic (all required) and
//      optional T_cn_cnm1, T_cam_imu (4 rows of 4 values that must all be doubles, :35-47; anything else falls
//      back to the default, :133-141), fix_intrinsics, fix_extrinsics, active, flip_code, rotate_code; top-level
//      T_imu_body.  cam0's T_cn_cnm1 is forced to identity (:171-175); an exact-identity rotation for cam>0 is
//      replaced by a 1e-5 rad rotation about x (:176-183) because the rotation Jacobian vanishes at identity.
// Out: writeCalibration (:413-453) with print_tf (:360-371) and vec2str (:347-358) formatting, including the stereo
//      rectification_matrix / projection_matrix entries of cameras 0 and 1 (compute_RPQ, stereo_rectify.h).
//
// The parser handles the YAML subset these files use: nested block mappings by indentation, scalars, flow
// sequences "[a, b]" and block sequences of flow sequences ("- [a, b, c, d]").
#ifndef MULTICAM_CALIBRATION_B200_KALIBR_YAML_H
#define MULTICAM_CALIBRATION_B200_KALIBR_YAML_H
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "multicam_calibration/calibration_data.h"
#include "multicam_calibration/stereo_rectify.h"

namespace multicam_calibration {
namespace yaml_detail {

struct Node {
  std::string scalar;                              // "key: value"
  std::vector<std::string> list;                   // "key: [a, b]"
  std::vector<std::vector<std::string>> rows;      // "key:" followed by "- [a, b]" lines
  std::vector<std::pair<std::string, Node>> kids;  // nested mapping
  bool has_list = false;
  const Node *find(const std::string &k) const {
    for (const auto &kv : kids) if (kv.first == k) return &kv.second;
    return nullptr;
  }
};

inline std::string trim(const std::string &s) {
  size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
  return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
}
inline std::string unquote(const std::string &s) {
  if (s.size() >= 2 && ((s.front() == '\'' && s.back() == '\'') || (s.front() == '"' && s.back() == '"'))) return s.substr(1, s.size() - 2);
  return s;
}
inline std::string strip_comment(const std::string &s) {
  bool sq = false, dq = false;
  for (size_t i = 0; i < s.size(); i++) {
    if (s[i] == '\'' && !dq) sq = !sq;
    else if (s[i] == '"' && !sq) dq = !dq;
    else if (s[i] == '#' && !sq && !dq && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '\t')) return s.substr(0, i);
  }
  return s;
}
inline std::vector<std::string> flow_list(const std::string &s) {
  std::vector<std::string> out;
  const size_t a = s.find('['), b = s.rfind(']');
  if (a == std::string::npos || b == std::string::npos || b < a) throw std::runtime_error("bad flow sequence: " + s);
  std::stringstream ss(s.substr(a + 1, b - a - 1));
  std::string item;
  while (std::getline(ss, item, ',')) { item = trim(item); if (!item.empty()) out.push_back(unquote(item)); }
  return out;
}

struct Line { int indent; std::string text; };

inline void parse_block(const std::vector<Line> &lines, size_t &i, int indent, Node &node) {
  while (i < lines.size() && lines[i].indent >= indent) {
    const Line &ln = lines[i];
    if (ln.indent > indent) throw std::runtime_error("unexpected indentation at: " + ln.text);
    if (ln.text[0] == '-') return;   // sequence items are consumed by the key that owns them
    const size_t colon = ln.text.find(':');
    if (colon == std::string::npos) throw std::runtime_error("expected 'key: value' at: " + ln.text);
    const std::string key = unquote(trim(ln.text.substr(0, colon)));
    const std::string val = trim(ln.text.substr(colon + 1));
    Node child;
    i++;
    if (!val.empty()) {
      if (val[0] == '[') { child.list = flow_list(val); child.has_list = true; }
      else child.scalar = unquote(val);
    } else if (i < lines.size() && lines[i].text[0] == '-' && lines[i].indent >= indent) {
      const int sind = lines[i].indent;
      while (i < lines.size() && lines[i].indent == sind && lines[i].text[0] == '-') {
        child.rows.push_back(flow_list(lines[i].text));
        i++;
      }
    } else if (i < lines.size() && lines[i].indent > indent) {
      parse_block(lines, i, lines[i].indent, child);
    }
    node.kids.push_back(std::make_pair(key, child));
  }
}

inline Node parse_file(const std::string &path) {
  std::ifstream in(path);
  if (!in.is_open()) throw std::runtime_error("cannot open " + path);
  std::vector<Line> lines;
  std::string raw;
  while (std::getline(in, raw)) {
    std::string s = strip_comment(raw);
    if (trim(s).empty() || trim(s) == "---") continue;
    int ind = 0;
    while (ind < (int)s.size() && s[ind] == ' ') ind++;
    lines.push_back(Line{ind, trim(s)});
  }
  Node root;
  size_t i = 0;
  if (!lines.empty()) parse_block(lines, i, lines[0].indent, root);
  return root;
}

inline bool looks_like_double(const std::string &s) {
  // XmlRpc only yields TypeDouble for literals with a '.', exponent, or inf/nan; "1" is an int and is rejected
  // by the reference's transform reader (calibration_nodelet.cpp:42-44)
  char *end = nullptr;
  std::strtod(s.c_str(), &end);
  if (end == s.c_str() || *end != 0) return false;
  return s.find_first_of(".eEnN") != std::string::npos;
}

inline CameraExtrinsics get_transform(const Node *n, const CameraExtrinsics &def, const std::string &what) {
  if (!n) return def;
  if (n->rows.size() != 4) { std::cerr << "invalid transform " << what << ", using default" << std::endl; return def; }
  CameraExtrinsics T;
  for (int i = 0; i < 4; i++) {
    if (n->rows[i].size() != 4) { std::cerr << "bad line for transform " << what << ", using default" << std::endl; return def; }
    for (int j = 0; j < 4; j++) {
      if (!looks_like_double(n->rows[i][j])) { std::cerr << "bad value for transform " << what << " (all entries must be doubles), using default" << std::endl; return def; }
      T(i, j) = std::strtod(n->rows[i][j].c_str(), nullptr);
    }
  }
  return T;
}
inline bool to_bool(const std::string &s, bool def) {
  if (s == "true" || s == "True" || s == "TRUE" || s == "1") return true;
  if (s == "false" || s == "False" || s == "FALSE" || s == "0") return false;
  return def;
}
}  // namespace yaml_detail

// parseCameras (calibration_nodelet.cpp:143-194)
inline CalibDataVec read_kalibr_yaml(const std::string &path, CameraExtrinsics *T_imu_body = nullptr) {
  using namespace yaml_detail;
  const Node root = parse_file(path);
  CalibDataVec cams;
  for (int idx = 0;; idx++) {
    const std::string cam = "cam" + std::to_string(idx);
    const Node *n = root.find(cam);
    if (!n) break;
    auto need = [&](const char *field) -> const Node & {
      const Node *f = n->find(field);
      if (!f) throw std::runtime_error(std::string("cannot find ") + field + " for cam " + cam);
      return *f;
    };
    CalibrationData cd;
    cd.name = cam;
    CameraIntrinsics &ci = cd.intrinsics;
    ci.camera_model = need("camera_model").scalar;
    ci.distortion_model = need("distortion_model").scalar;
    ci.distortion_coeffs.clear();
    for (const auto &s : need("distortion_coeffs").list) ci.distortion_coeffs.push_back(std::strtod(s.c_str(), nullptr));
    ci.intrinsics.clear();
    for (const auto &s : need("intrinsics").list) ci.intrinsics.push_back(std::strtod(s.c_str(), nullptr));
    ci.resolution.clear();
    for (const auto &s : need("resolution").list) ci.resolution.push_back(std::atoi(s.c_str()));
    cd.rostopic = need("rostopic").scalar;
    cd.T_cam_imu = get_transform(n->find("T_cam_imu"), zeros(), cam + "/T_cam_imu");
    cd.T_cn_cnm1 = get_transform(n->find("T_cn_cnm1"), identity(), cam + "/T_cn_cnm1");
    if (const Node *f = n->find("fix_intrinsics")) cd.fixIntrinsics = to_bool(f->scalar, false);
    if (const Node *f = n->find("fix_extrinsics")) cd.fixExtrinsics = to_bool(f->scalar, false);
    if (const Node *f = n->find("active")) cd.active = to_bool(f->scalar, true);
    if (const Node *f = n->find("flip_code")) cd.flipCode = std::atoi(f->scalar.c_str());
    if (const Node *f = n->find("rotate_code")) cd.rotateCode = std::atoi(f->scalar.c_str());
    if (idx == 0) {
      if (cd.T_cn_cnm1 != identity()) {
        std::cerr << "Cam0 had a non-identity T_cn_cnm1 specified!" << std::endl;
        cd.T_cn_cnm1 = identity();
      }
    } else {
      bool isI = true;
      for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) isI = isI && cd.T_cn_cnm1(i, j) == (i == j ? 1.0 : 0.0);
      if (isI) {
        std::cerr << cam << " cannot have an identity rotation in T_cn_cnm1. Perturb values as needed." << std::endl;
        const double a = 0.00001, c = std::cos(a), s = std::sin(a);   // AngleAxis(1e-5, UnitX)
        cd.T_cn_cnm1(1, 1) = c; cd.T_cn_cnm1(1, 2) = -s; cd.T_cn_cnm1(2, 1) = s; cd.T_cn_cnm1(2, 2) = c;
      }
    }
    cams.push_back(cd);
  }
  if (T_imu_body) *T_imu_body = get_transform(root.find("T_imu_body"), zeros(), "T_imu_body");
  return cams;
}
inline CalibDataVec CalibrationData::parse_cameras(const std::string &yaml_file) { return read_kalibr_yaml(yaml_file); }

// print_tf (calibration_nodelet.cpp:360-371)
inline void print_tf(std::ostream &os, const std::string &p, const CameraExtrinsics &tf) {
  std::ios::fmtflags f(os.flags());
  const std::streamsize prec = os.precision();
  os << std::fixed << std::setprecision(11);
  for (int i = 0; i < 4; i++) {
    os << p << "- [";
    for (int j = 0; j < 3; j++) os << std::setw(14) << tf(i, j) << ", ";
    os << std::setw(14) << tf(i, 3) << "] " << std::endl;
  }
  os.flags(f);
  os.precision(prec);
}
// the same for the 3x3 / 3x4 rectification matrices (row-major arrays)
inline void print_mat(std::ostream &os, const std::string &p, const double *m, int rows, int cols) {
  std::ios::fmtflags f(os.flags());
  const std::streamsize prec = os.precision();
  os << std::fixed << std::setprecision(11);
  for (int i = 0; i < rows; i++) {
    os << p << "- [";
    for (int j = 0; j < cols - 1; j++) os << std::setw(14) << m[i * cols + j] << ", ";
    os << std::setw(14) << m[i * cols + cols - 1] << "] " << std::endl;
  }
  os.flags(f);
  os.precision(prec);
}
// vec2str (calibration_nodelet.cpp:347-358): default ostream formatting, 6 significant digits
inline std::string vec2str(const std::vector<double> &v) {
  std::stringstream ss;
  ss << "[";
  for (size_t i = 0; i < v.size(); i++) ss << (i ? ", " : "") << v[i];
  ss << "]";
  return ss.str();
}
// writeCalibration (calibration_nodelet.cpp:413-453)
inline void write_calibration(std::ostream &os, const CalibDataVec &results, const CameraExtrinsics &T_imu_body = zeros()) {
  Rectification rp;
  bool hasRP = false;
  try { hasRP = compute_RPQ(0, 1, results, &rp); } catch (...) { hasRP = false; }
  for (size_t k = 0; k < results.size(); k++) {
    const CalibrationData &cd = results[k];
    os << cd.name << ":" << std::endl;
    if (isNonZero(cd.T_cam_imu)) { os << "  T_cam_imu:" << std::endl; print_tf(os, "  ", cd.T_cam_imu); }
    if (k > 0) { os << "  T_cn_cnm1:" << std::endl; print_tf(os, "  ", cd.T_cn_cnm1); }
    if (hasRP && k < 2) {
      os << "  rectification_matrix:" << std::endl;
      print_mat(os, "  ", rp.R[k], 3, 3);
      os << "  projection_matrix:" << std::endl;
      print_mat(os, "  ", rp.P[k], 3, 4);
    }
    const CameraIntrinsics &ci = cd.intrinsics;
    os << "  camera_model: " << ci.camera_model << std::endl;
    os << "  intrinsics: " << vec2str(ci.intrinsics) << std::endl;
    os << "  distortion_model: " << ci.distortion_model << std::endl;
    os << "  distortion_coeffs: " << vec2str(ci.distortion_coeffs) << std::endl;
    os << "  resolution: [" << ci.resolution[0] << ", " << ci.resolution[1] << "]" << std::endl;
    os << "  rostopic: " << cd.rostopic << std::endl;
  }
  if (isNonZero(T_imu_body)) { os << "T_imu_body:" << std::endl; print_tf(os, "  ", T_imu_body); }
}

}  // namespace multicam_calibration
#endif
