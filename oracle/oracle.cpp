// oracle/oracle.cpp
//
// TEST INFRASTRUCTURE ONLY -- NOT PART OF THE PRODUCT PATH.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load this library.  The product (libmcc.so) never links or calls it.
//
// CPU restatement of the multicam_calibration solve hot path:
//   * residual model       : /root/reference/src/calibrator.cpp:74-155 (FrameResidual),
//                            /root/reference/include/multicam_calibration/utils.h:29-39, 62-82, 101-130
//   * parameter packing    : calibrator.cpp:36-71, 157-170, 251-283
//   * problem structure    : calibrator.cpp:353-370, 466-523 (one Huber(1.0) block per frame, constant blocks)
//   * result unpacking     : calibrator.cpp:415-463
//   * residual-only passes : calibrator.cpp:316-351, 372-413
//   * solver               : Ceres Solver (third party, NOT in /root/reference, version unpinned in
//                            package.xml:27; the logs were made with 1.12.0).  Its documented algorithm
//                            is restated here: DynamicAutoDiffCostFunction (stride-4 forward Jets),
//                            HuberLoss + Corrector, Jacobi scaling, Levenberg-Marquardt trust region,
//                            exact normal-equation solve (done here by Schur elimination of the 6x6
//                            pose blocks + dense Cholesky instead of CHOLMOD).
//
// Parity pin: tests/test_oracle_golden.py checks this oracle against the reference's own
// logs test/radtan/correct_output.txt (12-row iteration table, termination, final cost,
// coefficients) and test/equidistant/correct_output.txt (final cost, coefficients).
//
// Plain C++17 + OpenMP, no third-party dependency.

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// ----------------------------------------------------------------------------------------
// Forward-mode dual numbers (what ceres::Jet<double,N> is).
// ----------------------------------------------------------------------------------------
template <int N>
struct Jet {
  double a;
  double v[N];
  Jet() : a(0.0) { for (int i = 0; i < N; i++) v[i] = 0.0; }
  Jet(double s) : a(s) { for (int i = 0; i < N; i++) v[i] = 0.0; }  // NOLINT (implicit on purpose)
};
template <int N> inline Jet<N> operator+(const Jet<N>& f, const Jet<N>& g) { Jet<N> h; h.a = f.a + g.a; for (int i = 0; i < N; i++) h.v[i] = f.v[i] + g.v[i]; return h; }
template <int N> inline Jet<N> operator-(const Jet<N>& f, const Jet<N>& g) { Jet<N> h; h.a = f.a - g.a; for (int i = 0; i < N; i++) h.v[i] = f.v[i] - g.v[i]; return h; }
template <int N> inline Jet<N> operator-(const Jet<N>& f) { Jet<N> h; h.a = -f.a; for (int i = 0; i < N; i++) h.v[i] = -f.v[i]; return h; }
template <int N> inline Jet<N> operator*(const Jet<N>& f, const Jet<N>& g) { Jet<N> h; h.a = f.a * g.a; for (int i = 0; i < N; i++) h.v[i] = f.a * g.v[i] + f.v[i] * g.a; return h; }
template <int N> inline Jet<N> operator/(const Jet<N>& f, const Jet<N>& g) {
  // (f/g)' = (f' - f/g g')/g
  Jet<N> h; const double gi = 1.0 / g.a; const double q = f.a * gi; h.a = q;
  for (int i = 0; i < N; i++) h.v[i] = (f.v[i] - q * g.v[i]) * gi;
  return h;
}
template <int N> inline Jet<N> operator+(const Jet<N>& f, double s) { Jet<N> h = f; h.a += s; return h; }
template <int N> inline Jet<N> operator+(double s, const Jet<N>& f) { return f + s; }
template <int N> inline Jet<N> operator-(const Jet<N>& f, double s) { Jet<N> h = f; h.a -= s; return h; }
template <int N> inline Jet<N> operator-(double s, const Jet<N>& f) { Jet<N> h = -f; h.a += s; return h; }
template <int N> inline Jet<N> operator*(const Jet<N>& f, double s) { Jet<N> h; h.a = f.a * s; for (int i = 0; i < N; i++) h.v[i] = f.v[i] * s; return h; }
template <int N> inline Jet<N> operator*(double s, const Jet<N>& f) { return f * s; }
template <int N> inline Jet<N> operator/(const Jet<N>& f, double s) { const double si = 1.0 / s; return f * si; }
template <int N> inline Jet<N>& operator+=(Jet<N>& f, const Jet<N>& g) { f = f + g; return f; }
template <int N> inline bool operator>(const Jet<N>& f, double s) { return f.a > s; }
template <int N> inline bool operator==(const Jet<N>& f, const Jet<N>& g) { return f.a == g.a; }
template <int N> inline bool operator!=(const Jet<N>& f, const Jet<N>& g) { return f.a != g.a; }

template <int N> inline Jet<N> sqrt(const Jet<N>& f) {
  Jet<N> h; h.a = std::sqrt(f.a); const double t = 1.0 / (2.0 * h.a);
  for (int i = 0; i < N; i++) h.v[i] = f.v[i] * t;
  return h;
}
template <int N> inline Jet<N> sin(const Jet<N>& f) { Jet<N> h; h.a = std::sin(f.a); const double c = std::cos(f.a); for (int i = 0; i < N; i++) h.v[i] = c * f.v[i]; return h; }
template <int N> inline Jet<N> cos(const Jet<N>& f) { Jet<N> h; h.a = std::cos(f.a); const double s = -std::sin(f.a); for (int i = 0; i < N; i++) h.v[i] = s * f.v[i]; return h; }
// atan2(g, f): d = (f dg - g df)/(f^2+g^2)
template <int N> inline Jet<N> atan2(const Jet<N>& g, const Jet<N>& f) {
  Jet<N> h; h.a = std::atan2(g.a, f.a); const double t = 1.0 / (f.a * f.a + g.a * g.a);
  for (int i = 0; i < N; i++) h.v[i] = t * (f.a * g.v[i] - g.a * f.v[i]);
  return h;
}
// pow(f, p) with a plain exponent
template <int N> inline Jet<N> pow(const Jet<N>& f, double p) {
  Jet<N> h; h.a = std::pow(f.a, p); const double t = p * std::pow(f.a, p - 1.0);
  for (int i = 0; i < N; i++) h.v[i] = t * f.v[i];
  return h;
}
using std::atan2; using std::cos; using std::pow; using std::sin; using std::sqrt;

// ----------------------------------------------------------------------------------------
// Geometry, generic over T in {double, Jet<N>}
// ----------------------------------------------------------------------------------------

// utils.h:29-39 : rotation vector -> matrix through (angle, axis); below 1e-8 the angle is the
// constant 0 about z, i.e. identity with NO dependence on the vector.  R is row-major 3x3.
template <typename T>
void rotation_matrix(const T* vec, T* R) {
  T angle = T(0.0);
  T ax[3] = {T(0.0), T(0.0), T(1.0)};
  const T nrm = sqrt(vec[0] * vec[0] + vec[1] * vec[1] + vec[2] * vec[2]);
  if (nrm > 1e-8) {
    ax[0] = vec[0] / nrm; ax[1] = vec[1] / nrm; ax[2] = vec[2] / nrm;
    angle = nrm;
  }
  // angle-axis -> matrix (the closed form Eigen's AngleAxis::toRotationMatrix evaluates)
  const T s = sin(angle), c = cos(angle);
  const T sa[3] = {s * ax[0], s * ax[1], s * ax[2]};
  const T ca[3] = {(T(1.0) - c) * ax[0], (T(1.0) - c) * ax[1], (T(1.0) - c) * ax[2]};
  T t;
  t = ca[0] * ax[1]; R[0 * 3 + 1] = t - sa[2]; R[1 * 3 + 0] = t + sa[2];
  t = ca[0] * ax[2]; R[0 * 3 + 2] = t + sa[1]; R[2 * 3 + 0] = t - sa[1];
  t = ca[1] * ax[2]; R[1 * 3 + 2] = t - sa[0]; R[2 * 3 + 1] = t + sa[0];
  R[0] = ca[0] * ax[0] + c; R[4] = ca[1] * ax[1] + c; R[8] = ca[2] * ax[2] + c;
}

// utils.h:62-82
template <typename T>
void project_point_equidistant(const double* Xw, const T* R, const T* t, const T* K4, const T* D, int nd, T* uv) {
  T Xc[3];
  for (int i = 0; i < 3; i++) Xc[i] = R[i * 3 + 0] * Xw[0] + R[i * 3 + 1] * Xw[1] + R[i * 3 + 2] * Xw[2] + t[i];
  const T n2 = Xc[0] * Xc[0] + Xc[1] * Xc[1];
  const T hyp = (n2 != T(0.0)) ? sqrt(n2) : T(0.0);
  const T theta = atan2(hyp, Xc[2]);
  const T psi = (Xc[0] == T(0.0) && Xc[1] == T(0.0)) ? T(0.0) : atan2(Xc[1], Xc[0]);
  T fac = T(1.0);
  for (int i = 0; i < nd; i++) fac += D[i] * pow(theta, (double)(2 * (i + 1)));
  const T r = theta * fac;
  uv[0] = K4[0] * (r * cos(psi)) + K4[2];
  uv[1] = K4[1] * (r * sin(psi)) + K4[3];
}

// utils.h:101-130 ; coefficients beyond the ones given are zero; OpenCV order k1,k2,p1,p2,k3,k4,k5,k6
template <typename T>
void project_point_radtan(const double* Xw, const T* R, const T* t, const T* K4, const T* Din, int nd, T* uv) {
  T Xc[3];
  for (int i = 0; i < 3; i++) Xc[i] = R[i * 3 + 0] * Xw[0] + R[i * 3 + 1] * Xw[1] + R[i * 3 + 2] * Xw[2] + t[i];
  const T x = Xc[0] / Xc[2], y = Xc[1] / Xc[2];
  const T x2 = x * x, y2 = y * y, xy = x * y;
  const T rsq = x2 + y2, r4 = rsq * rsq, r6 = rsq * r4;
  T D[8];
  for (int i = 0; i < 8; i++) D[i] = (i < nd) ? Din[i] : T(0.0);
  const T num = T(1.0) + D[0] * rsq + D[1] * r4 + D[4] * r6;
  const T den = T(1.0) + D[5] * rsq + D[6] * r4 + D[7] * r6;
  const T ratio = num / den;
  const T xpp = x * ratio + T(2.0) * D[2] * xy + D[3] * (rsq + T(2.0) * x2);
  const T ypp = y * ratio + D[2] * (rsq + T(2.0) * y2) + D[3] * T(2.0) * xy;
  uv[0] = K4[0] * xpp + K4[2];
  uv[1] = K4[1] * ypp + K4[3];
}

// ----------------------------------------------------------------------------------------
// Problem description (mirrors Calibrator's members, calibrator.h:42-48)
// ----------------------------------------------------------------------------------------
enum { MODEL_EQUIDISTANT = 0, MODEL_RADTAN = 1 };

struct Cam {
  int model = 0, n_intr = 4, n_dist = 4;
  double intr[4] = {0, 0, 0, 0};
  double dist[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  double T_cn_cnm1[16];
  bool fix_intr = false, fix_extr = false, active = true;
  int nparam() const { return n_intr + n_dist; }
};
struct FrameCam {
  std::vector<double> w;       // 3 per point
  std::vector<double> ip;      // 2 per point
  std::vector<uint8_t> mask;   // 1 = valid
};
struct Problem {
  std::vector<Cam> cams;
  bool fix_all_intr = false;
  std::vector<std::vector<FrameCam>> frames;   // [frame][cam]
  std::vector<double> pose_guess;              // 16 per frame, row-major 4x4
  std::vector<double> params;                  // calibrator.cpp:36-46 layout
  int ncam() const { return (int)cams.size(); }
  int nframes() const { return (int)frames.size(); }
  int num_intr_total() const { int n = 0; for (auto& c : cams) n += c.nparam(); return n; }
  int cam_params_total() const { return num_intr_total() + 6 * (ncam() - 1); }
  int intr_offset(int k) const { int n = 0; for (int i = 0; i < k; i++) n += cams[i].nparam(); return n; }
  int ext_offset(int k) const { return num_intr_total() + 6 * (k - 1); }   // k >= 1
  int pose_offset(int f) const { return cam_params_total() + 6 * f; }
  // calibrator.cpp:353-370: valid points of ACTIVE cameras
  size_t frame_num_valid(int f) const {
    size_t n = 0;
    for (int k = 0; k < ncam(); k++)
      if (cams[k].active) for (uint8_t m : frames[f][k].mask) n += (m != 0);
    return n;
  }
};

// calibrator.cpp:86-148.  blocks = [intr_0..intr_{n-1}, ext_1..ext_{n-1}, pose]  (calibrator.cpp:48-71)
template <typename T>
bool frame_residual(const Problem& P, int f, T const* const* blocks, T* residual) {
  const int ncam = P.ncam();
  const T* pose = blocks[2 * ncam - 1];
  T Rf[9];
  rotation_matrix(pose, Rf);
  const T tf[3] = {pose[3], pose[4], pose[5]};
  size_t cnt = 0;
  for (int k = 0; k < ncam; k++) {
    const Cam& cam = P.cams[k];
    if (!cam.active) continue;
    const T* ib = blocks[k];
    const T K4[4] = {ib[0], ib[1], ib[2], ib[3]};
    const T* D = ib + cam.n_intr;
    T Rc[9], tc[3];
    if (k == 0) {
      for (int i = 0; i < 9; i++) Rc[i] = T((i % 4 == 0) ? 1.0 : 0.0);
      tc[0] = tc[1] = tc[2] = T(0.0);
    } else {
      const T* eb = blocks[ncam + k - 1];
      rotation_matrix(eb, Rc);
      tc[0] = eb[3]; tc[1] = eb[4]; tc[2] = eb[5];
    }
    T R[9], t[3];
    for (int i = 0; i < 3; i++) {
      for (int j = 0; j < 3; j++)
        R[i * 3 + j] = Rc[i * 3 + 0] * Rf[0 * 3 + j] + Rc[i * 3 + 1] * Rf[1 * 3 + j] + Rc[i * 3 + 2] * Rf[2 * 3 + j];
      t[i] = Rc[i * 3 + 0] * tf[0] + Rc[i * 3 + 1] * tf[1] + Rc[i * 3 + 2] * tf[2] + tc[i];
    }
    const FrameCam& fc = P.frames[f][k];
    const size_t n = fc.mask.size();
    if (cam.model != MODEL_EQUIDISTANT && cam.model != MODEL_RADTAN) return false;
    for (size_t i = 0; i < n; i++) {
      if (!fc.mask[i]) continue;   // (the reference projects then skips; same values)
      T uv[2];
      if (cam.model == MODEL_EQUIDISTANT) project_point_equidistant(&fc.w[3 * i], R, t, K4, D, cam.n_dist, uv);
      else project_point_radtan(&fc.w[3 * i], R, t, K4, D, cam.n_dist, uv);
      residual[cnt++] = uv[0] - T(fc.ip[2 * i + 0]);
      residual[cnt++] = uv[1] - T(fc.ip[2 * i + 1]);
    }
  }
  return true;
}

// block table of a frame: pointer offsets into params and sizes
struct Blocks {
  std::vector<int> off, size;
};
Blocks frame_blocks(const Problem& P, int f) {
  Blocks b;
  const int ncam = P.ncam();
  for (int k = 0; k < ncam; k++) { b.off.push_back(P.intr_offset(k)); b.size.push_back(P.cams[k].nparam()); }
  for (int k = 1; k < ncam; k++) { b.off.push_back(P.ext_offset(k)); b.size.push_back(6); }
  b.off.push_back(P.pose_offset(f)); b.size.push_back(6);
  return b;
}

void eval_frame_double(const Problem& P, int f, const double* x, double* r) {
  Blocks b = frame_blocks(P, f);
  std::vector<const double*> ptr(b.off.size());
  for (size_t i = 0; i < ptr.size(); i++) ptr[i] = x + b.off[i];
  frame_residual<double>(P, f, ptr.data(), r);
}

// Forward autodiff the way ceres::DynamicAutoDiffCostFunction does it: repeated passes of the
// whole functor, STRIDE derivative slots per pass, over the parameters flagged in `want`.
// J is row-major nres x ncols_out; col_of[global param offset within the frame's block list] = output col or -1.
template <int STRIDE>
void eval_frame_jets(const Problem& P, int f, const double* x, const std::vector<int>& out_col /* per flat block-param */,
                     int ncols_out, size_t nres, double* r, double* J) {
  using JT = Jet<STRIDE>;
  Blocks b = frame_blocks(P, f);
  int nflat = 0;
  for (int s : b.size) nflat += s;
  std::vector<JT> flat(nflat);
  std::vector<const JT*> ptr(b.off.size());
  {
    int o = 0;
    for (size_t i = 0; i < b.off.size(); i++) {
      ptr[i] = flat.data() + o;
      for (int j = 0; j < b.size[i]; j++) flat[o + j] = JT(x[b.off[i] + j]);
      o += b.size[i];
    }
  }
  std::vector<int> active;
  for (int i = 0; i < nflat; i++) if (out_col[i] >= 0) active.push_back(i);
  std::vector<JT> res(nres);
  if (J) std::fill(J, J + nres * (size_t)ncols_out, 0.0);
  size_t pos = 0;
  bool first = true;
  while (pos < active.size() || first) {
    const size_t cnt = std::min<size_t>(STRIDE, active.size() - pos);
    for (size_t s = 0; s < cnt; s++) flat[active[pos + s]].v[s] = 1.0;
    frame_residual<JT>(P, f, ptr.data(), res.data());
    if (first && r) for (size_t i = 0; i < nres; i++) r[i] = res[i].a;
    if (J)
      for (size_t s = 0; s < cnt; s++) {
        const int c = out_col[active[pos + s]];
        for (size_t i = 0; i < nres; i++) J[i * ncols_out + c] = res[i].v[s];
      }
    for (size_t s = 0; s < cnt; s++) flat[active[pos + s]].v[s] = 0.0;
    pos += cnt;
    first = false;
  }
}

// ----------------------------------------------------------------------------------------
// Transforms (calibrator.cpp:157-170 via Eigen AngleAxisd(Matrix3d): matrix -> quaternion ->
// angle-axis, angle in [0, pi])
// ----------------------------------------------------------------------------------------
void transform_to_rvec_tvec(const double* T, double* out) {
  auto m = [&](int i, int j) { return T[i * 4 + j]; };
  double q[4];  // x y z w
  double t = m(0, 0) + m(1, 1) + m(2, 2);
  if (t > 0.0) {
    t = std::sqrt(t + 1.0);
    q[3] = 0.5 * t; t = 0.5 / t;
    q[0] = (m(2, 1) - m(1, 2)) * t; q[1] = (m(0, 2) - m(2, 0)) * t; q[2] = (m(1, 0) - m(0, 1)) * t;
  } else {
    int i = 0;
    if (m(1, 1) > m(0, 0)) i = 1;
    if (m(2, 2) > m(i, i)) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(m(i, i) - m(j, j) - m(k, k) + 1.0);
    q[i] = 0.5 * t; t = 0.5 / t;
    q[3] = (m(k, j) - m(j, k)) * t; q[j] = (m(j, i) + m(i, j)) * t; q[k] = (m(k, i) + m(i, k)) * t;
  }
  double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
  double angle = 0.0, ax[3] = {1.0, 0.0, 0.0};
  if (n != 0.0) {
    angle = 2.0 * std::atan2(n, std::fabs(q[3]));
    if (q[3] < 0.0) n = -n;
    ax[0] = q[0] / n; ax[1] = q[1] / n; ax[2] = q[2] / n;
  }
  out[0] = ax[0] * angle; out[1] = ax[1] * angle; out[2] = ax[2] * angle;
  out[3] = m(0, 3); out[4] = m(1, 3); out[5] = m(2, 3);
}

void rvec_tvec_to_transform(const double* in, double* T) {
  double R[9];
  rotation_matrix<double>(in, R);
  for (int i = 0; i < 3; i++) { for (int j = 0; j < 3; j++) T[i * 4 + j] = R[i * 3 + j]; T[i * 4 + 3] = in[3 + i]; }
  T[12] = T[13] = T[14] = 0.0; T[15] = 1.0;
}
void mat4_mul(const double* A, const double* B, double* C) {
  double tmp[16];
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { double s = 0; for (int k = 0; k < 4; k++) s += A[i * 4 + k] * B[k * 4 + j]; tmp[i * 4 + j] = s; }
  std::memcpy(C, tmp, sizeof(tmp));
}
// general 4x4 inverse by Gauss-Jordan with partial pivoting (Eigen's .inverse() on a Matrix4d is a
// cofactor formula; both are exact to rounding on rigid transforms)
void mat4_inv(const double* A, double* Ai) {
  double a[4][8];
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { a[i][j] = A[i * 4 + j]; a[i][4 + j] = (i == j); }
  for (int c = 0; c < 4; c++) {
    int p = c;
    for (int r = c + 1; r < 4; r++) if (std::fabs(a[r][c]) > std::fabs(a[p][c])) p = r;
    if (p != c) for (int j = 0; j < 8; j++) std::swap(a[c][j], a[p][j]);
    const double d = 1.0 / a[c][c];
    for (int j = 0; j < 8; j++) a[c][j] *= d;
    for (int r = 0; r < 4; r++) if (r != c) { const double fct = a[r][c]; for (int j = 0; j < 8; j++) a[r][j] -= fct * a[c][j]; }
  }
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) Ai[i * 4 + j] = a[i][4 + j];
}

// calibrator.cpp:251-283
void initialize_variables(Problem& P) {
  P.params.clear();
  for (auto& c : P.cams) {
    for (int i = 0; i < c.n_intr; i++) P.params.push_back(c.intr[i]);
    for (int i = 0; i < c.n_dist; i++) P.params.push_back(c.dist[i]);
  }
  double Tprev[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  for (int k = 1; k < P.ncam(); k++) {
    double Tn0[16];
    mat4_mul(P.cams[k].T_cn_cnm1, Tprev, Tn0);
    double rt[6];
    transform_to_rvec_tvec(Tn0, rt);
    P.params.insert(P.params.end(), rt, rt + 6);
    std::memcpy(Tprev, Tn0, sizeof(Tn0));
  }
  for (int f = 0; f < P.nframes(); f++) {
    double rt[6];
    transform_to_rvec_tvec(&P.pose_guess[16 * f], rt);
    P.params.insert(P.params.end(), rt, rt + 6);
  }
}

// ----------------------------------------------------------------------------------------
// Reduced program: column maps (Ceres removes constant blocks, calibrator.cpp:510-521)
// ----------------------------------------------------------------------------------------
struct Reduced {
  int C = 0;                         // reduced camera columns
  std::vector<int> cam_col;          // [cam_params_total] -> reduced col or -1
  std::vector<int> cam_param_of_col; // [C] -> param index
  std::vector<int> frames;           // frames with >=1 valid active point
  std::vector<int> frame_slot;       // [nframes] -> index in frames or -1
  std::vector<size_t> nres;          // per active frame: 2*nvalid
  std::vector<size_t> row0;          // residual offset
  size_t total_res = 0;
  int nred() const { return C + 6 * (int)frames.size(); }
};
Reduced build_reduced(const Problem& P) {
  Reduced R;
  const int ct = P.cam_params_total();
  R.cam_col.assign(ct, -1);
  for (int k = 0; k < P.ncam(); k++) {
    const bool fixed = P.fix_all_intr || P.cams[k].fix_intr;
    if (!fixed) for (int j = 0; j < P.cams[k].nparam(); j++) { R.cam_col[P.intr_offset(k) + j] = R.C++; }
  }
  for (int k = 1; k < P.ncam(); k++)
    if (!P.cams[k].fix_extr) for (int j = 0; j < 6; j++) R.cam_col[P.ext_offset(k) + j] = R.C++;
  R.cam_param_of_col.resize(R.C);
  for (int i = 0; i < ct; i++) if (R.cam_col[i] >= 0) R.cam_param_of_col[R.cam_col[i]] = i;
  R.frame_slot.assign(P.nframes(), -1);
  for (int f = 0; f < P.nframes(); f++) {
    const size_t nv = P.frame_num_valid(f);
    if (nv == 0) continue;
    R.frame_slot[f] = (int)R.frames.size();
    R.frames.push_back(f);
    R.row0.push_back(R.total_res);
    R.nres.push_back(2 * nv);
    R.total_res += 2 * nv;
  }
  return R;
}

// Huber(a=1): rho, rho'
inline void huber(double s, double* rho) {
  if (s > 1.0) { const double r = std::sqrt(s); rho[0] = 2.0 * r - 1.0; rho[1] = std::max(std::numeric_limits<double>::min(), 1.0 / r); rho[2] = -rho[1] / (2.0 * s); }
  else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
}

struct Options {
  int max_num_iterations = 200;
  double function_tolerance = 1e-12, parameter_tolerance = 1e-12, gradient_tolerance = 1e-10;
  double initial_trust_region_radius = 1e4, max_trust_region_radius = 1e16, min_trust_region_radius = 1e-32;
  double min_relative_decrease = 1e-3, min_lm_diagonal = 1e-6, max_lm_diagonal = 1e32;
  int max_num_consecutive_invalid_steps = 5;
  int jacobi_scaling = 1;
  int num_threads = 4;
  int use_loss = 1;            // Huber(1.0) per frame block
  int autodiff_all_params = 1; // 1: Jets over every non-constant block of the frame (what Ceres does)
  int verbose = 0;
};
enum { TERM_CONVERGENCE = 0, TERM_NO_CONVERGENCE = 1, TERM_FAILURE = 2 };
enum { WHY_GRADIENT = 0, WHY_PARAMETER = 1, WHY_FUNCTION = 2, WHY_MAX_ITER = 3, WHY_INVALID = 4, WHY_MIN_RADIUS = 5, WHY_LINEAR = 6, WHY_NODATA = 7 };

#define ORC_MAX_TRACE 256
struct Summary {
  double initial_cost, final_cost;
  int num_iterations;        // logged rows, row 0 included ("Minimizer iterations")
  int num_successful, num_unsuccessful;
  int termination, why;
  double why_value;          // the quantity compared with its tolerance
  int num_params, num_reduced_params;
  long long num_residuals;
  double t_jacobian, t_linear, t_residual, t_total;
  int n_jacobian_evals, n_residual_evals, n_linear_solves;
  int trace_len;
  double tr_cost[ORC_MAX_TRACE], tr_cost_change[ORC_MAX_TRACE], tr_gradient_max[ORC_MAX_TRACE],
      tr_step_norm[ORC_MAX_TRACE], tr_ratio[ORC_MAX_TRACE], tr_radius[ORC_MAX_TRACE];
};

double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// dense Cholesky in place (lower), returns false if not positive definite
bool chol_inplace(double* A, int n) {
  for (int j = 0; j < n; j++) {
    double d = A[j * n + j];
    for (int k = 0; k < j; k++) d -= A[j * n + k] * A[j * n + k];
    if (!(d > 0.0)) return false;
    d = std::sqrt(d);
    A[j * n + j] = d;
    for (int i = j + 1; i < n; i++) {
      double s = A[i * n + j];
      for (int k = 0; k < j; k++) s -= A[i * n + k] * A[j * n + k];
      A[i * n + j] = s / d;
    }
  }
  return true;
}
void chol_solve(const double* L, int n, double* b) {
  for (int i = 0; i < n; i++) { double s = b[i]; for (int k = 0; k < i; k++) s -= L[i * n + k] * b[k]; b[i] = s / L[i * n + i]; }
  for (int i = n - 1; i >= 0; i--) { double s = b[i]; for (int k = i + 1; k < n; k++) s -= L[k * n + i] * b[k]; b[i] = s / L[i * n + i]; }
}

// Evaluation state at a point
struct Eval {
  double cost = 0;
  std::vector<double> r;   // corrected residuals, total_res
  std::vector<double> J;   // corrected jacobian, per active frame nres x (C+6), concatenated
  std::vector<size_t> j0;  // offset per active frame
  std::vector<double> g;   // gradient over reduced params
};

struct Solver {
  const Problem& P;
  Reduced R;
  Options o;
  explicit Solver(const Problem& p, const Options& opt) : P(p), R(build_reduced(p)), o(opt) {}

  // reduced vector <-> params
  void gather(const std::vector<double>& params, std::vector<double>& x) const {
    x.resize(R.nred());
    for (int c = 0; c < R.C; c++) x[c] = params[R.cam_param_of_col[c]];
    for (size_t s = 0; s < R.frames.size(); s++)
      for (int j = 0; j < 6; j++) x[R.C + 6 * s + j] = params[P.pose_offset(R.frames[s]) + j];
  }
  void scatter(const std::vector<double>& x, std::vector<double>& params) const {
    for (int c = 0; c < R.C; c++) params[R.cam_param_of_col[c]] = x[c];
    for (size_t s = 0; s < R.frames.size(); s++)
      for (int j = 0; j < 6; j++) params[P.pose_offset(R.frames[s]) + j] = x[R.C + 6 * s + j];
  }

  // cost only (candidate evaluation): 1/2 sum rho(s_f) on raw residuals
  double cost_only(const std::vector<double>& params) const {
    const int nf = (int)R.frames.size();
    std::vector<double> cf(nf);
#pragma omp parallel for schedule(dynamic, 4) num_threads(o.num_threads)
    for (int s = 0; s < nf; s++) {
      std::vector<double> r(R.nres[s]);
      eval_frame_double(P, R.frames[s], params.data(), r.data());
      double sq = 0;
      for (double v : r) sq += v * v;
      double rho[3];
      if (o.use_loss) huber(sq, rho); else rho[0] = sq;
      cf[s] = 0.5 * rho[0];
    }
    double c = 0;
    for (double v : cf) c += v;
    return c;
  }

  void evaluate(const std::vector<double>& params, Eval& E) const {
    const int nf = (int)R.frames.size();
    const int W = R.C + 6;
    E.r.assign(R.total_res, 0.0);
    E.j0.resize(nf);
    size_t tot = 0;
    for (int s = 0; s < nf; s++) { E.j0[s] = tot; tot += R.nres[s] * (size_t)W; }
    E.J.resize(tot);
    std::vector<double> cf(nf);
    const int ncam = P.ncam();
#pragma omp parallel for schedule(dynamic, 4) num_threads(o.num_threads)
    for (int s = 0; s < nf; s++) {
      const int f = R.frames[s];
      // flat block-param -> output column
      std::vector<int> out_col;
      for (int k = 0; k < ncam; k++)
        for (int j = 0; j < P.cams[k].nparam(); j++) out_col.push_back(R.cam_col[P.intr_offset(k) + j]);
      for (int k = 1; k < ncam; k++)
        for (int j = 0; j < 6; j++) out_col.push_back(R.cam_col[P.ext_offset(k) + j]);
      for (int j = 0; j < 6; j++) out_col.push_back(R.C + j);
      if (!o.autodiff_all_params) {
        // skip blocks of cameras that contribute nothing to this frame (cheaper, same numbers)
        int idx = 0;
        for (int k = 0; k < ncam; k++) {
          bool seen = P.cams[k].active && std::any_of(P.frames[f][k].mask.begin(), P.frames[f][k].mask.end(), [](uint8_t m) { return m != 0; });
          for (int j = 0; j < P.cams[k].nparam(); j++, idx++) if (!seen) out_col[idx] = -1;
        }
        for (int k = 1; k < ncam; k++) {
          bool seen = P.cams[k].active && std::any_of(P.frames[f][k].mask.begin(), P.frames[f][k].mask.end(), [](uint8_t m) { return m != 0; });
          for (int j = 0; j < 6; j++, idx++) if (!seen) out_col[idx] = -1;
        }
      }
      double* r = &E.r[R.row0[s]];
      double* J = &E.J[E.j0[s]];
      eval_frame_jets<4>(P, f, params.data(), out_col, W, R.nres[s], r, J);
      double sq = 0;
      for (size_t i = 0; i < R.nres[s]; i++) sq += r[i] * r[i];
      double rho[3] = {sq, 1.0, 0.0};
      if (o.use_loss) huber(sq, rho);
      cf[s] = 0.5 * rho[0];
      // Corrector: rho'' <= 0 for Huber -> residuals and Jacobian rows both scaled by sqrt(rho')
      const double sc = std::sqrt(rho[1]);
      if (sc != 1.0) {
        for (size_t i = 0; i < R.nres[s]; i++) r[i] *= sc;
        for (size_t i = 0; i < R.nres[s] * (size_t)W; i++) J[i] *= sc;
      }
    }
    E.cost = 0;
    for (double v : cf) E.cost += v;
    // gradient = J^T r
    E.g.assign(R.nred(), 0.0);
    for (int s = 0; s < nf; s++) {
      const double* r = &E.r[R.row0[s]];
      const double* J = &E.J[E.j0[s]];
      for (size_t i = 0; i < R.nres[s]; i++) {
        const double ri = r[i];
        const double* Ji = J + i * W;
        for (int c = 0; c < R.C; c++) E.g[c] += Ji[c] * ri;
        for (int j = 0; j < 6; j++) E.g[R.C + 6 * s + j] += Ji[R.C + j] * ri;
      }
    }
  }

  void squared_column_norms(const Eval& E, std::vector<double>& n2) const {
    const int nf = (int)R.frames.size(), W = R.C + 6;
    n2.assign(R.nred(), 0.0);
    for (int s = 0; s < nf; s++) {
      const double* J = &E.J[E.j0[s]];
      for (size_t i = 0; i < R.nres[s]; i++) {
        const double* Ji = J + i * W;
        for (int c = 0; c < R.C; c++) n2[c] += Ji[c] * Ji[c];
        for (int j = 0; j < 6; j++) n2[R.C + 6 * s + j] += Ji[R.C + j] * Ji[R.C + j];
      }
    }
  }
  void scale_columns(Eval& E, const std::vector<double>& scale) const {
    const int nf = (int)R.frames.size(), W = R.C + 6;
#pragma omp parallel for schedule(static) num_threads(o.num_threads)
    for (int s = 0; s < nf; s++) {
      double* J = &E.J[E.j0[s]];
      for (size_t i = 0; i < R.nres[s]; i++) {
        double* Ji = J + i * W;
        for (int c = 0; c < R.C; c++) Ji[c] *= scale[c];
        for (int j = 0; j < 6; j++) Ji[R.C + j] *= scale[R.C + 6 * s + j];
      }
    }
  }

  // normal-equation blocks of the (scaled, corrected) Jacobian held in E
  struct NormalBlocks {
    std::vector<double> Hcc, gc;         // C x C, C
    std::vector<double> Hff, gf, Hfc;    // per frame 36, 6, 6 x C
  };
  void normal_blocks(const Eval& E, NormalBlocks& N) const {
    const int nf = (int)R.frames.size(), W = R.C + 6, C = R.C;
    N.Hff.assign((size_t)nf * 36, 0.0); N.gf.assign((size_t)nf * 6, 0.0); N.Hfc.assign((size_t)nf * 6 * C, 0.0);
    std::vector<double> Hcc_f((size_t)nf * C * C, 0.0), gc_f((size_t)nf * C, 0.0);
#pragma omp parallel for schedule(dynamic, 4) num_threads(o.num_threads)
    for (int s = 0; s < nf; s++) {
      const double* r = &E.r[R.row0[s]];
      const double* J = &E.J[E.j0[s]];
      double* Hff = &N.Hff[(size_t)s * 36]; double* gf = &N.gf[(size_t)s * 6]; double* Hfc = &N.Hfc[(size_t)s * 6 * C];
      double* Hcc = &Hcc_f[(size_t)s * C * C]; double* gc = &gc_f[(size_t)s * C];
      for (size_t i = 0; i < R.nres[s]; i++) {
        const double* Ji = J + i * W;
        const double* Jp = Ji + C;
        for (int a = 0; a < 6; a++) {
          for (int b = 0; b < 6; b++) Hff[a * 6 + b] += Jp[a] * Jp[b];
          gf[a] += Jp[a] * r[i];
          for (int c = 0; c < C; c++) Hfc[a * C + c] += Jp[a] * Ji[c];
        }
        for (int a = 0; a < C; a++) {
          if (Ji[a] == 0.0) continue;
          for (int b = 0; b < C; b++) Hcc[a * C + b] += Ji[a] * Ji[b];
          gc[a] += Ji[a] * r[i];
        }
      }
    }
    N.Hcc.assign((size_t)C * C, 0.0); N.gc.assign(C, 0.0);
    for (int s = 0; s < nf; s++) {
      for (int i = 0; i < C * C; i++) N.Hcc[i] += Hcc_f[(size_t)s * C * C + i];
      for (int i = 0; i < C; i++) N.gc[i] += gc_f[(size_t)s * C + i];
    }
  }

  // Solve (H + D^2) y = g by eliminating the pose blocks. D is the LM diagonal (sqrt form).
  bool schur_solve(const NormalBlocks& N, const std::vector<double>& D, std::vector<double>& y) const {
    const int nf = (int)R.frames.size(), C = R.C;
    std::vector<double> S(N.Hcc), b(N.gc);
    for (int c = 0; c < C; c++) S[c * C + c] += D[c] * D[c];
    std::vector<double> Lf((size_t)nf * 36), Wf((size_t)nf * 6 * C), zf((size_t)nf * 6);
    bool ok = true;
#pragma omp parallel for schedule(static) num_threads(o.num_threads)
    for (int s = 0; s < nf; s++) {
      double* L = &Lf[(size_t)s * 36];
      std::memcpy(L, &N.Hff[(size_t)s * 36], 36 * sizeof(double));
      for (int j = 0; j < 6; j++) { const double d = D[C + 6 * s + j]; L[j * 6 + j] += d * d; }
      if (!chol_inplace(L, 6)) { ok = false; continue; }
      double* Wm = &Wf[(size_t)s * 6 * C];
      const double* Hfc = &N.Hfc[(size_t)s * 6 * C];
      for (int c = 0; c < C; c++)
        for (int i = 0; i < 6; i++) {
          double v = Hfc[i * C + c];
          for (int k = 0; k < i; k++) v -= L[i * 6 + k] * Wm[k * C + c];
          Wm[i * C + c] = v / L[i * 6 + i];
        }
      double* z = &zf[(size_t)s * 6];
      for (int i = 0; i < 6; i++) {
        double v = N.gf[(size_t)s * 6 + i];
        for (int k = 0; k < i; k++) v -= L[i * 6 + k] * z[k];
        z[i] = v / L[i * 6 + i];
      }
    }
    if (!ok) return false;
    if (C > 0) {
      for (int s = 0; s < nf; s++) {
        const double* Wm = &Wf[(size_t)s * 6 * C];
        const double* z = &zf[(size_t)s * 6];
        for (int i = 0; i < 6; i++) {
          const double* wi = Wm + i * C;
          for (int a = 0; a < C; a++) {
            const double wa = wi[a];
            if (wa == 0.0) continue;
            double* Sa = &S[a * C];
            for (int c = 0; c < C; c++) Sa[c] -= wa * wi[c];
            b[a] -= wa * z[i];
          }
        }
      }
      if (!chol_inplace(S.data(), C)) return false;
      chol_solve(S.data(), C, b.data());
    }
    y.assign(R.nred(), 0.0);
    for (int c = 0; c < C; c++) y[c] = b[c];
    for (int s = 0; s < nf; s++) {
      const double* L = &Lf[(size_t)s * 36];
      const double* Wm = &Wf[(size_t)s * 6 * C];
      double q[6];
      for (int i = 0; i < 6; i++) {
        double v = zf[(size_t)s * 6 + i];
        for (int c = 0; c < C; c++) v -= Wm[i * C + c] * b[c];
        q[i] = v;
      }
      for (int i = 5; i >= 0; i--) {
        double v = q[i];
        for (int k = i + 1; k < 6; k++) v -= L[k * 6 + i] * q[k];
        q[i] = v / L[i * 6 + i];
      }
      for (int i = 0; i < 6; i++) y[C + 6 * s + i] = q[i];
    }
    return true;
  }

  // m = J_s * step  ->  model_cost_change = -m.(r + m/2)
  double model_cost_change(const Eval& E, const std::vector<double>& step) const {
    const int nf = (int)R.frames.size(), W = R.C + 6;
    std::vector<double> part(nf, 0.0);
#pragma omp parallel for schedule(static) num_threads(o.num_threads)
    for (int s = 0; s < nf; s++) {
      const double* r = &E.r[R.row0[s]];
      const double* J = &E.J[E.j0[s]];
      double acc = 0;
      for (size_t i = 0; i < R.nres[s]; i++) {
        const double* Ji = J + i * W;
        double m = 0;
        for (int c = 0; c < R.C; c++) m += Ji[c] * step[c];
        for (int j = 0; j < 6; j++) m += Ji[R.C + j] * step[R.C + 6 * s + j];
        acc += m * (r[i] + m / 2.0);
      }
      part[s] = acc;
    }
    double tot = 0;
    for (double v : part) tot += v;
    return -tot;
  }

  static double max_norm_gradient(const std::vector<double>& x, const std::vector<double>& g) {
    // Ceres: || x - Plus(x, -g) ||_inf
    double m = 0;
    for (size_t i = 0; i < x.size(); i++) { const double p = x[i] + (-g[i]); m = std::max(m, std::fabs(x[i] - p)); }
    return m;
  }
  static double norm2(const std::vector<double>& v) { double s = 0; for (double a : v) s += a * a; return std::sqrt(s); }

  void solve(std::vector<double>& params, Summary& S) {
    std::memset(&S, 0, sizeof(S));
    const double t_start = now_s();
    S.num_params = (int)params.size();
    S.num_reduced_params = R.nred();
    S.num_residuals = (long long)R.total_res;
    auto log_row = [&](double cost, double cc, double gm, double sn, double ratio, double radius) {
      if (S.trace_len < ORC_MAX_TRACE) {
        const int i = S.trace_len++;
        S.tr_cost[i] = cost; S.tr_cost_change[i] = cc; S.tr_gradient_max[i] = gm; S.tr_step_norm[i] = sn; S.tr_ratio[i] = ratio; S.tr_radius[i] = radius;
      }
      S.num_iterations++;
      if (o.verbose) std::printf("%4d  %.6e  %9.2e  %9.2e  %9.2e  %9.2e  %9.2e\n", S.num_iterations - 1, cost, cc, gm, sn, ratio, radius);
    };
    if (R.frames.empty()) { S.termination = TERM_FAILURE; S.why = WHY_NODATA; return; }

    std::vector<double> x, x_plus, cand_params(params);
    gather(params, x);
    double x_norm = norm2(x);
    Eval E;
    double t0 = now_s();
    evaluate(params, E);
    S.t_jacobian += now_s() - t0; S.n_jacobian_evals++;
    double cost = E.cost;
    S.initial_cost = cost;
    double gmax = max_norm_gradient(x, E.g);
    std::vector<double> scale(R.nred(), 1.0);
    if (o.jacobi_scaling) {
      std::vector<double> n2;
      squared_column_norms(E, n2);
      for (int i = 0; i < R.nred(); i++) scale[i] = 1.0 / (1.0 + std::sqrt(n2[i]));
      scale_columns(E, scale);
    }
    double radius = o.initial_trust_region_radius, decrease_factor = 2.0;
    bool reuse_diagonal = false;
    std::vector<double> diagonal;
    NormalBlocks N;
    bool blocks_valid = false;
    log_row(cost, 0, gmax, 0, 0, radius);
    S.num_successful = 1;
    auto finish = [&](int term, int why, double val) {
      S.termination = term; S.why = why; S.why_value = val; S.final_cost = cost;
      scatter(x, params);
      S.t_total = now_s() - t_start;
    };
    if (gmax <= o.gradient_tolerance) { finish(TERM_CONVERGENCE, WHY_GRADIENT, gmax); return; }
    int invalid_run = 0;
    int iter = 0;
    while (true) {
      if (iter >= o.max_num_iterations) { finish(TERM_NO_CONVERGENCE, WHY_MAX_ITER, iter); return; }
      iter++;
      t0 = now_s();
      if (!blocks_valid) { normal_blocks(E, N); blocks_valid = true; }
      if (!reuse_diagonal) {
        // squared column norms of the scaled Jacobian = diag of the blocks
        diagonal.resize(R.nred());
        for (int c = 0; c < R.C; c++) diagonal[c] = N.Hcc[c * R.C + c];
        for (size_t s = 0; s < R.frames.size(); s++) for (int j = 0; j < 6; j++) diagonal[R.C + 6 * s + j] = N.Hff[s * 36 + j * 6 + j];
        for (double& d : diagonal) d = std::min(std::max(d, o.min_lm_diagonal), o.max_lm_diagonal);
      }
      std::vector<double> D(R.nred());
      for (int i = 0; i < R.nred(); i++) D[i] = std::sqrt(diagonal[i] / radius);
      // right-hand side J_s^T r
      std::vector<double> y;
      {
        // gradient in scaled space
        NormalBlocks& NN = N;
        // N.gc / N.gf already are J_s^T r (E.J is scaled)
        (void)NN;
      }
      bool lin_ok = schur_solve(N, D, y);
      reuse_diagonal = true;
      S.t_linear += now_s() - t0; S.n_linear_solves++;
      bool step_valid = lin_ok;
      std::vector<double> step(R.nred());
      double mcc = 0;
      if (lin_ok) {
        for (int i = 0; i < R.nred(); i++) { step[i] = -y[i]; if (!std::isfinite(step[i])) step_valid = false; }
      }
      if (step_valid) {
        mcc = model_cost_change(E, step);
        if (!(mcc > 0.0)) step_valid = false;
      }
      if (!step_valid) {
        invalid_run++;
        if (invalid_run >= o.max_num_consecutive_invalid_steps) { finish(TERM_FAILURE, WHY_INVALID, invalid_run); return; }
        radius /= decrease_factor; decrease_factor *= 2.0; reuse_diagonal = true;
        S.num_unsuccessful++;
        log_row(cost, 0, gmax, 0, 0, radius);
        if (radius < o.min_trust_region_radius) { finish(TERM_CONVERGENCE, WHY_MIN_RADIUS, radius); return; }
        continue;
      }
      invalid_run = 0;
      x_plus.resize(R.nred());
      for (int i = 0; i < R.nred(); i++) x_plus[i] = x[i] + step[i] * scale[i];
      double sn2 = 0;
      for (int i = 0; i < R.nred(); i++) { const double d = x[i] - x_plus[i]; sn2 += d * d; }
      const double step_norm = std::sqrt(sn2);
      scatter(x_plus, cand_params);
      t0 = now_s();
      const double new_cost = cost_only(cand_params);
      S.t_residual += now_s() - t0; S.n_residual_evals++;
      if (step_norm <= o.parameter_tolerance * (x_norm + o.parameter_tolerance)) {
        finish(TERM_CONVERGENCE, WHY_PARAMETER, step_norm / (x_norm + o.parameter_tolerance)); return;
      }
      const double cost_change = cost - new_cost;
      if (std::fabs(cost_change) <= o.function_tolerance * cost) {
        finish(TERM_CONVERGENCE, WHY_FUNCTION, std::fabs(cost_change) / cost); return;
      }
      const double ratio = cost_change / mcc;
      if (ratio > o.min_relative_decrease) {
        x = x_plus; x_norm = norm2(x);
        scatter(x, params);
        radius = radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * ratio - 1.0, 3));
        radius = std::min(o.max_trust_region_radius, radius);
        decrease_factor = 2.0; reuse_diagonal = false;
        t0 = now_s();
        evaluate(params, E);
        S.t_jacobian += now_s() - t0; S.n_jacobian_evals++;
        cost = E.cost;
        gmax = max_norm_gradient(x, E.g);
        if (o.jacobi_scaling) scale_columns(E, scale);
        blocks_valid = false;
        S.num_successful++;
        log_row(cost, cost_change, gmax, step_norm, ratio, radius);
        if (gmax <= o.gradient_tolerance) { finish(TERM_CONVERGENCE, WHY_GRADIENT, gmax); return; }
      } else {
        radius /= decrease_factor; decrease_factor *= 2.0; reuse_diagonal = true;
        S.num_unsuccessful++;
        log_row(cost, cost_change, 0.0, step_norm, ratio, radius);
      }
      if (radius < o.min_trust_region_radius) { finish(TERM_CONVERGENCE, WHY_MIN_RADIUS, radius); return; }
    }
  }
};

}  // namespace

// ========================================================================================
// C interface (ctypes-friendly)
// ========================================================================================
extern "C" {

struct orc_options {
  int max_num_iterations;
  double function_tolerance, parameter_tolerance, gradient_tolerance;
  double initial_trust_region_radius, max_trust_region_radius, min_trust_region_radius;
  double min_relative_decrease, min_lm_diagonal, max_lm_diagonal;
  int max_num_consecutive_invalid_steps;
  int jacobi_scaling;
  int num_threads;
  int use_loss;
  int autodiff_all_params;
  int verbose;
};

void orc_default_options(orc_options* o) {
  Options d;
  o->max_num_iterations = d.max_num_iterations;
  o->function_tolerance = d.function_tolerance; o->parameter_tolerance = d.parameter_tolerance; o->gradient_tolerance = d.gradient_tolerance;
  o->initial_trust_region_radius = d.initial_trust_region_radius; o->max_trust_region_radius = d.max_trust_region_radius; o->min_trust_region_radius = d.min_trust_region_radius;
  o->min_relative_decrease = d.min_relative_decrease; o->min_lm_diagonal = d.min_lm_diagonal; o->max_lm_diagonal = d.max_lm_diagonal;
  o->max_num_consecutive_invalid_steps = d.max_num_consecutive_invalid_steps;
  o->jacobi_scaling = d.jacobi_scaling; o->num_threads = d.num_threads; o->use_loss = d.use_loss;
  o->autodiff_all_params = d.autodiff_all_params; o->verbose = d.verbose;
}
static Options to_options(const orc_options* o) {
  Options d;
  if (!o) return d;
  d.max_num_iterations = o->max_num_iterations;
  d.function_tolerance = o->function_tolerance; d.parameter_tolerance = o->parameter_tolerance; d.gradient_tolerance = o->gradient_tolerance;
  d.initial_trust_region_radius = o->initial_trust_region_radius; d.max_trust_region_radius = o->max_trust_region_radius; d.min_trust_region_radius = o->min_trust_region_radius;
  d.min_relative_decrease = o->min_relative_decrease; d.min_lm_diagonal = o->min_lm_diagonal; d.max_lm_diagonal = o->max_lm_diagonal;
  d.max_num_consecutive_invalid_steps = o->max_num_consecutive_invalid_steps;
  d.jacobi_scaling = o->jacobi_scaling; d.num_threads = std::max(1, o->num_threads); d.use_loss = o->use_loss;
  d.autodiff_all_params = o->autodiff_all_params; d.verbose = o->verbose;
  return d;
}

typedef Summary orc_summary;
int orc_summary_size() { return (int)sizeof(Summary); }
int orc_max_trace() { return ORC_MAX_TRACE; }

void* orc_create(int ncam, const int* model, const int* n_dist, const double* intr /*4 per cam*/, const double* dist /*8 per cam*/,
                 const double* T_cn_cnm1 /*16 per cam*/, const int* fix_intr, const int* fix_extr, const int* active, int fix_all_intr) {
  Problem* P = new Problem();
  P->cams.resize(ncam);
  for (int k = 0; k < ncam; k++) {
    Cam& c = P->cams[k];
    c.model = model[k]; c.n_dist = n_dist[k];
    for (int i = 0; i < 4; i++) c.intr[i] = intr[4 * k + i];
    for (int i = 0; i < 8; i++) c.dist[i] = dist[8 * k + i];
    std::memcpy(c.T_cn_cnm1, T_cn_cnm1 + 16 * k, 16 * sizeof(double));
    c.fix_intr = fix_intr[k] != 0; c.fix_extr = fix_extr[k] != 0; c.active = active[k] != 0;
  }
  P->fix_all_intr = fix_all_intr != 0;
  return P;
}
void orc_destroy(void* h) { delete (Problem*)h; }

// calibrator.cpp:229-249. world: 3 doubles/pt, image: 2 doubles/pt, cameras concatenated in order.
void orc_add_frame(void* h, const int* npts, const double* world, const double* image, const double* T_c0_w) {
  Problem& P = *(Problem*)h;
  std::vector<FrameCam> fr(P.ncam());
  size_t o = 0;
  for (int k = 0; k < P.ncam(); k++) {
    const size_t n = npts[k];
    fr[k].w.assign(world + 3 * o, world + 3 * (o + n));
    fr[k].ip.assign(image + 2 * o, image + 2 * (o + n));
    fr[k].mask.assign(n, 1);
    o += n;
  }
  P.frames.push_back(std::move(fr));
  P.pose_guess.insert(P.pose_guess.end(), T_c0_w, T_c0_w + 16);
}
void orc_set_flags(void* h, const int* fix_intr, const int* fix_extr, const int* active, int fix_all_intr) {
  Problem& P = *(Problem*)h;
  for (int k = 0; k < P.ncam(); k++) { P.cams[k].fix_intr = fix_intr[k] != 0; P.cams[k].fix_extr = fix_extr[k] != 0; P.cams[k].active = active[k] != 0; }
  P.fix_all_intr = fix_all_intr != 0;
}
void orc_init_params(void* h) { initialize_variables(*(Problem*)h); }
int orc_num_params(void* h) { Problem& P = *(Problem*)h; return P.cam_params_total() + 6 * P.nframes(); }
int orc_num_cam_params(void* h) { return ((Problem*)h)->cam_params_total(); }
void orc_set_params(void* h, const double* p) { Problem& P = *(Problem*)h; P.params.assign(p, p + orc_num_params(h)); }
void orc_get_params(void* h, double* p) { Problem& P = *(Problem*)h; std::memcpy(p, P.params.data(), P.params.size() * sizeof(double)); }
long long orc_num_residuals(void* h) {
  Problem& P = *(Problem*)h; long long n = 0;
  for (int f = 0; f < P.nframes(); f++) n += 2 * (long long)P.frame_num_valid(f);
  return n;
}
int orc_num_reduced_cam_cols(void* h) { return build_reduced(*(Problem*)h).C; }
void orc_reduced_cam_cols(void* h, int* col_of_param) {
  Problem& P = *(Problem*)h; Reduced R = build_reduced(P);
  for (int i = 0; i < P.cam_params_total(); i++) col_of_param[i] = R.cam_col[i];
}
void orc_get_mask(void* h, uint8_t* out) {
  Problem& P = *(Problem*)h; size_t o = 0;
  for (int f = 0; f < P.nframes(); f++) for (int k = 0; k < P.ncam(); k++) for (uint8_t m : P.frames[f][k].mask) out[o++] = m;
}
void orc_set_mask(void* h, const uint8_t* in) {
  Problem& P = *(Problem*)h; size_t o = 0;
  for (int f = 0; f < P.nframes(); f++) for (int k = 0; k < P.ncam(); k++) for (uint8_t& m : P.frames[f][k].mask) m = in[o++];
}

// calibrator.cpp:316-351: raw residuals (no loss), frame -> active cam -> valid point order, (x,y) pairs
void orc_eval_residuals(void* h, double* out) {
  Problem& P = *(Problem*)h; size_t o = 0;
  for (int f = 0; f < P.nframes(); f++) {
    const size_t nv = P.frame_num_valid(f);
    if (!nv) continue;
    eval_frame_double(P, f, P.params.data(), out + o);
    o += 2 * nv;
  }
}
// Huber cost 1/2 sum_f rho(s_f) at the current parameters
double orc_cost(void* h, int use_loss, double* per_frame /*nframes or NULL*/) {
  Problem& P = *(Problem*)h; double tot = 0;
  for (int f = 0; f < P.nframes(); f++) {
    const size_t nv = P.frame_num_valid(f);
    double c = 0;
    if (nv) {
      std::vector<double> r(2 * nv);
      eval_frame_double(P, f, P.params.data(), r.data());
      double s = 0; for (double v : r) s += v * v;
      double rho[3] = {s, 1, 0};
      if (use_loss) huber(s, rho);
      c = 0.5 * rho[0];
    }
    if (per_frame) per_frame[f] = c;
    tot += c;
  }
  return tot;
}
// Autodiff Jacobian of one frame w.r.t. ALL its block parameters, raw (no loss):
// J row-major (2*nvalid) x (cam_params_total + 6), columns = params layout then the frame's pose.
void orc_frame_jacobian(void* h, int f, double* J, double* r) {
  Problem& P = *(Problem*)h;
  const size_t nv = P.frame_num_valid(f);
  if (!nv) return;
  const int ct = P.cam_params_total();
  std::vector<int> out_col(ct + 6);
  for (int i = 0; i < ct + 6; i++) out_col[i] = i;
  eval_frame_jets<4>(P, f, P.params.data(), out_col, ct + 6, 2 * nv, r, J);
}
long long orc_frame_num_valid(void* h, int f) { return (long long)((Problem*)h)->frame_num_valid(f); }

// Normal-equation blocks at the current parameters: Huber-corrected, UNSCALED Jacobian.
// Hcc [C*C], gc [C], per ACTIVE frame slot: Hff [36], gf [6], Hfc [6*C]; returns number of active frames.
int orc_normal_blocks(void* h, int use_loss, int num_threads, double* cost, double* Hcc, double* gc, double* Hff, double* gf, double* Hfc, int* frame_ids) {
  Problem& P = *(Problem*)h;
  Options o; o.use_loss = use_loss; o.num_threads = std::max(1, num_threads);
  Solver S(P, o);
  Eval E;
  S.evaluate(P.params, E);
  Solver::NormalBlocks N;
  S.normal_blocks(E, N);
  if (cost) *cost = E.cost;
  const int C = S.R.C, nf = (int)S.R.frames.size();
  if (Hcc) std::memcpy(Hcc, N.Hcc.data(), sizeof(double) * C * C);
  if (gc) std::memcpy(gc, N.gc.data(), sizeof(double) * C);
  if (Hff) std::memcpy(Hff, N.Hff.data(), sizeof(double) * 36 * nf);
  if (gf) std::memcpy(gf, N.gf.data(), sizeof(double) * 6 * nf);
  if (Hfc) std::memcpy(Hfc, N.Hfc.data(), sizeof(double) * 6 * C * nf);
  if (frame_ids) for (int s = 0; s < nf; s++) frame_ids[s] = S.R.frames[s];
  return nf;
}

// calibrator.cpp:285-313 (+ Ceres). Starts from the CURRENT params (call orc_init_params first to
// reproduce runCalibration's params_.clear(); initializeVariables()).
int orc_solve(void* h, const orc_options* opt, orc_summary* sum) {
  Problem& P = *(Problem*)h;
  Solver S(P, to_options(opt));
  S.solve(P.params, *sum);
  return sum->termination;
}

// calibrator.cpp:372-413 ; returns number of points dropped
long long orc_remove_outliers(void* h, double max_err) {
  Problem& P = *(Problem*)h;
  const double m2 = max_err * max_err;
  long long dropped = 0;
  for (int f = 0; f < P.nframes(); f++) {
    const size_t nv = P.frame_num_valid(f);
    if (!nv) continue;
    std::vector<double> r(2 * nv);
    eval_frame_double(P, f, P.params.data(), r.data());
    size_t idx = 0;
    for (int k = 0; k < P.ncam(); k++) {
      if (!P.cams[k].active) continue;
      for (uint8_t& m : P.frames[f][k].mask) {
        if (!m) continue;
        const double e = r[2 * idx] * r[2 * idx] + r[2 * idx + 1] * r[2 * idx + 1];
        if (e > m2) { m = 0; dropped++; }
        idx++;
      }
    }
  }
  return dropped;
}

// calibrator.cpp:415-463 : camera k -> intr[4], dist[n_dist], T_cn_cnm1[16]
void orc_get_camera(void* h, int k, double* intr, double* dist, double* T_cn_cnm1) {
  Problem& P = *(Problem*)h;
  const int o = P.intr_offset(k);
  for (int i = 0; i < P.cams[k].n_intr; i++) intr[i] = P.params[o + i];
  for (int i = 0; i < P.cams[k].n_dist; i++) dist[i] = P.params[o + P.cams[k].n_intr + i];
  double Tk[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}, Tkm1[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  if (k > 0) {
    rvec_tvec_to_transform(&P.params[P.ext_offset(k)], Tk);
    if (k > 1) rvec_tvec_to_transform(&P.params[P.ext_offset(k - 1)], Tkm1);
    double inv[16];
    mat4_inv(Tkm1, inv);
    mat4_mul(Tk, inv, T_cn_cnm1);
  } else {
    std::memcpy(T_cn_cnm1, P.cams[0].T_cn_cnm1, sizeof(Tk));   // cam0 keeps what it was given
  }
}

void orc_transform_to_rvec_tvec(const double* T, double* out) { transform_to_rvec_tvec(T, out); }
void orc_rvec_tvec_to_transform(const double* in, double* T) { rvec_tvec_to_transform(in, T); }
void orc_rotation_matrix(const double* v, double* R) { rotation_matrix<double>(v, R); }
int orc_hw_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

}  // extern "C"
