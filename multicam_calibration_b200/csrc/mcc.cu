// mcc.cu -- handle, host-side Levenberg-Marquardt driver and the C ABI of include/mcc.h.
//
// Host logic mirrors what the reference does around ceres::Solve:
//   * parameter packing / unpacking : /root/reference/src/calibrator.cpp:36-71, 157-170, 251-283, 415-463
//   * problem structure             : calibrator.cpp:353-370, 466-523 (one Huber(1.0) block per frame with >=1
//                                     valid point of an active camera; constant intrinsic / extrinsic blocks)
//   * trust-region loop             : Ceres' TrustRegionMinimizer + LevenbergMarquardtStrategy with the options of
//                                     calibrator.cpp:300-305
// All arithmetic of the solve runs in the kernels of kernels.cuh; the host only sequences launches, reads a
// handful of scalars per iteration and takes the accept / reject decision.  There is no CPU fallback.
#include "../../include/mcc.h"

#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "kernels.cuh"

namespace {

using namespace mcc;

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// ------------------------------------------------------------------------------------------------
// NCCL, resolved at run time from the libnccl already loaded in the process (torch's) or the system one
// ------------------------------------------------------------------------------------------------
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess_ = 0 };
enum { ncclSum_ = 0, ncclMax_ = 2 };
enum { ncclInt32_ = 2, ncclFloat64_ = 8 };
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(ncclUniqueId*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  // symmetric-memory windows (NCCL >= 2.27), optional
  int (*MemAlloc)(void**, size_t) = nullptr;
  int (*MemFree)(void*) = nullptr;
  int (*CommWindowRegister)(ncclComm_t, void*, size_t, void**, int) = nullptr;
  int (*CommWindowDeregister)(ncclComm_t, void*) = nullptr;
  bool has_windows() const { return MemAlloc && MemFree && CommWindowRegister && CommWindowDeregister; }
  bool ok() const { return lib && GetUniqueId && CommInitRank && AllReduce && GroupStart && GroupEnd; }
};
NcclApi& nccl() {
  static NcclApi api;
  static bool tried = false;
  if (!tried) {
    tried = true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (api.lib) break;
    }
    if (api.lib) {
      api.GetUniqueId = (int (*)(ncclUniqueId*))dlsym(api.lib, "ncclGetUniqueId");
      api.CommInitRank = (int (*)(ncclComm_t*, int, ncclUniqueId, int))dlsym(api.lib, "ncclCommInitRank");
      api.CommDestroy = (int (*)(ncclComm_t))dlsym(api.lib, "ncclCommDestroy");
      api.AllReduce = (int (*)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t))dlsym(api.lib, "ncclAllReduce");
      api.GroupStart = (int (*)())dlsym(api.lib, "ncclGroupStart");
      api.GroupEnd = (int (*)())dlsym(api.lib, "ncclGroupEnd");
      api.GetErrorString = (const char* (*)(int))dlsym(api.lib, "ncclGetErrorString");
      api.MemAlloc = (int (*)(void**, size_t))dlsym(api.lib, "ncclMemAlloc");
      api.MemFree = (int (*)(void*))dlsym(api.lib, "ncclMemFree");
      api.CommWindowRegister = (int (*)(ncclComm_t, void*, size_t, void**, int))dlsym(api.lib, "ncclCommWindowRegister");
      api.CommWindowDeregister = (int (*)(ncclComm_t, void*))dlsym(api.lib, "ncclCommWindowDeregister");
    }
  }
  return api;
}

// ------------------------------------------------------------------------------------------------
// host-side transforms
// ------------------------------------------------------------------------------------------------
// utils.h:29-39
void h_rotation_matrix(const double* v, double* R) {
  const double n = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  double ang = 0.0, a[3] = {0.0, 0.0, 1.0};
  if (n > 1e-8) { a[0] = v[0] / n; a[1] = v[1] / n; a[2] = v[2] / n; ang = n; }
  const double s = std::sin(ang), c = std::cos(ang), k = 1.0 - c;
  R[0] = k * a[0] * a[0] + c;        R[1] = k * a[0] * a[1] - s * a[2]; R[2] = k * a[0] * a[2] + s * a[1];
  R[3] = k * a[0] * a[1] + s * a[2]; R[4] = k * a[1] * a[1] + c;        R[5] = k * a[1] * a[2] - s * a[0];
  R[6] = k * a[0] * a[2] - s * a[1]; R[7] = k * a[1] * a[2] + s * a[0]; R[8] = k * a[2] * a[2] + c;
}
// calibrator.cpp:157-170: rotation block -> unit quaternion (largest-component branch) -> angle in [0,pi] * axis
void h_transform_to_rvec_tvec(const double* T, double* out) {
  const double m00 = T[0], m01 = T[1], m02 = T[2], m10 = T[4], m11 = T[5], m12 = T[6], m20 = T[8], m21 = T[9], m22 = T[10];
  double qw, qx, qy, qz;
  const double tr = m00 + m11 + m22;
  if (tr > 0.0) {
    const double t = std::sqrt(tr + 1.0), h = 0.5 / t;
    qw = 0.5 * t; qx = (m21 - m12) * h; qy = (m02 - m20) * h; qz = (m10 - m01) * h;
  } else if (m00 >= m11 && m00 >= m22) {
    const double t = std::sqrt(m00 - m11 - m22 + 1.0), h = 0.5 / t;
    qx = 0.5 * t; qw = (m21 - m12) * h; qy = (m10 + m01) * h; qz = (m20 + m02) * h;
  } else if (m11 > m00 && m11 >= m22) {
    const double t = std::sqrt(m11 - m22 - m00 + 1.0), h = 0.5 / t;
    qy = 0.5 * t; qw = (m02 - m20) * h; qz = (m21 + m12) * h; qx = (m01 + m10) * h;
  } else {
    const double t = std::sqrt(m22 - m00 - m11 + 1.0), h = 0.5 / t;
    qz = 0.5 * t; qw = (m10 - m01) * h; qx = (m02 + m20) * h; qy = (m12 + m21) * h;
  }
  double n = std::sqrt(qx * qx + qy * qy + qz * qz);
  double ang = 0.0, ax = 1.0, ay = 0.0, az = 0.0;
  if (n != 0.0) {
    ang = 2.0 * std::atan2(n, std::fabs(qw));
    if (qw < 0.0) n = -n;
    ax = qx / n; ay = qy / n; az = qz / n;
  }
  out[0] = ax * ang; out[1] = ay * ang; out[2] = az * ang;
  out[3] = T[3]; out[4] = T[7]; out[5] = T[11];
}
void h_rvec_tvec_to_transform(const double* in, double* T) {
  double R[9];
  h_rotation_matrix(in, R);
  for (int i = 0; i < 3; i++) { T[i * 4 + 0] = R[i * 3 + 0]; T[i * 4 + 1] = R[i * 3 + 1]; T[i * 4 + 2] = R[i * 3 + 2]; T[i * 4 + 3] = in[3 + i]; }
  T[12] = T[13] = T[14] = 0.0; T[15] = 1.0;
}
void h_mat4_mul(const double* A, const double* B, double* C) {
  double t[16];
  for (int i = 0; i < 4; i++)
    for (int j = 0; j < 4; j++) t[i * 4 + j] = A[i * 4 + 0] * B[0 * 4 + j] + A[i * 4 + 1] * B[1 * 4 + j] + A[i * 4 + 2] * B[2 * 4 + j] + A[i * 4 + 3] * B[3 * 4 + j];
  std::memcpy(C, t, sizeof(t));
}
// 4x4 inverse (Gauss-Jordan, partial pivoting)
void h_mat4_inv(const double* A, double* Ai) {
  double a[4][8];
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { a[i][j] = A[i * 4 + j]; a[i][4 + j] = i == j ? 1.0 : 0.0; }
  for (int c = 0; c < 4; c++) {
    int p = c;
    for (int r = c + 1; r < 4; r++) if (std::fabs(a[r][c]) > std::fabs(a[p][c])) p = r;
    if (p != c) for (int j = 0; j < 8; j++) std::swap(a[c][j], a[p][j]);
    const double d = 1.0 / a[c][c];
    for (int j = 0; j < 8; j++) a[c][j] *= d;
    for (int r = 0; r < 4; r++) if (r != c) { const double f = a[r][c]; if (f != 0.0) for (int j = 0; j < 8; j++) a[r][j] -= f * a[c][j]; }
  }
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) Ai[i * 4 + j] = a[i][4 + j];
}

struct CamHost {
  int model = 0, n_intr = 4, n_dist = 4;
  double intr[4] = {0, 0, 0, 0}, dist[MCC_MAX_DIST] = {0, 0, 0, 0, 0, 0, 0, 0}, T[16];
  int fix_intr = 0, fix_extr = 0, active = 1;
  int nparam() const { return n_intr + n_dist; }
};

// device buffer that only grows
template <typename T>
struct DBuf {
  T* p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t n) {
    if (n <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    cudaError_t e = cudaMalloc((void**)&p, std::max<size_t>(n, 1) * sizeof(T));
    if (e == cudaSuccess) cap = n;
    return e;
  }
  // grow keeping the first `keep` elements (device-to-device copy on stream st)
  cudaError_t reserve_keep(size_t n, size_t keep, cudaStream_t st) {
    if (n <= cap) return cudaSuccess;
    const size_t ncap = std::max(n, cap + cap / 2);
    T* q = nullptr;
    cudaError_t e = cudaMalloc((void**)&q, std::max<size_t>(ncap, 1) * sizeof(T));
    if (e != cudaSuccess) return e;
    if (p && keep) {
      e = cudaMemcpyAsync(q, p, keep * sizeof(T), cudaMemcpyDeviceToDevice, st);
      if (e == cudaSuccess) e = cudaStreamSynchronize(st);
      if (e != cudaSuccess) { cudaFree(q); return e; }
    }
    if (p) cudaFree(p);
    p = q; cap = ncap;
    return cudaSuccess;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
template <typename T>
struct PinBuf {
  T* p = nullptr;
  size_t cap = 0;
  bool reserve(size_t n, size_t keep = 0) {
    if (n <= cap) return true;
    size_t ncap = std::max(n, cap + cap / 2);
    T* q = nullptr;
    if (cudaHostAlloc((void**)&q, std::max<size_t>(ncap, 1) * sizeof(T), cudaHostAllocDefault) != cudaSuccess) {
      cudaGetLastError();
      // no CUDA runtime / device: plain memory keeps the host-side API (add / get / set) usable
      q = (T*)std::malloc(std::max<size_t>(ncap, 1) * sizeof(T));
      if (!q) return false;
      pinned = false;
    }
    if (p && keep) std::memcpy(q, p, keep * sizeof(T));
    release();
    p = q; cap = ncap;
    return true;
  }
  void release() { if (p) { if (pinned) cudaFreeHost(p); else std::free(p); } p = nullptr; cap = 0; }
  bool pinned = true;
};

void parallel_memcpy(void* dst, const void* src, size_t bytes) {
  const size_t kMin = 8u << 20;
  unsigned nt = std::min<unsigned>(8, std::max<unsigned>(1, std::thread::hardware_concurrency()));
  if (bytes < 2 * kMin || nt <= 1) { std::memcpy(dst, src, bytes); return; }
  nt = (unsigned)std::min<size_t>(nt, bytes / kMin);
  std::vector<std::thread> th;
  const size_t per = (bytes + nt - 1) / nt;
  for (unsigned i = 0; i < nt; i++) {
    const size_t o = i * per;
    if (o >= bytes) break;
    const size_t n = std::min(per, bytes - o);
    th.emplace_back([=] { std::memcpy((char*)dst + o, (const char*)src + o, n); });
  }
  for (auto& t : th) t.join();
}

// kernel categories for the optional per-launch CUDA-event timing (mcc_set_kernel_timing)
enum { KC_SETUP = 0, KC_TILE_NORMAL, KC_ASSEMBLE, KC_SCHUR, KC_SOLVE, KC_BACKSUB, KC_TILE_COST, KC_REDUCE, KC_OTHER, KC_EXCHANGE, KC_COUNT };
struct KPending { cudaEvent_t a, b; int cat; };

// FP64 FMA throughput probe: 8 independent dependent-chains per thread
__global__ void k_dfma_peak(double* out, int iters) {
  const double b = 1.0000000001, c = 1e-12;
  double a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  for (int i = 0; i < iters; i++) {
    a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c);
    a4 = fma(a4, b, c); a5 = fma(a5, b, c); a6 = fma(a6, b, c); a7 = fma(a7, b, c);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
}

// AoS -> SoA on the device (observations arrive as Point3 / Point2 arrays, types.h:17-47)
__global__ void k_aos_to_soa(size_t n, const double* __restrict__ w, const double* __restrict__ im, double* __restrict__ X,
                             double* __restrict__ Y, double* __restrict__ Z, double* __restrict__ U, double* __restrict__ V) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  X[i] = w[3 * i]; Y[i] = w[3 * i + 1]; Z[i] = w[3 * i + 2];
  U[i] = im[2 * i]; V[i] = im[2 * i + 1];
}

}  // namespace

// =================================================================================================
struct mcc_handle {
  // ---- host problem (what Calibrator keeps, calibrator.h:42-48)
  std::vector<CamHost> cams;
  bool fix_all = false;
  std::vector<int> npts;               // [F * ncam]
  PinBuf<double> world, image;         // AoS staging: 3N, 2N
  std::vector<double> poses;           // 16 F
  std::vector<uint8_t> mask;           // N, materialised lazily: while mask_all_one every point is valid
  bool mask_all_one = true;
  size_t nobs = 0;
  size_t host_count = 0;               // observations [0, host_count) are present in the host staging buffers
  size_t dev_count = 0;                // observations [0, dev_count) are present in the device AoS buffers
  int nframes = 0;
  std::vector<double> params;          // calibrator.cpp:36-46 layout
  bool params_valid = false;
  std::vector<double> params_init;     // what initializeVariables() made of the current cameras and pose guesses (runCalibration
  bool init_valid = false;             // starts from it every time; recomputed only when cameras or frames change)
  PinBuf<double> pose_stage;           // pinned staging of the frame poses for the parameter upload / download
  std::string err;
  int device = -1;

  // ---- structure (rebuilt when cameras / flags / frames / masks change)
  bool structure_dirty = true, obs_dirty = true, mask_dirty = true, params_dirty = true;
  int C_full = 0, C = 0, E = 0, nslots = 0, ntiles = 0, ngroups = 0, gstride = 0, rowstride = 0, npairs = 0, LD = 0;
  std::vector<int> intr_off, ext_off, cam_col, col_param, slot_frame, frame_slot;
  std::vector<size_t> group_obs0;      // [F*ncam] observation offset
  long long nres = 0;

  // ---- device
  cudaStream_t stream = nullptr;
  DBuf<double> d_aos_w, d_aos_i, d_X, d_Y, d_Z, d_U, d_V, d_res;
  DBuf<uint8_t> d_mask;
  DBuf<Tile> d_tiles;
  DBuf<int> d_gto, d_col_param, d_tab, d_gs_tab;
  DBuf<int2> d_dist;                   // hand-out lists of the Jacobian pass (per camera: pose x camera, camera x camera, pose x pose)
  DBuf<uint4> d_cls_li;                // lane records per camera class
  DBuf<double> d_Linv, d_zf;
  bool k1_temp_global = false;         // frames too large for shared memory: their tile Grams go through d_k1_temp
  DBuf<double> d_k1_temp;
  int ncls = 0, maxft = 1, k1_grid = 1, k1_fs = 0, cc_rows = 1;   // Jacobian pass: grid, statically dealt frames, rows of cc_part
  unsigned int* d_ticket = nullptr;    // [0] frame counter of the Jacobian pass, [1..] entry-block counters of k_reduce_misc
  DBuf<double> d_cc_part2;
  DBuf<CamMeta> d_cams;
  DBuf<CcEntry> d_cc;
  DBuf<double> d_camp[2], d_pose[2];   // current / candidate
  int cur = 0;
  DBuf<double> d_Hff2[2], d_gf2[2], d_Hfc2[2], d_Hcc2[2], d_gc2[2];   // block sets: current point / speculative candidate
  int blk = 0;
  DBuf<double> d_frot, d_camrec;       // rotation records of the Jacobian pass (k_rot_setup)
  DBuf<double> d_setup, d_tile_s, d_Lf, d_cost_f, d_gmax_f, d_xn2_f, d_newcost_f, d_mcc_f, d_dx2_f, d_xn2c_f;
  DBuf<double> d_cc_part, d_cc_sum, d_scale_c, d_scale_f, d_S_part, d_S_sum, d_M, d_yc, d_ysc, d_scal;
  // several ranks: the two all-reduced buffers live in NCCL symmetric-memory windows when the library has them
  double* win_S = nullptr; double* win_cc = nullptr;
  void* wh_S = nullptr; void* wh_cc = nullptr;
  size_t win_S_n = 0, win_cc_n = 0;
  // ... or, on one node with peer access, in a CUDA IPC window that every rank maps (k_peer_allreduce)
  PeerWin pw{};
  uint4* peer_own = nullptr;            // this rank's window (cudaMalloc)
  size_t peer_nS = 0, peer_nC = 0;      // slot sizes the window was laid out for
  unsigned long long peer_seq[3] = {0, 0, 0};
  DBuf<double> d_small;                 // [4 scalars | failure flag per rank] of a non-speculative step (several ranks)
  bool peer_on = false, peer_failed = false;
  long long peer_timeout_cycles = 0;    // k_peer_allreduce gives up after this many SM clocks (MCC_PEER_TIMEOUT_S, default 60 s)
  // iterations queued ahead of their accept / reject decision (solve(): pipelined section)
  LmState* d_state = nullptr;
  const LmState* st_cur = nullptr;      // what every launch passes: d_state while the pipeline runs, else nullptr
  DecideArgs decide{};                  // st == nullptr: k_scatter_cc takes no decision
  double* d_rec = nullptr;              // [2][LM_REC] decision records
  double* h_rec = nullptr;              // pinned mirror
  cudaEvent_t ev_rec[2] = {nullptr, nullptr};
  // one captured CUDA graph per (parameter set, block set, record slot) of a queued iteration; rebuilt when the
  // structure or the options change (single rank only: the peer exchange takes a new sequence number per launch)
  cudaGraphExec_t it_graph[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  long long it_graph_kernels[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // kernels inside each graph, for the launch count
  double it_graph_key[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};   // the option values baked into the graphs
  double* h_scal = nullptr;            // pinned mirror of d_scal
  int sy_gx = 0, sy_gy = 0;
  int sms = 0;
  bool attrs_set = false;
  long long launches = 0, h2d = 0, d2h = 0;
  // optional per-launch timing
  bool ktiming = false;
  std::vector<cudaEvent_t> ev_pool;
  std::vector<KPending> ev_pending;
  double kc_ms[KC_COUNT] = {0};
  long long kc_n[KC_COUNT] = {0};
  cudaEvent_t tm_a = nullptr, tm_b = nullptr;

  // ---- comm
  ncclComm_t comm = nullptr;
  int nranks = 1, rank = 0;
  // ---- several devices behind ONE handle (mcc_set_devices): this handle is then only a front that deals the frames out to one
  // member handle per device and runs them on one host thread each; a member finds its peers through `member_of`
  struct mcc_group* group = nullptr;      // front
  struct mcc_group* member_of = nullptr;  // member

  int ncam() const { return (int)cams.size(); }
};

// The members of a multi-device handle and what they need to meet each other inside one process.
struct mcc_group {
  std::vector<mcc_handle*> kids;
  std::vector<int> frame_kid, frame_local;          // per frame of the front: which member holds it, and as which of its frames
  std::vector<std::vector<int>> kid_frames;         // per member: its frames (front numbering), in the member's order
  std::vector<long long> kid_obs;                   // observations per member (balance)
  bool cams_dirty = true, params_pushed = false;
  // rendezvous of the member threads (peer-window set-up): a counting barrier and a table of window addresses
  std::mutex mu;
  std::condition_variable cv;
  int arrived = 0;
  long generation = 0;
  std::vector<void*> win;
  int bad = 0;
  void barrier() {
    std::unique_lock<std::mutex> lk(mu);
    const long g = generation;
    if (++arrived == (int)kids.size()) { arrived = 0; generation++; cv.notify_all(); }
    else cv.wait(lk, [&] { return generation != g; });
  }
};

namespace {

mcc_status fail(mcc_handle* h, mcc_status st, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (h) h->err = buf;
  return st;
}
#define CU(call)                                                                                         \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess) return fail(h, MCC_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)
#define ST(call)                       \
  do {                                 \
    mcc_status s_ = (call);            \
    if (s_ != MCC_OK) return s_;       \
  } while (0)
#define NC(call)                                                                                                      \
  do {                                                                                                                \
    int r_ = (call);                                                                                                  \
    if (r_ != 0) return fail(h, MCC_ERR_NCCL, "%s failed: %s", #call, nccl().GetErrorString ? nccl().GetErrorString(r_) : "?"); \
  } while (0)

// host mask, created on first need (addPoints marks every point valid, calibrator.cpp:246)
void materialize_mask(mcc_handle* h) {
  if (h->mask.size() != h->nobs) h->mask.resize(h->nobs, 1);
}

// Launch of a kernel of the LM iteration.  (Programmatic dependent launch between them -- cudaLaunchAttributeProgrammaticStreamSerialization
// plus griddepcontrol.wait / launch_dependents at the top of every kernel -- was measured: 6.00 instead of 5.85 ms per step; dropped.)
template <typename... KArgs, typename... Args>
cudaError_t launch_k(mcc_handle* h, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = h->stream;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

struct KScope {
  mcc_handle* h;
  KPending p;
  bool on;
  KScope(mcc_handle* hh, int cat) : h(hh), on(hh->ktiming) {
    h->launches++;
    if (!on) return;
    auto get = [&]() {
      cudaEvent_t e = nullptr;
      if (!h->ev_pool.empty()) { e = h->ev_pool.back(); h->ev_pool.pop_back(); }
      else cudaEventCreate(&e);
      return e;
    };
    p.a = get(); p.b = get(); p.cat = cat;
    cudaEventRecord(p.a, h->stream);
  }
  ~KScope() {
    if (!on) return;
    cudaEventRecord(p.b, h->stream);
    h->ev_pending.push_back(p);
  }
};
void resolve_ktiming(mcc_handle* h) {   // call after a stream synchronize
  for (KPending& p : h->ev_pending) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) { h->kc_ms[p.cat] += ms; h->kc_n[p.cat]++; }
    h->ev_pool.push_back(p.a); h->ev_pool.push_back(p.b);
  }
  h->ev_pending.clear();
}

mcc_status ensure_device(mcc_handle* h) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail(h, MCC_ERR_NO_DEVICE, "no CUDA device available (%s); this library has no CPU fallback", e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
  }
  if (h->device < 0) CU(cudaGetDevice(&h->device));
  CU(cudaSetDevice(h->device));
  if (!h->stream) CU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  if (!h->h_scal) CU(cudaHostAlloc((void**)&h->h_scal, 33 * sizeof(double), cudaHostAllocDefault));
  if (!h->d_state) {
    CU(cudaMalloc((void**)&h->d_state, sizeof(LmState)));
    CU(cudaMalloc((void**)&h->d_ticket, 64 * sizeof(unsigned int)));
    CU(cudaMemset(h->d_ticket, 0, 64 * sizeof(unsigned int)));   // both kernels leave their counters at zero
    CU(cudaHostAlloc((void**)&h->h_rec, 2 * LM_REC * sizeof(double), cudaHostAllocMapped));   // decide_body writes the records straight to the host
    CU(cudaHostGetDevicePointer((void**)&h->d_rec, h->h_rec, 0));
    for (int i = 0; i < 2; i++) CU(cudaEventCreateWithFlags(&h->ev_rec[i], cudaEventDisableTiming));
  }
  if (!h->attrs_set) {
    CU((cudaFuncSetAttribute(k_frame_normal<15>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)));
    CU((cudaFuncSetAttribute(k_frame_normal<21>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)));
    CU(cudaFuncSetAttribute(k_group_setup, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SETUP_SMEM_BYTES));
    CU((cudaFuncSetAttribute(k_schur_syrk<true, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)));
    CU((cudaFuncSetAttribute(k_schur_syrk<true, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)));
    CU((cudaFuncSetAttribute(k_schur_syrk<false, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)));
    CU(cudaFuncSetAttribute(k_solve_cam<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 184 * 1024));
    h->attrs_set = true;
  }
  return MCC_OK;
}

// calibrator.cpp:251-283
void init_params(mcc_handle* h) {
  std::vector<double>& p = h->params;
  if (h->init_valid) {      // same cameras, same pose guesses: same initial values (0.5 ms of host time per run on 10 000 frames)
    p = h->params_init;
    h->params_valid = true;
    h->params_dirty = true;
    return;
  }
  p.clear();
  for (const CamHost& c : h->cams) {
    p.insert(p.end(), c.intr, c.intr + c.n_intr);
    p.insert(p.end(), c.dist, c.dist + c.n_dist);
  }
  double prev[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  for (int k = 1; k < h->ncam(); k++) {
    double Tn0[16], rt[6];
    h_mat4_mul(h->cams[k].T, prev, Tn0);
    h_transform_to_rvec_tvec(Tn0, rt);
    p.insert(p.end(), rt, rt + 6);
    std::memcpy(prev, Tn0, sizeof(prev));
  }
  for (int f = 0; f < h->nframes; f++) {
    double rt[6];
    h_transform_to_rvec_tvec(&h->poses[16 * (size_t)f], rt);
    p.insert(p.end(), rt, rt + 6);
  }
  h->params_init = p;
  h->init_valid = true;
  h->params_valid = true;
  h->params_dirty = true;
}


// Everything that depends on cameras, flags, frame list and masks: reduced column maps
// (calibrator.cpp:510-521), frames with residual blocks (calibrator.cpp:482-485), tiles.
void drop_iteration_graphs(mcc_handle* h) {
  for (cudaGraphExec_t& g : h->it_graph) { if (g) cudaGraphExecDestroy(g); g = nullptr; }
}

// contraction units of the reduced (n_k + 7)-column Gram matrix: unit u = block column u with itself and with the block
// columns at circular distance 1 .. (nb-1)/2; for an even block count the distance nb/2 pairs are taken by the lower
// half of the units.  Together the units cover every block pair exactly once.
static void build_units(CamMeta& m, int n_dist) {
  const int rnb = (4 + n_dist + 7 + 2) / 3;
  m.nunits = rnb;
  for (int d = 0; d < 3; d++) for (int u = 0; u < 8; u++) m.unit_o[d][u] = -1;
  for (int u = 0; u < rnb; u++) {
    int nd_ = 0;
    for (int d = 1; d <= (rnb - 1) / 2; d++) m.unit_o[nd_++][u] = (int8_t)((u + d) % rnb);
    if (rnb % 2 == 0 && u < rnb / 2) m.unit_o[nd_++][u] = (int8_t)(u + rnb / 2);
  }
}

// Lane -> (unit, slice) map of the K1 contraction.  A 64-bit shared load is served one half-warp at a time, so the map
// is searched (backtracking, bounded) for one where the <= 16 lanes of each half-warp hit different 8-byte banks for
// the unit's own block column and for every partner column; slice q reads rows q, q + ksl, ... and the row stride rs
// is odd.  Slice counts of the form 2^n or 3 * 2^n only (the reduction handles those).  Falls back to
// lane = q * nunits + u when no conflict-free map exists.
struct LaneSearch {
  int n;
  int bank[64][4];
  int half_of[64];
  int cnt[2];
  bool used[2][4][16];
  long nodes;
  bool go(int i) {
    if (++nodes > 200000) return false;
    if (i == n) return true;
    for (int hh = 0; hh < 2; hh++) {
      if (cnt[hh] >= 16) continue;
      bool ok = true;
      for (int d = 0; d < 4; d++) if (used[hh][d][bank[i][d]]) ok = false;
      if (!ok) continue;
      for (int d = 0; d < 4; d++) used[hh][d][bank[i][d]] = true;
      cnt[hh]++; half_of[i] = hh;
      if (go(i + 1)) return true;
      cnt[hh]--;
      for (int d = 0; d < 4; d++) used[hh][d][bank[i][d]] = false;
    }
    return false;
  }
};

static void build_lane_map(CamMeta& m, int rs) {
  const int nun = m.nunits;
  static const int cand[] = {8, 6, 4, 3, 2, 1};
  int lane_of[8][8];
  bool found = false;
  int ksl = 1;
  for (int ci = 0; ci < 6 && !found && (rs & 1); ci++) {
    ksl = cand[ci];
    if (nun * ksl > 32) continue;
    LaneSearch ls;
    ls.n = nun * ksl; ls.nodes = 0; ls.cnt[0] = ls.cnt[1] = 0;
    memset(ls.used, 0, sizeof(ls.used));
    for (int u = 0; u < nun; u++)
      for (int q = 0; q < ksl; q++)
        for (int d = 0; d < 4; d++) {
          int b = u;
          if (d > 0 && m.unit_o[d - 1][u] >= 0) b = m.unit_o[d - 1][u];
          ls.bank[u * ksl + q][d] = (rs * q + 3 * b) & 15;
        }
    if (ls.go(0)) {
      found = true;
      int next[2] = {0, 16};
      for (int u = 0; u < nun; u++)
        for (int q = 0; q < ksl; q++) lane_of[u][q] = next[ls.half_of[u * ksl + q]]++;
    }
  }
  if (!found) {
    ksl = 32 / nun;
    if (ksl == 5 || ksl == 7) ksl--;
    for (int u = 0; u < nun; u++) for (int q = 0; q < ksl; q++) lane_of[u][q] = q * nun + u;
  }
  m.ksl = ksl;
  for (int l = 0; l < 32; l++) { m.lane_u[l] = -1; m.lane_q[l] = -1; }
  for (int u = 0; u < nun; u++) for (int q = 0; q < ksl; q++) { m.lane_u[lane_of[u][q]] = (int8_t)u; m.lane_q[lane_of[u][q]] = (int8_t)q; }
  int mm = ksl;
  m.nhalve = 0;
  for (int r = 0; r < 3; r++) for (int l = 0; l < 32; l++) m.red_src[r][l] = (int8_t)l;
  for (int r = 0; r < 2; r++) for (int l = 0; l < 32; l++) m.fin_src[r][l] = (int8_t)l;
  while ((mm & 1) == 0) {
    mm >>= 1;
    for (int u = 0; u < nun; u++) for (int q = 0; q < mm; q++) m.red_src[m.nhalve][lane_of[u][q]] = (int8_t)lane_of[u][q + mm];
    m.nhalve++;
  }
  m.mfin = mm;
  if (mm == 3) for (int u = 0; u < nun; u++) { m.fin_src[0][lane_of[u][0]] = (int8_t)lane_of[u][1]; m.fin_src[1][lane_of[u][0]] = (int8_t)lane_of[u][2]; }
  for (int l = 0; l < 32; l++) {
    int8_t b[16];
    std::memset(b, 0, sizeof(b));
    b[0] = m.lane_u[l]; b[1] = m.lane_q[l];
    for (int d = 0; d < 3; d++) b[2 + d] = m.lane_u[l] >= 0 ? m.unit_o[d][m.lane_u[l]] : (int8_t)-1;
    for (int r = 0; r < 3; r++) b[5 + r] = m.red_src[r][l];
    b[8] = m.fin_src[0][l]; b[9] = m.fin_src[1][l];
    {
      int bi = 0, rem = l;   // l-th block of the upper triangle, row by row
      while (bi < m.nb && rem >= m.nb - bi) { rem -= m.nb - bi; bi++; }
      b[10] = (int8_t)bi; b[11] = (int8_t)(bi + rem);
    }
    std::memcpy(&m.lane_info[l], b, 16);
  }
}

// Entry (i, j) of a camera's Gram matrix over the full column layout [ext | pose | intr | r] as the hand-out phase of
// the Jacobian pass finds it (kernels.cuh): geometry x geometry (ext / pose columns) is computed there and coded
// a | b << 8 with a <= b; everything else already sits in the warp's shared memory and is coded by its offset from
// `rows`: geometry x other = Y[reduced column of the other][geometry column] (Y follows the reduced Gram matrix Gs),
// other x other = Gs[reduced][reduced]; the reduced column of intrinsic c is c - ng, of the residual column n_k + 6.
int ab_code(const CamMeta& m, int i, int j, int rs) {
  const int ng = m.intr_col;   // geometry columns come first
  if (i < ng && j < ng) { if (i > j) std::swap(i, j); return i | (j << 8); }
  if (j < ng) std::swap(i, j);
  const int nk = 4 + m.nd;
  const int rj = j < m.r_col ? j - ng : nk + 6;
  if (i < ng) return k1_gsrows(rs) * k1_gsld(rs) + rj * 12 + i;
  const int ri = i < m.r_col ? i - ng : nk + 6;
  return ri * k1_gsld(rs) + rj;
}
inline bool ab_is_geo(const CamMeta& m, int i, int j) { return i < m.intr_col && j < m.intr_col; }

mcc_status build_structure(mcc_handle* h) {
  const int ncam = h->ncam();
  const int F = h->nframes;
  h->intr_off.assign(ncam, 0); h->ext_off.assign(ncam, -1);
  int off = 0;
  for (int k = 0; k < ncam; k++) { h->intr_off[k] = off; off += h->cams[k].nparam(); }
  for (int k = 1; k < ncam; k++) { h->ext_off[k] = off; off += 6; }
  h->C_full = off;
  h->cam_col.assign(h->C_full, -1);
  h->col_param.clear();
  for (int k = 0; k < ncam; k++)
    if (!(h->fix_all || h->cams[k].fix_intr))
      for (int j = 0; j < h->cams[k].nparam(); j++) { h->cam_col[h->intr_off[k] + j] = (int)h->col_param.size(); h->col_param.push_back(h->intr_off[k] + j); }
  for (int k = 1; k < ncam; k++)
    if (!h->cams[k].fix_extr)
      for (int j = 0; j < 6; j++) { h->cam_col[h->ext_off[k] + j] = (int)h->col_param.size(); h->col_param.push_back(h->ext_off[k] + j); }
  h->C = (int)h->col_param.size();
  const int C = h->C;

  // camera metadata
  std::vector<CamMeta> cm(ncam);
  int max_nb = 1, max_nblk = 1, max_red_cols = 11;
  for (int k = 0; k < ncam; k++) {
    const CamHost& c = h->cams[k];
    CamMeta& m = cm[k];
    std::memset(&m, 0, sizeof(m));
    m.model = c.model; m.nd = c.n_dist; m.intr_off = h->intr_off[k]; m.ext_off = h->ext_off[k];
    // full local layout [ext 6 (k > 0) | pose 6 | intr n_k | r]: the geometry columns come first so that every
    // 3-wide block is purely "geometry" or purely "intrinsics / residual" (the expansion epilogue of K1 relies on it)
    m.ext_col = (k > 0) ? 0 : -1;
    m.pose_col = (k > 0) ? 6 : 0;
    m.intr_col = m.pose_col + 6;
    m.r_col = m.intr_col + 4 + c.n_dist;
    m.ncols = m.r_col + 1;
    m.nb = (m.ncols + 2) / 3;
    m.nblk = m.nb * (m.nb + 1) / 2;
    m.active = c.active;
    for (int j = 0; j < MAXC; j++) m.red_col[j] = -1;
    for (int j = 0; j < c.nparam(); j++) m.red_col[m.intr_col + j] = h->cam_col[h->intr_off[k] + j];
    if (k > 0) for (int j = 0; j < 6; j++) m.red_col[m.ext_col + j] = h->cam_col[h->ext_off[k] + j];
    build_units(m, c.n_dist);
    if (c.active) { max_nb = std::max(max_nb, m.nb); max_nblk = std::max(max_nblk, m.nblk); max_red_cols = std::max(max_red_cols, 4 + c.n_dist + 7); }
  }
  h->gstride = 9 * max_nblk;
  // k_frame_normal<RS> works on reduced rows of (4 + n_dist) + 7 columns
  h->rowstride = (max_red_cols + 2) / 3 <= 5 ? 15 : 21;
  for (int k = 0; k < ncam; k++) build_lane_map(cm[k], h->rowstride);
  // lane-record classes: the record depends on the unit layout (n_dist) and on the block count of the full layout
  // (n_dist, extrinsics or not); cameras of one class share a copy in the shared memory of the Jacobian pass
  std::vector<uint4> cls_li;
  std::vector<int> gs_tab;
  {
    std::vector<std::pair<int, int>> classes;
    for (int k = 0; k < ncam; k++) {
      const std::pair<int, int> key(h->cams[k].n_dist, k > 0 ? 1 : 0);
      int ci = -1;
      for (size_t i = 0; i < classes.size(); i++) if (classes[i] == key) ci = (int)i;
      if (ci < 0) {
        ci = (int)classes.size(); classes.push_back(key); cls_li.insert(cls_li.end(), cm[k].lane_info, cm[k].lane_info + 32);
        // where element (unit u, entry e) of a parked tile Gram goes in the symmetric matrix Gs of the expansion:
        // e < 6 the upper triangle of the unit's own 3x3 block, then 9 entries per partner block column
        const int nun = h->rowstride / 3, nacc = k1_nacc(h->rowstride), ld = k1_gsld(h->rowstride);
        for (int u = 0; u < nun; u++)
          for (int e = 0; e < nacc; e++) {
            int i = -1, j = -1;
            if (u < cm[k].nunits) {
              if (e < 6) { static const int ti[6] = {0, 0, 0, 1, 1, 2}, tj[6] = {0, 1, 2, 1, 2, 2}; i = 3 * u + ti[e]; j = 3 * u + tj[e]; }
              else { const int d = (e - 6) / 9, r = (e - 6) % 9; if (cm[k].unit_o[d][u] >= 0) { i = 3 * u + r / 3; j = 3 * cm[k].unit_o[d][u] + r % 3; } }
            }
            gs_tab.push_back(i < 0 ? -1 : ((i * ld + j) | ((j * ld + i) << 16)));
          }
      }
      cm[k].cls = ci;
    }
    h->ncls = (int)classes.size();
  }

  // camera-camera entries
  std::vector<CcEntry> cc;
  for (int k = 0; k < ncam; k++) {
    if (!h->cams[k].active) continue;
    const int ncc = cm[k].r_col;   // every local column but the residual; pose columns carry red_col = -1
    // two rounds: the geometry x geometry entries of a camera (extrinsics x extrinsics) come first in its list
    for (int geo = 1; geo >= 0; geo--)
      for (int i = 0; i < ncc; i++) {
        if (cm[k].red_col[i] < 0) continue;
        for (int j = i; j < ncc; j++) {
          if (cm[k].red_col[j] < 0 || (int)ab_is_geo(cm[k], i, j) != geo) continue;
          cc.push_back(CcEntry{k, ab_code(cm[k], i, j, h->rowstride), cm[k].red_col[i], cm[k].red_col[j]});
          cm[k].n_cc_g += geo;
        }
        if (!geo) cc.push_back(CcEntry{k, ab_code(cm[k], i, cm[k].r_col, h->rowstride), cm[k].red_col[i], -1});
      }
  }
  h->E = (int)cc.size();
  // entry range of every camera in cc[] (entries were pushed camera by camera)
  std::vector<int> cc_e0(ncam + 1, 0);
  {
    size_t e = 0;
    for (int k = 0; k < ncam; k++) { cc_e0[k] = (int)e; while (e < cc.size() && cc[e].cam == k) e++; }
    cc_e0[ncam] = (int)e;
  }
  // hand-out lists of the Jacobian pass: which entries (a, b) of a camera's Gram matrix over the full column layout
  // [ext | pose | intr | r] go where (ab_code packs and orders the pair the way the kernel's `entry` expects it).
  //   pose x camera columns  -> H_fc[frame][a * C + reduced column]
  //   camera x camera (+ camera gradient) entries -> index into the camera-camera entry vector
  //   pose x pose (+ pose gradient) -> entry a * 7 + b of the frame's 6 x 7 block (b == 6: gradient)
  std::vector<int2> dist;
  for (int k = 0; k < ncam; k++) {
    CamMeta& m = cm[k];
    m.dist_off = (int)dist.size();
    for (int geo = 1; geo >= 0; geo--)
      for (int a = 0; a < 6; a++)
        for (int j = 0; j < m.r_col; j++)
          if (m.red_col[j] >= 0 && (int)ab_is_geo(m, j, m.pose_col + a) == geo) {
            dist.push_back(make_int2(ab_code(m, j, m.pose_col + a, h->rowstride), a * C + m.red_col[j]));
            m.n_hfc_g += geo;
          }
    m.n_hfc = (int)dist.size() - m.dist_off;
    for (int e = cc_e0[k]; e < cc_e0[k + 1]; e++) dist.push_back(make_int2(cc[e].gram_off, e));
    m.n_cc = cc_e0[k + 1] - cc_e0[k];
    m.e0 = cc_e0[k];
    for (int grad = 0; grad < 2; grad++)       // the 36 pose x pose entries first, then the 6 gradient entries
      for (int ent = 0; ent < 42; ent++) {
        const int a = ent / 7, b2 = ent % 7;
        if ((b2 == 6) == (grad == 1)) dist.push_back(make_int2(ab_code(m, m.pose_col + a, b2 < 6 ? m.pose_col + b2 : m.r_col, h->rowstride), ent));
      }
  }

  // observation offsets, active frames, tiles
  h->group_obs0.assign((size_t)F * ncam + 1, 0);
  {
    size_t o = 0;
    for (size_t g = 0; g < (size_t)F * ncam; g++) { h->group_obs0[g] = o; o += h->npts[g]; }
    h->group_obs0[(size_t)F * ncam] = o;
  }
  h->slot_frame.clear();
  h->frame_slot.assign(F, -1);
  h->nres = 0;
  for (int f = 0; f < F; f++) {
    long long nv = 0;
    for (int k = 0; k < ncam; k++) {
      if (!h->cams[k].active) continue;
      const size_t o0 = h->group_obs0[(size_t)f * ncam + k], n = h->npts[(size_t)f * ncam + k];
      if (h->mask_all_one) nv += (long long)n;
      else for (size_t i = 0; i < n; i++) nv += h->mask[o0 + i] != 0;
    }
    if (nv > 0) { h->frame_slot[f] = (int)h->slot_frame.size(); h->slot_frame.push_back(f); h->nres += 2 * nv; }
  }
  h->nslots = (int)h->slot_frame.size();
  h->ngroups = h->nslots * ncam;
  std::vector<Tile> tiles;
  std::vector<int> gto((size_t)h->ngroups + 1, 0);
  const int chunks_per_tile = TILE_OBS / 32;
  for (int s = 0; s < h->nslots; s++) {
    const int f = h->slot_frame[s];
    for (int k = 0; k < ncam; k++) {
      gto[(size_t)s * ncam + k] = (int)tiles.size();
      const int n = h->npts[(size_t)f * ncam + k];
      if (!h->cams[k].active || n == 0) continue;
      const size_t o0 = h->group_obs0[(size_t)f * ncam + k];
      if (o0 + n > 0x7fffffffull) return fail(h, MCC_ERR_TOO_LARGE, "more than 2^31 observations on one device");
      const int nchunks = (n + 31) / 32;
      const int nt = (nchunks + chunks_per_tile - 1) / chunks_per_tile;
      const int base = nchunks / nt, rem = nchunks % nt;
      int done = 0;
      for (int t = 0; t < nt; t++) {
        const int cnt = std::min(n - done, (base + (t < rem ? 1 : 0)) * 32);
        tiles.push_back(Tile{(int)(o0 + done), cnt, s * ncam + k, k});
        done += cnt;
      }
    }
  }
  gto[h->ngroups] = (int)tiles.size();
  h->ntiles = (int)tiles.size();
  h->maxft = 1;
  for (int sl = 0; sl < h->nslots; sl++) h->maxft = std::max(h->maxft, gto[(size_t)(sl + 1) * ncam] - gto[(size_t)sl * ncam]);

  const int NT = (C + 1 + SY_T - 1) / SY_T;
  h->npairs = NT * (NT + 1) / 2;
  h->LD = NT * SY_TS;
  h->sy_gy = (h->npairs + SY_THREADS - 1) / SY_THREADS;
  if (h->sms <= 0) CU(cudaDeviceGetAttribute(&h->sms, cudaDevAttrMultiProcessorCount, h->device));
  const int sms = h->sms;
  h->sy_gx = std::max(1, std::min((h->nslots + SY_FB - 1) / SY_FB, std::max(1, (2 * sms) / h->sy_gy)));
  // the Jacobian pass: persistent CTAs, as many as fit at once (shared memory: rows of 4 warps + the frame's tile Grams)
  {
    // frames whose tile Grams would take more than ~100 KB park them in global memory instead (k1_temp_global)
    h->k1_temp_global = k1_smem_bytes(h->rowstride, ncam, h->maxft) > 100 * 1024;
    const size_t smem = k1_smem_bytes(h->rowstride, ncam, h->k1_temp_global ? 0 : h->maxft);
    if (smem > 200 * 1024) return fail(h, MCC_ERR_TOO_LARGE, "%d cameras need %zu bytes of shared memory in the Jacobian pass", ncam, smem);
    int occ = 0;
    if (h->rowstride == 15) CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_frame_normal<15>, K1_WARPS * 32, smem));
    else CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_frame_normal<21>, K1_WARPS * 32, smem));
    h->k1_grid = std::max(1, std::min(std::max(h->nslots, 1), std::max(occ, 1) * sms));
    // half of the frames in whole static rounds, the rest (and the remainder) from the counter: measured on c4, static
    // fraction 1 / 0.75 / 0.5 / 0.25 / 0: Jacobian pass 0.757 / 0.727 / 0.720 / 0.712 / 0.706 ms, but the row reduction grows
    // with the number of per-frame rows (whole step 6.03 / 5.89 / 5.85 / 5.86 / 5.89 ms)
    double frac = 0.5;
    if (const char* e = std::getenv("MCC_K1_STATIC_FRAC")) frac = std::min(1.0, std::max(0.0, std::atof(e)));
    h->k1_fs = (int)(frac * h->nslots / h->k1_grid) * h->k1_grid;
    h->cc_rows = h->k1_grid + (h->nslots - h->k1_fs);
    if (h->k1_temp_global) CU(h->d_k1_temp.reserve((size_t)h->k1_grid * ((size_t)h->maxft * (h->rowstride / 3) * k1_nacc(h->rowstride) + h->maxft)));
  }

  // upload structure
  auto up = [&](auto& buf, const auto& vec) -> cudaError_t {
    cudaError_t e = buf.reserve(vec.size());
    if (e != cudaSuccess) return e;
    if (vec.empty()) return cudaSuccess;
    h->h2d += vec.size() * sizeof(vec[0]);
    return cudaMemcpyAsync(buf.p, vec.data(), vec.size() * sizeof(vec[0]), cudaMemcpyHostToDevice, h->stream);
  };
  CU(up(h->d_cams, cm));
  CU(up(h->d_cc, cc));
  CU(up(h->d_dist, dist));
  CU(up(h->d_cls_li, cls_li));
  CU(up(h->d_gs_tab, gs_tab));
  CU(h->d_Linv.reserve((size_t)std::max(h->nslots, 1) * 6)); CU(h->d_zf.reserve((size_t)std::max(h->nslots, 1) * 6));
  CU(up(h->d_tiles, tiles));
  CU(up(h->d_gto, gto));
  CU(up(h->d_col_param, h->col_param));
  const size_t ns = std::max(h->nslots, 1);
  CU(h->d_setup.reserve((size_t)std::max(h->ngroups, 1) * SETUP_STRIDE));
  CU(h->d_frot.reserve(ns * K1_FROT)); CU(h->d_camrec.reserve((size_t)std::max(ncam, 1) * K1_CAMREC));
  CU(h->d_tile_s.reserve(std::max(h->ntiles, 1)));
  for (int b = 0; b < 2; b++) { CU(h->d_Hff2[b].reserve(ns * 36)); CU(h->d_gf2[b].reserve(ns * 6)); CU(h->d_Hfc2[b].reserve(ns * 6 * std::max(C, 1))); }
  CU(h->d_Lf.reserve(ns * 36));
  CU(h->d_cost_f.reserve(ns)); CU(h->d_gmax_f.reserve(ns)); CU(h->d_xn2_f.reserve(ns)); CU(h->d_newcost_f.reserve(ns));
  CU(h->d_mcc_f.reserve(ns)); CU(h->d_dx2_f.reserve(ns)); CU(h->d_xn2c_f.reserve(ns));
  CU(h->d_cc_part.reserve((size_t)h->cc_rows * std::max(h->E, 1))); CU(h->d_cc_part2.reserve((size_t)CCR_RS * std::max(h->E, 1)));
  if ((h->E + CCR_E - 1) / CCR_E > 63) return fail(h, MCC_ERR_TOO_LARGE, "too many camera-camera entries (%d)", h->E); CU(h->d_cc_sum.reserve((size_t)h->E + COMM_SCAL + 2 * 64));
  for (int b = 0; b < 2; b++) { CU(h->d_Hcc2[b].reserve((size_t)std::max(C, 1) * std::max(C, 1))); CU(h->d_gc2[b].reserve(std::max(C, 1))); }
  CU(h->d_scale_c.reserve(std::max(C, 1))); CU(h->d_scale_f.reserve(ns * 6));
  CU(h->d_S_part.reserve((size_t)h->sy_gx * h->npairs * SY_T * SY_T)); CU(h->d_S_sum.reserve((size_t)h->npairs * SY_T * SY_T));
  CU(h->d_M.reserve((size_t)(C + 1) * std::max(C, 1))); CU(h->d_yc.reserve(std::max(C, 1))); CU(h->d_ysc.reserve(std::max(C, 1)));
  CU(h->d_scal.reserve(40)); CU(h->d_small.reserve(4 + 64));   // 32 scalars + the failure flags (an int at [32]) so that one copy brings both back
  for (int i = 0; i < 2; i++) { CU(h->d_camp[i].reserve(std::max(h->C_full, 1))); CU(h->d_pose[i].reserve(ns * 6)); }
  CU(cudaStreamSynchronize(h->stream));   // the staging vectors above die with this scope
  h->structure_dirty = false;
  h->params_dirty = true;
  drop_iteration_graphs(h);
  return MCC_OK;
}

mcc_status upload_obs(mcc_handle* h) {
  const size_t n = h->nobs;
  if (h->obs_dirty) {
    CU(h->d_X.reserve(n)); CU(h->d_Y.reserve(n)); CU(h->d_Z.reserve(n)); CU(h->d_U.reserve(n)); CU(h->d_V.reserve(n));
    CU(h->d_aos_w.reserve_keep(3 * n, 3 * h->dev_count, h->stream)); CU(h->d_aos_i.reserve_keep(2 * n, 2 * h->dev_count, h->stream));
    if (n) {
      if (h->dev_count < n) {
        // only what is not on the device yet (frames added from pinned memory went there directly)
        const size_t o = h->dev_count, m = n - o;
        CU(cudaMemcpyAsync(h->d_aos_w.p + 3 * o, h->world.p + 3 * o, 3 * m * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        CU(cudaMemcpyAsync(h->d_aos_i.p + 2 * o, h->image.p + 2 * o, 2 * m * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        h->h2d += 5 * m * sizeof(double);
        h->dev_count = n;
      }
      {
        KScope ks(h, KC_OTHER);
        k_aos_to_soa<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(n, h->d_aos_w.p, h->d_aos_i.p, h->d_X.p, h->d_Y.p, h->d_Z.p, h->d_U.p, h->d_V.p);
      }
      CU(cudaGetLastError());
    }
    h->obs_dirty = false;
  }
  if (h->mask_dirty) {
    CU(h->d_mask.reserve(n));
    if (n) {
      if (h->mask_all_one) CU(cudaMemsetAsync(h->d_mask.p, 1, n, h->stream));
      else { CU(cudaMemcpyAsync(h->d_mask.p, h->mask.data(), n, cudaMemcpyHostToDevice, h->stream)); h->h2d += n; }
    }
    h->mask_dirty = false;
  }
  return MCC_OK;
}

mcc_status upload_params(mcc_handle* h) {
  if (!h->params_dirty) return MCC_OK;
  h->cur = 0;
  if (h->C_full) CU(cudaMemcpyAsync(h->d_camp[0].p, h->params.data(), h->C_full * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  const size_t np = (size_t)h->nslots * 6;
  CU(cudaStreamSynchronize(h->stream));      // nothing may still be reading the staging buffer (cheap: the stream is idle between runs)
  if (!h->pose_stage.reserve(std::max<size_t>(np, 1))) return fail(h, MCC_ERR_BAD_ARG, "out of host memory");
  double* pose = h->pose_stage.p;
  for (int s = 0; s < h->nslots; s++) std::memcpy(&pose[(size_t)s * 6], &h->params[h->C_full + 6 * (size_t)h->slot_frame[s]], 6 * sizeof(double));
  if (h->nslots) CU(cudaMemcpyAsync(h->d_pose[0].p, pose, np * sizeof(double), cudaMemcpyHostToDevice, h->stream));   // pinned: queued, no wait
  h->h2d += (h->C_full + np) * sizeof(double);
  h->params_dirty = false;
  return MCC_OK;
}

mcc_status download_params(mcc_handle* h) {
  if (h->C_full) CU(cudaMemcpyAsync(h->params.data(), h->d_camp[h->cur].p, h->C_full * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  const size_t np = (size_t)h->nslots * 6;
  if (!h->pose_stage.reserve(std::max<size_t>(np, 1))) return fail(h, MCC_ERR_BAD_ARG, "out of host memory");
  double* pose = h->pose_stage.p;
  if (h->nslots) CU(cudaMemcpyAsync(pose, h->d_pose[h->cur].p, np * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  h->d2h += (h->C_full + np) * sizeof(double);
  for (int s = 0; s < h->nslots; s++) std::memcpy(&h->params[h->C_full + 6 * (size_t)h->slot_frame[s]], &pose[(size_t)s * 6], 6 * sizeof(double));
  return MCC_OK;
}

mcc_status ensure_comm_windows(mcc_handle* h);
mcc_status ensure_peer_window(mcc_handle* h);

mcc_status prepare(mcc_handle* h) {
  // several ranks: a rank whose shard is empty still takes part in every collective (with empty launches)
  if (h->cams.empty() || (h->nframes == 0 && h->nranks <= 1)) return fail(h, MCC_ERR_NO_DATA, "no data to run on!");
  ST(ensure_device(h));
  for (const CamHost& c : h->cams)
    if (c.active && c.model != MCC_MODEL_EQUIDISTANT && c.model != MCC_MODEL_RADTAN) return fail(h, MCC_ERR_BAD_MODEL, "unknown distortion model: %d", c.model);
  if (!h->params_valid || (long long)h->params.size() != mcc_num_params(h)) init_params(h);
  if (h->structure_dirty) ST(build_structure(h));
  ST(ensure_peer_window(h));
  if (!h->peer_on) ST(ensure_comm_windows(h));
  ST(upload_obs(h));
  ST(upload_params(h));
  return MCC_OK;
}

// ---- kernel sequences ---------------------------------------------------------------------------
mcc_status launch_setup(mcc_handle* h, int which) {
  if (h->ngroups == 0) return MCC_OK;
  {
    KScope ks(h, KC_SETUP);
    const int fpb = std::min(32, 128 / h->ncam());
    k_group_setup<<<(h->nslots + fpb - 1) / fpb, 128, SETUP_SMEM_BYTES, h->stream>>>(h->nslots, h->ncam(), fpb, h->d_cams.p, h->d_camp[which].p, h->d_pose[which].p, h->d_setup.p, h->st_cur);
  }
  CU(cudaGetLastError());
  return MCC_OK;
}

// K5 at parameter set `which`: per-frame costs into cost_f_out; optional raw residuals
mcc_status launch_cost(mcc_handle* h, int which, int use_loss, double* cost_f_out, double* res_out) {
  ST(launch_setup(h, which));
  if (h->ntiles) {
    {
      KScope ks(h, KC_TILE_COST);
      k_tile_cost<<<(h->ntiles + K1_WARPS - 1) / K1_WARPS, K1_WARPS * 32, 0, h->stream>>>(
        h->ntiles, h->ncam(), h->d_tiles.p, h->d_cams.p, h->d_setup.p, h->d_camp[which].p, h->d_X.p, h->d_Y.p, h->d_Z.p, h->d_U.p, h->d_V.p,
        h->d_mask.p, h->d_tile_s.p, res_out);
    }
    CU(cudaGetLastError());
  }
  if (h->nslots && cost_f_out) {
    {
      KScope ks(h, KC_REDUCE);
      k_frame_cost<<<(h->nslots + 127) / 128, 128, 0, h->stream>>>(h->nslots, h->ncam(), h->d_gto.p, h->d_tile_s.p, use_loss, cost_f_out);
    }
    CU(cudaGetLastError());
  }
  return MCC_OK;
}

inline int* fail_ptr(mcc_handle* h) { return reinterpret_cast<int*>(h->d_scal.p + 32); }
inline double* S_sum_buf(mcc_handle* h) { return h->win_S ? h->win_S : h->d_S_sum.p; }
inline double* cc_sum_buf(mcc_handle* h) { return h->win_cc ? h->win_cc : h->d_cc_sum.p; }

void free_comm_windows(mcc_handle* h) {
  if (h->wh_S) nccl().CommWindowDeregister(h->comm, h->wh_S);
  if (h->wh_cc) nccl().CommWindowDeregister(h->comm, h->wh_cc);
  if (h->win_S) nccl().MemFree(h->win_S);
  if (h->win_cc) nccl().MemFree(h->win_cc);
  h->wh_S = h->wh_cc = nullptr; h->win_S = h->win_cc = nullptr; h->win_S_n = h->win_cc_n = 0;
}

// Collective over the ranks (every rank prepares the same structure): the Schur tile pairs and the merged
// camera-camera / scalar buffer move into ncclMemAlloc'ed memory registered as symmetric windows, which lets NCCL
// pick its low-latency NVLink kernels for these ~50 KB all-reduces.  Only with MCC_NCCL_SYMMETRIC=1.
mcc_status ensure_comm_windows(mcc_handle* h) {
  if (h->nranks <= 1 || !nccl().has_windows()) return MCC_OK;
  const char* env = std::getenv("MCC_NCCL_SYMMETRIC");
  if (!env || env[0] != '1') return MCC_OK;   // opt-in: measured no gain at 2 and 8 GPUs
  const size_t nS = (size_t)h->npairs * SY_T * SY_T, nC = (size_t)h->E + COMM_SCAL + 2 * 64;
  if (h->win_S && h->win_S_n == nS && h->win_cc_n == nC) return MCC_OK;
  CU(cudaStreamSynchronize(h->stream));
  free_comm_windows(h);
  auto round_up = [](size_t b) { return (b + 4095) / 4096 * 4096; };
  void *pS = nullptr, *pC = nullptr;
  if (nccl().MemAlloc(&pS, round_up(nS * sizeof(double))) != 0 || nccl().MemAlloc(&pC, round_up(nC * sizeof(double))) != 0) {
    if (pS) nccl().MemFree(pS);
    return MCC_OK;   // plain buffers
  }
  h->win_S = (double*)pS; h->win_cc = (double*)pC; h->win_S_n = nS; h->win_cc_n = nC;
  const int rS = nccl().CommWindowRegister(h->comm, pS, round_up(nS * sizeof(double)), &h->wh_S, 1 /* NCCL_WIN_COLL_SYMMETRIC */);
  const int rC = nccl().CommWindowRegister(h->comm, pC, round_up(nC * sizeof(double)), &h->wh_cc, 1);
  if (rS != 0) h->wh_S = nullptr;    // the memory is still usable as an ordinary buffer
  if (rC != 0) h->wh_cc = nullptr;
  if (std::getenv("MCC_COMM_DEBUG"))
    std::fprintf(stderr, "[mcc rank %d] symmetric windows: S %zu B -> %s, cc %zu B -> %s\n", h->rank, round_up(nS * sizeof(double)), rS == 0 ? "registered" : "plain",
                 round_up(nC * sizeof(double)), rC == 0 ? "registered" : "plain");
  CU(cudaMemsetAsync(pS, 0, round_up(nS * sizeof(double)), h->stream));
  CU(cudaMemsetAsync(pC, 0, round_up(nC * sizeof(double)), h->stream));
  return MCC_OK;
}

void free_peer_window(mcc_handle* h) {
  if (!h->peer_own) return;
  // No collective here (a handle may be destroyed on one rank only, e.g. on an error path).  It is safe after any completed
  // solve or evaluation: an exchange returns on this rank only once EVERY rank's cells of that exchange have arrived, and
  // a peer stores into this window again only in a later exchange, i.e. in a later call that this rank takes part in.
  cudaStreamSynchronize(h->stream);
  if (!h->member_of)
    for (int r = 0; r < h->pw.nranks; r++)
      if (r != h->pw.rank && h->pw.base[r]) cudaIpcCloseMemHandle(h->pw.base[r]);
  cudaFree(h->peer_own);
  h->peer_own = nullptr; h->peer_on = false; h->peer_nS = h->peer_nC = 0;
  h->pw = PeerWin{};
}

// Collective over the ranks: (re)creates the peer windows for the current sizes of the two exchanges.  Any failure
// (more than PEER_MAX ranks, ranks on different nodes, no peer access, CUDA IPC not permitted) leaves the NCCL path
// in place -- on ALL ranks, the outcome is agreed upon by an all-reduce.  MCC_P2P=0 disables the attempt.
mcc_status ensure_peer_window(mcc_handle* h) {
  if (h->nranks <= 1) return MCC_OK;
  const char* env = std::getenv("MCC_P2P");
  if (((env && env[0] == '0') && !h->member_of) || h->nranks > PEER_MAX) return MCC_OK;   // (members of one handle have no NCCL to fall back to)
  const size_t nS = (size_t)h->npairs * SY_T * SY_T, nC = (size_t)h->E + COMM_SCAL + 2 * 64;
  if ((h->peer_own || h->peer_failed) && h->peer_nS == nS && h->peer_nC == nC) return MCC_OK;
  free_peer_window(h);
  h->peer_failed = false;
  const int R = h->nranks;
  if (h->member_of) {
    // Members of a multi-device handle live in one process: no CUDA IPC, no NCCL.  Every member allocates its window,
    // enables peer access to the other devices and publishes the address; two barriers of the member threads replace the
    // two agreement all-reduces.  All members call this at the same time (the front runs them on a thread each).
    mcc_group* g = h->member_of;
    const size_t cells_ip = 2 * (size_t)R * (nS + nC + PEER_SMALL);
    int bad_ip = 0;
    if (cudaMalloc((void**)&h->peer_own, cells_ip * sizeof(uint4)) != cudaSuccess) { h->peer_own = nullptr; bad_ip = 1; }
    if (!bad_ip && cudaMemsetAsync(h->peer_own, 0, cells_ip * sizeof(uint4), h->stream) != cudaSuccess) bad_ip = 1;
    for (int r = 0; r < R && !bad_ip; r++) {
      if (r == h->rank) continue;
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, h->device, g->kids[r]->device) != cudaSuccess || !can) { bad_ip = 1; break; }
      const cudaError_t pe = cudaDeviceEnablePeerAccess(g->kids[r]->device, 0);
      if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) bad_ip = 1;
    }
    if (cudaStreamSynchronize(h->stream) != cudaSuccess) bad_ip = 1;
    cudaGetLastError();
    { std::lock_guard<std::mutex> lk(g->mu); g->win[h->rank] = h->peer_own; if (h->rank == 0) g->bad = 0; }
    g->barrier();
    { std::lock_guard<std::mutex> lk(g->mu); g->bad |= bad_ip; }
    g->barrier();
    const int any_bad = g->bad;
    PeerWin pwi{};
    pwi.nranks = R; pwi.rank = h->rank;
    for (int r = 0; r < R; r++) pwi.base[r] = (uint4*)g->win[r];
    g->barrier();            // everybody has read the table and the verdict before anybody may start the next round
    h->peer_nS = nS; h->peer_nC = nC;
    if (any_bad) {
      if (h->peer_own) cudaFree(h->peer_own);
      h->peer_own = nullptr; h->pw = PeerWin{}; h->peer_on = false; h->peer_failed = true;
      return fail(h, MCC_ERR_NCCL, "no peer access between the devices of this multi-device handle (device %d)", h->device);
    }
    h->pw = pwi;
    h->peer_seq[0] = h->peer_seq[1] = h->peer_seq[2] = 0; h->peer_on = true;
    double secs = 60.0;
    if (const char* e = std::getenv("MCC_PEER_TIMEOUT_S")) { const double v = std::atof(e); if (v > 0.0) secs = v; }
    h->peer_timeout_cycles = (long long)(secs * 2.0e9);
    return MCC_OK;
  }
  const size_t cells = 2 * (size_t)R * (nS + nC + PEER_SMALL);   // 2 parities x R slots per exchange class, 16 bytes per value
  int bad = 0;
  if (cudaMalloc((void**)&h->peer_own, cells * sizeof(uint4)) != cudaSuccess) { h->peer_own = nullptr; bad = 1; }
  cudaIpcMemHandle_t mine;
  std::memset(&mine, 0, sizeof(mine));
  if (!bad) {
    if (cudaMemsetAsync(h->peer_own, 0, cells * sizeof(uint4), h->stream) != cudaSuccess) bad = 1;
    if (cudaIpcGetMemHandle(&mine, h->peer_own) != cudaSuccess) bad = 1;
  }
  cudaGetLastError();
  // gather the handles: every rank fills its own 64 bytes of a zeroed table, the int32 sum is the concatenation.
  // This is a handshake: a rank that fails locally must still take part in both all-reduces (with bad = 1), or the
  // others would wait for it forever; only a failing NCCL call itself ends it early.
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  const size_t tab_ints = (size_t)R * 16 + 1;   // + the failure count
  std::vector<int> tab(tab_ints, 0);
  std::memcpy(&tab[(size_t)h->rank * 16], &mine, 64);
  CU(h->d_tab.reserve(tab_ints));                // grow-only member buffer: freed with the handle, never leaked
  int* d_tab = h->d_tab.p;
  PeerWin pw{};
  pw.nranks = R; pw.rank = h->rank;
  int one = 1;
  mcc_status hs = MCC_OK;
  auto agree = [&](size_t n) -> bool {           // sum over the ranks of tab[0..n); false = the transport itself failed
    bool local_ok = cudaMemcpyAsync(d_tab, tab.data(), n * sizeof(int), cudaMemcpyHostToDevice, h->stream) == cudaSuccess;
    if (nccl().AllReduce(d_tab, d_tab, n, ncclInt32_, ncclSum_, h->comm, h->stream) != 0) return false;
    local_ok = local_ok && cudaMemcpyAsync(tab.data(), d_tab, n * sizeof(int), cudaMemcpyDeviceToHost, h->stream) == cudaSuccess;
    local_ok = cudaStreamSynchronize(h->stream) == cudaSuccess && local_ok;
    if (!local_ok) { cudaGetLastError(); tab[n - 1] = 1; }   // count as a failure of this rank; the next agreement spreads it
    return true;
  };
  tab[(size_t)R * 16] = bad;
  if (!agree(tab_ints)) hs = fail(h, MCC_ERR_NCCL, "peer-window handshake: ncclAllReduce failed");
  if (hs == MCC_OK) {
    bad = tab[(size_t)R * 16];
    if (!bad) {
      for (int r = 0; r < R; r++) {
        if (r == h->rank) { pw.base[r] = h->peer_own; continue; }
        cudaIpcMemHandle_t hd;
        std::memcpy(&hd, &tab[(size_t)r * 16], 64);
        void* p = nullptr;
        if (cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { bad = 1; cudaGetLastError(); break; }
        pw.base[r] = (uint4*)p;
      }
    }
    // second agreement: did everybody manage to map everybody?
    tab[0] = bad;
    if (!agree(1)) hs = fail(h, MCC_ERR_NCCL, "peer-window handshake: ncclAllReduce failed");
    else one = tab[0];
  }
  h->pw = pw;
  if (std::getenv("MCC_COMM_DEBUG")) std::fprintf(stderr, "[mcc rank %d] peer window %zu B: %s\n", h->rank, cells * sizeof(uint4), one == 0 ? "mapped on all ranks" : "not available, NCCL");
  if (one != 0 || hs != MCC_OK) {
    for (int r = 0; r < R; r++)
      if (r != h->rank && pw.base[r]) cudaIpcCloseMemHandle(pw.base[r]);
    if (h->peer_own) cudaFree(h->peer_own);
    cudaGetLastError();
    h->peer_own = nullptr; h->pw = PeerWin{}; h->peer_on = false;
    // remember the sizes so that the attempt is not repeated on every run
    h->peer_nS = nS; h->peer_nC = nC; h->peer_failed = true;
    return hs;
  }
  {
    // how long an exchange waits for a peer before it gives up (SM clocks; clock64 runs at the SM clock, <= ~2 GHz)
    double secs = 60.0;
    if (const char* e = std::getenv("MCC_PEER_TIMEOUT_S")) { const double v = std::atof(e); if (v > 0.0) secs = v; }
    h->peer_timeout_cycles = (long long)(secs * 2.0e9);
  }
  h->peer_nS = nS; h->peer_nC = nC; h->peer_seq[0] = h->peer_seq[1] = h->peer_seq[2] = 0; h->peer_on = true;
  return MCC_OK;
}

// exchange 0: Schur tile pairs (n = peer_nS), exchange 1: camera-camera entries + scalars (n <= peer_nC).
// The descriptor of the next exchange of class ex; queued iterations (st_cur set) take the sequence number from the device
// state, which the host mirrors from the decision records.
PeerEx peer_ex(mcc_handle* h, int ex) {
  PeerEx px;
  px.pw = h->pw; px.ex = ex;
  px.slot_n = ex == 0 ? h->peer_nS : (ex == 1 ? h->peer_nC : (size_t)PEER_SMALL);
  px.ex_base = ex == 0 ? 0 : 2 * (size_t)h->nranks * (ex == 1 ? h->peer_nS : h->peer_nS + h->peer_nC);
  px.seq = (h->st_cur && ex < 2) ? h->peer_seq[ex] + 1 : ++h->peer_seq[ex];
  px.timeout = h->peer_timeout_cycles;
  return px;
}
mcc_status peer_allreduce(mcc_handle* h, int ex, double* p, size_t n) {
  const PeerEx px = peer_ex(h, ex);
  if (n > px.slot_n) return fail(h, MCC_ERR_NCCL, "peer window laid out for another size");
  {
    KScope ks(h, KC_EXCHANGE);   // includes the wait for the slowest rank
    CU(launch_k(h, k_peer_allreduce, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, px, p, (int)n, fail_ptr(h), h->st_cur));
  }
  CU(cudaGetLastError());
  return MCC_OK;
}

mcc_status allreduce(mcc_handle* h, double* p, size_t n, int op) {
  if (h->nranks <= 1 || n == 0) return MCC_OK;
  KScope ks(h, KC_EXCHANGE);
  if (!h->comm) return fail(h, MCC_ERR_NCCL, "no NCCL communicator and no peer windows between the devices");
  NC(nccl().AllReduce(p, p, n, ncclFloat64_, op, h->comm, h->stream));
  return MCC_OK;
}
// sum over the ranks of a few doubles (<= PEER_SMALL), outside queued iterations: peer windows if there are any, else NCCL
mcc_status allreduce_small(mcc_handle* h, double* p, size_t n) {
  if (h->nranks <= 1 || n == 0) return MCC_OK;
  if (h->peer_on) return peer_allreduce(h, 2, p, n);
  return allreduce(h, p, n, ncclSum_);
}

// K1 at the current parameters: blocks, cost, gradient norm, |x|^2 -> d_scal[0..2], [8..9]
// blocks of parameter set `which` into block set `bo`
// Several ranks: the camera-camera entries and every scalar of the iteration travel in ONE all-reduce (sum) over
// d_cc_sum = [entries E | 6 scalars | gradient max per rank | failure flag per rank]; k_scatter_cc unpacks it.
// extra: further per-frame arrays to reduce in the same launch (launch_step's scalars of a speculative step).
mcc_status launch_jacobian(mcc_handle* h, int use_loss, int which, int bo, const ReduceJob* extra = nullptr, int nextra = 0, bool rot_ready = false) {
  const bool multi = h->nranks > 1;
  double* tail = cc_sum_buf(h) + h->E;
  if (h->nslots) {
    K1Args ka;
    ka.nslots = h->nslots; ka.ncam = h->ncam(); ka.C = h->C; ka.maxft = h->maxft; ka.E = h->E; ka.use_loss = use_loss; ka.ncls = h->ncls; ka.gstride = h->gstride;
    ka.tiles = h->d_tiles.p; ka.gto = h->d_gto.p; ka.cams = h->d_cams.p; ka.cls_lane_info = h->d_cls_li.p; ka.dist = h->d_dist.p; ka.gs_tab = h->d_gs_tab.p;
    ka.camrec = h->d_camrec.p; ka.frot = h->d_frot.p; ka.pose = h->d_pose[which].p;
    ka.oX = h->d_X.p; ka.oY = h->d_Y.p; ka.oZ = h->d_Z.p; ka.oU = h->d_U.p; ka.oV = h->d_V.p; ka.omask = h->d_mask.p;
    ka.Hff = h->d_Hff2[bo].p; ka.gf = h->d_gf2[bo].p; ka.Hfc = h->d_Hfc2[bo].p; ka.cost_f = h->d_cost_f.p; ka.gmax_f = h->d_gmax_f.p; ka.xn2_f = h->d_xn2_f.p;
    ka.cc_part = h->d_cc_part.p; ka.st = h->st_cur; ka.fs = h->k1_fs; ka.ticket = h->d_ticket;
    const size_t smem = k1_smem_bytes(h->rowstride, h->ncam(), h->k1_temp_global ? 0 : h->maxft);
    ka.temp_g = h->k1_temp_global ? h->d_k1_temp.p : nullptr;
    if (!rot_ready) {     // inside an LM iteration k_solve_cam and k_backsub have already made the records of the candidate
      KScope ks(h, KC_SETUP);
      CU(launch_k(h, k_rot_setup, dim3((h->nslots + h->ncam() + 127) / 128), dim3(128), 0, h->nslots, h->ncam(), h->d_cams.p, h->d_camp[which].p, h->d_pose[which].p, h->d_frot.p,
                                                                              h->d_camrec.p, h->st_cur));
    }
    KScope ks(h, KC_TILE_NORMAL);
    if (h->rowstride == 15) CU(launch_k(h, k_frame_normal<15>, dim3(h->k1_grid), dim3(K1_WARPS * 32), smem, ka));
    else CU(launch_k(h, k_frame_normal<21>, dim3(h->k1_grid), dim3(K1_WARPS * 32), smem, ka));
  }
  CU(cudaGetLastError());
  // the tail of the pass in one launch (k_reduce_misc): row and array reductions, with several ranks the exchange of the
  // camera-camera entries and scalars through the peer windows, dense H_cc / g_c and the decision of a queued iteration.
  // Without peer windows the exchange is an NCCL all-reduce between two launches.
  static const bool fused_env = !(std::getenv("MCC_FUSED_TAIL") && std::getenv("MCC_FUSED_TAIL")[0] == '0');
  const bool fused = fused_env && (!multi || h->peer_on);
  if (multi && !fused) CU(cudaMemsetAsync(cc_sum_buf(h), 0, ((size_t)h->E + COMM_SCAL + 2 * 64) * sizeof(double), h->stream));   // the other ranks' slots
  TailArgs ta;
  std::memset(&ta, 0, sizeof(ta));
  ta.jobs.j[0] = ReduceJob{h->d_cost_f.p, 0, multi ? tail + 0 : nullptr};
  ta.jobs.j[1] = ReduceJob{h->d_xn2_f.p, 0, multi ? tail + 1 : nullptr};
  ta.jobs.j[2] = ReduceJob{h->d_gmax_f.p, 1, multi ? tail + COMM_SCAL + h->rank : nullptr};
  for (int i = 0; i < nextra && i < 5; i++) ta.jobs.j[3 + i] = extra[i];
  ta.ncc = (h->E + CCR_E - 1) / CCR_E * CCR_RS;
  ta.E = h->E; ta.nrows = h->nslots ? h->cc_rows : 0; ta.n = h->nslots; ta.tail = fused; ta.has_flag_job = nextra > 0;
  ta.part = h->d_cc_part.p; ta.part2 = h->d_cc_part2.p; ta.cc_sum = cc_sum_buf(h); ta.counters = h->d_ticket + 1; ta.out = h->d_scal.p;
  ta.rank = h->rank;
  ta.px.pw.nranks = 1;
  if (multi && fused) {
    if ((size_t)h->E + COMM_SCAL + 2 * h->nranks > h->peer_nC) return fail(h, MCC_ERR_NCCL, "peer window laid out for another size");
    ta.px = peer_ex(h, 1);
  }
  ScatterArgs& sc = ta.sc;
  sc.C = h->C; sc.E = h->E; sc.nranks = h->nranks; sc.cc = h->d_cc.p; sc.cc_sum = cc_sum_buf(h); sc.col_param = h->d_col_param.p; sc.camp = h->d_camp[which].p;
  sc.Hcc = h->d_Hcc2[bo].p; sc.gc = h->d_gc2[bo].p; sc.scal = h->d_scal.p + 8; sc.scal_all = h->d_scal.p; sc.fail = fail_ptr(h); sc.da = h->decide;
  {
    KScope ks(h, KC_REDUCE);
    CU(launch_k(h, k_reduce_misc, dim3(ta.ncc + 3 + nextra), dim3(1024), 0, ta, h->st_cur));
  }
  if (!fused) {
    if (multi) ST(allreduce(h, cc_sum_buf(h), (size_t)h->E + COMM_SCAL + 2 * h->nranks, ncclSum_));
    KScope ks(h, KC_REDUCE);
    CU(launch_k(h, k_scatter_cc, dim3(1), dim3(1024), 0, sc, h->st_cur));
  }
  CU(cudaGetLastError());
  return MCC_OK;
}

// failure flags of an iteration (d_scal[32]): 1 pose block not positive definite, 2 reduced system not positive definite,
// 4 non-finite step -- all three mean "invalid step" to the trust-region loop -- and 16: a peer exchange gave up waiting
// for another rank, which is not a property of the step but a broken job: reported as an error, peer windows are dropped.
mcc_status check_flags(mcc_handle* h, int flags) {
  if (flags & 16) {
    h->peer_on = false; h->peer_failed = true;
    return fail(h, MCC_ERR_NCCL, "peer exchange timed out: rank %d waited more than %.0f s for another rank's part of an all-reduce (did every rank enter "
                "the same mcc_* call? MCC_PEER_TIMEOUT_S raises the limit, MCC_P2P=0 uses NCCL, which waits forever)", h->rank, (double)h->peer_timeout_cycles / 2.0e9);
  }
  return MCC_OK;
}

mcc_status read_scalars(mcc_handle* h) {
  CU(cudaMemcpyAsync(h->h_scal, h->d_scal.p, 33 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));   // scalars + failure flags
  CU(cudaStreamSynchronize(h->stream));
  h->d2h += 33 * sizeof(double);
  if (h->ktiming) resolve_ktiming(h);
  return check_flags(h, *reinterpret_cast<const int*>(h->h_scal + 32));
}

// K2..K5 for one trust-region radius: candidate parameters in the other buffer, scalars in d_scal[3..6], [10..12]
// K2..K4 for one trust-region radius, then the evaluation of the candidate: either the cheap residual-only pass
// (speculate = false) or straight away the full Jacobian pass into the OTHER block set (speculate = true), whose
// cost doubles as the candidate cost; an accepted step then needs no further pass.
mcc_status launch_step(mcc_handle* h, const mcc_options& o, double radius, bool speculate) {
  const int cur = h->cur, cand = 1 - cur;
  const int C = h->C;
  if (!h->st_cur) CU(cudaMemsetAsync(fail_ptr(h), 0, sizeof(int), h->stream));   // queued iterations: decide_body clears the flags
  {
    CU(cudaGetLastError());
    auto schur_smem = [&](int fb, bool pf) {
      size_t b = ((size_t)fb * 6 * h->LD + 2 * fb * (36 + 18)) * sizeof(double);
      if (pf) b += (size_t)fb * 6 * C * sizeof(double);
      return std::max(b, (size_t)SY_THREADS * (SY_T * SY_T + 1) * sizeof(double));
    };
    dim3 grid(h->sy_gx, h->sy_gy);
    {
      KScope ks(h, KC_SCHUR);
#define MCC_LAUNCH_SCHUR(PF, FB)                                                                                                                   \
      CU((launch_k(h, k_schur_syrk<PF, FB>, grid, dim3(SY_THREADS), schur_smem(FB, PF), h->nslots, C, h->LD, h->npairs, h->d_Hfc2[h->blk].p, h->d_scale_f.p,     \
                                                                                h->d_scale_c.p, h->d_Lf.p, h->d_Linv.p, h->d_zf.p, h->d_S_part.p,                                \
                                                                                radius, o.min_lm_diagonal, o.max_lm_diagonal, h->d_Hff2[h->blk].p, h->d_gf2[h->blk].p, fail_ptr(h), h->st_cur)))
      // prefetch variants keep two CTAs per SM (<= 112 KB each): 8 frames per batch, else 4; wider systems read Hfc from global memory
      if (schur_smem(8, true) <= 112 * 1024) MCC_LAUNCH_SCHUR(true, 8);
      else if (schur_smem(4, true) <= 112 * 1024) MCC_LAUNCH_SCHUR(true, 4);
      else MCC_LAUNCH_SCHUR(false, 8);
#undef MCC_LAUNCH_SCHUR
    }
    CU(cudaGetLastError());
    const int n = h->npairs * SY_T * SY_T;
    {
      // with peer windows the sum over the ranks (exchange 0) happens inside the row reduction
      PeerEx px;
      std::memset(&px, 0, sizeof(px));
      px.pw.nranks = 1;
      const bool peer = h->nranks > 1 && h->peer_on;
      if (peer) {
        if ((size_t)n > h->peer_nS) return fail(h, MCC_ERR_NCCL, "peer window laid out for another size");
        px = peer_ex(h, 0);
      }
      KScope ks(h, peer ? KC_EXCHANGE : KC_REDUCE);
      CU(launch_k(h, k_reduce_rows, dim3((n + RR_E - 1) / RR_E), dim3(RR_E * RR_S), 0, h->sy_gx, n, h->d_S_part.p, S_sum_buf(h), px, fail_ptr(h), h->st_cur));
    }
    CU(cudaGetLastError());
    if (h->nranks > 1 && !h->peer_on) ST(allreduce(h, S_sum_buf(h), n, ncclSum_));
  }
  {
    const size_t need = ((size_t)C * (C + 1) / 2 + C + 1) * sizeof(double);   // packed lower triangle + rhs row
    const bool packed = need <= 184 * 1024;   // + ~42 KB of static shared memory in k_solve_cam
    {
      KScope ks(h, KC_SOLVE);
      if (packed)
        CU(launch_k(h, k_solve_cam<true>, dim3(1), dim3(CH_THREADS), need, C, radius, o.min_lm_diagonal, o.max_lm_diagonal, h->d_Hcc2[h->blk].p, h->d_gc2[h->blk].p, S_sum_buf(h),
                                                              h->d_scale_c.p, h->d_col_param.p, h->C_full, h->d_camp[cur].p, h->d_camp[cand].p, h->d_yc.p,
                                                              h->d_ysc.p, h->d_M.p, h->d_scal.p + 10, fail_ptr(h), h->d_cams.p, h->ncam(), h->d_camrec.p, h->st_cur));
      else
        CU(launch_k(h, k_solve_cam<false>, dim3(1), dim3(CH_THREADS), 0, C, radius, o.min_lm_diagonal, o.max_lm_diagonal, h->d_Hcc2[h->blk].p, h->d_gc2[h->blk].p, S_sum_buf(h),
                                                            h->d_scale_c.p, h->d_col_param.p, h->C_full, h->d_camp[cur].p, h->d_camp[cand].p, h->d_yc.p,
                                                            h->d_ysc.p, h->d_M.p, h->d_scal.p + 10, fail_ptr(h), h->d_cams.p, h->ncam(), h->d_camrec.p, h->st_cur));
    }
    CU(cudaGetLastError());
  }
  if (h->nslots) {
    {
      KScope ks(h, KC_BACKSUB);
      CU(launch_k(h, k_backsub, dim3((h->nslots + 3) / 4), dim3(128), 0, h->nslots, C, h->d_Hff2[h->blk].p, h->d_gf2[h->blk].p, h->d_Hfc2[h->blk].p, h->d_Lf.p, h->d_Linv.p, h->d_scale_f.p, h->d_ysc.p,
                                                         h->d_pose[cur].p, h->d_pose[cand].p, h->d_mcc_f.p, h->d_dx2_f.p, h->d_xn2c_f.p, fail_ptr(h), h->d_frot.p, h->st_cur));
    }
    CU(cudaGetLastError());
  }
  ReduceJobs jobs;
  std::memset(&jobs, 0, sizeof(jobs));
  if (speculate) {
    // the step's scalars are reduced in the same launch as the Jacobian pass's; with several ranks they (and the
    // failure flag) ride in the tail of the buffer that pass all-reduces: two collectives per iteration, this and K2's
    const bool multi = h->nranks > 1;
    double* tail = cc_sum_buf(h) + h->E;
    ReduceJob extra[4];
    extra[0] = ReduceJob{h->d_mcc_f.p, 0, multi ? tail + 3 : h->d_scal.p + 4};
    extra[1] = ReduceJob{h->d_dx2_f.p, 0, multi ? tail + 4 : h->d_scal.p + 5};
    extra[2] = ReduceJob{h->d_xn2c_f.p, 0, multi ? tail + 5 : h->d_scal.p + 6};
    extra[3] = ReduceJob{reinterpret_cast<const double*>(fail_ptr(h)), 2, tail + COMM_SCAL + h->nranks + h->rank};
    ST(launch_jacobian(h, o.use_loss, cand, 1 - h->blk, extra, multi ? 4 : 3, true));   // cost, |x|^2, gradient norm -> d_scal[0..2], [8..9]
    return MCC_OK;
  }
  ST(launch_cost(h, cand, o.use_loss, h->d_newcost_f.p, nullptr));
  // several ranks: [candidate cost, model cost change, |dx|^2, |x+|^2 | failure flag per rank] summed over the ranks in ONE
  // small exchange (peer windows or NCCL), then unpacked (flags OR-ed)
  const bool multi = h->nranks > 1;
  double* sm4 = multi ? h->d_small.p : nullptr;
  if (multi) CU(cudaMemsetAsync(h->d_small.p, 0, (4 + 64) * sizeof(double), h->stream));
  jobs.j[0] = ReduceJob{h->d_newcost_f.p, 0, multi ? sm4 + 0 : nullptr};
  jobs.j[1] = ReduceJob{h->d_mcc_f.p, 0, multi ? sm4 + 1 : nullptr};
  jobs.j[2] = ReduceJob{h->d_dx2_f.p, 0, multi ? sm4 + 2 : nullptr};
  jobs.j[3] = ReduceJob{h->d_xn2c_f.p, 0, multi ? sm4 + 3 : nullptr};
  jobs.j[4] = ReduceJob{reinterpret_cast<const double*>(fail_ptr(h)), 2, multi ? sm4 + 4 + h->rank : nullptr};
  {
    KScope ks(h, KC_REDUCE);
    k_reduce_arrays<<<multi ? 5 : 4, 1024, 0, h->stream>>>(jobs, h->nslots, h->d_scal.p + 3);
  }
  CU(cudaGetLastError());
  if (multi) {
    ST(allreduce_small(h, sm4, 4 + h->nranks));
    KScope ks(h, KC_REDUCE);
    k_unpack_small<<<1, 32, 0, h->stream>>>(h->nranks, sm4, h->d_scal.p + 3, fail_ptr(h));
    CU(cudaGetLastError());
  }
  return MCC_OK;
}

// the LM loop
mcc_status solve(mcc_handle* h, const mcc_options& o, mcc_summary& S) {
  const double t_start = now_s();
  double t_iter = t_start;
  auto log_row = [&](double cost, double cc, double gm, double sn, double ratio, double radius) {
    const double t = now_s();
    if (S.trace_len < MCC_MAX_TRACE) {
      const int i = S.trace_len++;
      S.tr_cost[i] = cost; S.tr_cost_change[i] = cc; S.tr_gradient_max[i] = gm; S.tr_step_norm[i] = sn; S.tr_ratio[i] = ratio; S.tr_radius[i] = radius;
      S.tr_iter_time[i] = t - t_iter;
    }
    if (o.verbose) {
      if (S.num_iterations == 0) std::printf("iter      cost      cost_change  |gradient|   |step|    tr_ratio  tr_radius  ls_iter  iter_time  total_time\n");
      std::printf("%4d  %.6e    %.2e    %.2e   %.2e   %.2e  %.2e        %d    %.2e    %.2e\n", S.num_iterations, cost, cc, gm, sn, ratio, radius,
                  S.num_iterations ? 1 : 0, t - t_iter, t - t_start);
    }
    t_iter = t;
    S.num_iterations++;
  };
  double t0 = now_s();
  ST(launch_jacobian(h, o.use_loss, h->cur, h->blk));
  {
    KScope ks(h, KC_OTHER);
    k_jacobi_scale<<<(std::max(h->C, h->nslots * 6) + 127) / 128, 128, 0, h->stream>>>(h->nslots, h->C, o.jacobi_scaling, h->d_Hcc2[h->blk].p, h->d_Hff2[h->blk].p,
                                                                                       h->d_scale_c.p, h->d_scale_f.p);
  }
  CU(cudaGetLastError());
  ST(read_scalars(h));
  S.t_jacobian += now_s() - t0; S.n_jacobian_evals++;
  double cost = h->h_scal[0];
  double x_norm = std::sqrt(h->h_scal[1] + h->h_scal[9]);
  double gmax = std::max(h->h_scal[2], h->h_scal[8]);
  S.initial_cost = cost;
  double radius = o.initial_trust_region_radius, decrease_factor = 2.0;
  log_row(cost, 0, gmax, 0, 0, radius);
  S.num_successful = 1;
  auto finish = [&](int term, int why, double val) {
    S.termination = term; S.why = why; S.why_value = val; S.final_cost = cost;
    return MCC_OK;
  };
  if (gmax <= o.gradient_tolerance) return finish(0, 0, gmax);
  // Candidate evaluation is speculative while steps are being accepted: the full Jacobian pass at x+ goes into the
  // spare block set and its cost is the candidate cost, so an accepted iteration is ONE pass over the observations
  // and ONE host synchronisation.  After a rejected or invalid step the next candidate gets the cheap cost-only
  // pass first (Ceres' order), and the Jacobian pass follows only on acceptance.
  bool speculate = true;
  int invalid_run = 0, iter = 0;
  // Pipelined section: while steps are being accepted, iteration i+1 is queued before the host has seen the decision
  // of iteration i.  decide_body (k_scatter_cc) takes that decision on the device (same tests as below); if it is anything but
  // "accepted, go on" it sets a stop flag that turns the already queued iteration into no-ops, and the host picks
  // up from the scalars of the deciding iteration with the ordinary code below.  The device keeps the radius, so the
  // GPU never waits for the host between accepted iterations (MCC_PIPELINE=0: always the ordinary path).
  const char* pipe_env = std::getenv("MCC_PIPELINE");
  const bool pipeline = !(pipe_env && pipe_env[0] == '0') && !h->ktiming;
  const char* graph_env = std::getenv("MCC_GRAPHS");
  const bool use_graphs = pipeline && (h->nranks == 1 || h->peer_on) && !(graph_env && graph_env[0] == '0');
  if (use_graphs) {
    const double key[9] = {o.function_tolerance, o.parameter_tolerance, o.gradient_tolerance, o.max_trust_region_radius, o.min_relative_decrease,
                           o.min_lm_diagonal, o.max_lm_diagonal, (double)o.use_loss, 1.0};
    if (std::memcmp(h->it_graph_key, key, sizeof(key)) != 0) { drop_iteration_graphs(h); std::memcpy(h->it_graph_key, key, sizeof(key)); }
  }
  while (true) {
    if (iter >= o.max_num_iterations) return finish(1, 3, iter);
    bool have_scalars = false;
    t0 = now_s();
    if (speculate && pipeline) {
      LmState st0;
      st0.radius = radius; st0.cost = cost; st0.x_norm = x_norm; st0.gmax = gmax; st0.stop = 0; st0.pad = 0;
      st0.seq[0] = h->peer_seq[0]; st0.seq[1] = h->peer_seq[1];
      CU(cudaMemcpyAsync(h->d_state, &st0, sizeof(st0), cudaMemcpyHostToDevice, h->stream));
      CU(cudaMemsetAsync(fail_ptr(h), 0, sizeof(int), h->stream));
      int ecur = h->cur, eblk = h->blk, queued = 0, decided = 0;
      auto enqueue = [&]() -> mcc_status {
        const int slot = queued & 1;
        auto body = [&]() -> mcc_status {
          const int ccur = h->cur, cblk = h->blk;
          h->cur = ecur; h->blk = eblk; h->st_cur = h->d_state;
          h->decide = DecideArgs{h->d_state, h->d_rec + slot * LM_REC, fail_ptr(h), o.parameter_tolerance, o.function_tolerance, o.gradient_tolerance,
                                 o.min_relative_decrease, o.max_trust_region_radius};
          mcc_status r = launch_step(h, o, radius, true);
          h->cur = ccur; h->blk = cblk; h->st_cur = nullptr; h->decide = DecideArgs{};
          ST(r);
          return MCC_OK;
        };
        if (use_graphs) {
          const int gi = (ecur << 2) | (eblk << 1) | slot;
          cudaGraphExec_t& ge = h->it_graph[gi];
          if (!ge) {
            const long long l0 = h->launches;
            CU(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
            const mcc_status r = body();
            h->it_graph_kernels[gi] = h->launches - l0;   // captured, not launched
            h->launches = l0;
            cudaGraph_t g = nullptr;
            const cudaError_t ce = cudaStreamEndCapture(h->stream, &g);
            if (r != MCC_OK) { if (g) cudaGraphDestroy(g); return r; }
            CU(ce);
            const cudaError_t ie = cudaGraphInstantiate(&ge, g, 0);
            cudaGraphDestroy(g);
            CU(ie);
          }
          CU(cudaGraphLaunch(ge, h->stream));
          h->launches += h->it_graph_kernels[gi];
        } else {
          ST(body());
        }
        CU(cudaEventRecord(h->ev_rec[slot], h->stream));
        h->d2h += LM_REC * sizeof(double);
        ecur ^= 1; eblk ^= 1; queued++;
        return MCC_OK;
      };
      while (true) {
        while (queued - decided < 2 && iter + (queued - decided) < o.max_num_iterations) ST(enqueue());
        if (queued == decided) break;   // iteration limit
        const int slot = decided & 1;
        CU(cudaEventSynchronize(h->ev_rec[slot]));
        const double* rec = h->h_rec + slot * LM_REC;
        decided++;
        const int outcome = (int)rec[33];
        if (outcome != LM_SKIPPED) { h->peer_seq[0]++; h->peer_seq[1]++; }   // this iteration ran both exchanges
        if (outcome != LM_ACCEPT) {
          // the iteration that stopped the pipeline: hand its scalars to the ordinary code (what is queued behind it does nothing)
          CU(cudaStreamSynchronize(h->stream));
          std::memcpy(h->h_scal, rec, 33 * sizeof(double));
          ST(check_flags(h, *reinterpret_cast<const int*>(h->h_scal + 32)));
          have_scalars = true;
          break;
        }
        iter++;
        S.t_linear += now_s() - t0; t0 = now_s();
        S.n_linear_solves++; S.n_jacobian_evals++;
        invalid_run = 0;
        h->cur = 1 - h->cur; h->blk = 1 - h->blk;
        x_norm = rec[38]; radius = rec[34]; cost = rec[39]; gmax = rec[40];
        decrease_factor = 2.0;
        S.num_successful++;
        log_row(cost, rec[36], gmax, rec[37], rec[35], radius);
      }
      if (!have_scalars) continue;      // iteration limit reached inside the pipeline: the test at the top ends the solve
    }
    iter++;
    const bool spec = speculate;
    if (!have_scalars) {
      ST(launch_step(h, o, radius, spec));
      ST(read_scalars(h));
    }
    S.t_linear += now_s() - t0; S.n_linear_solves++;
    if (spec) S.n_jacobian_evals++; else S.n_residual_evals++;
    const int flags = *reinterpret_cast<const int*>(h->h_scal + 32);
    const double new_cost = spec ? h->h_scal[0] : h->h_scal[3];
    const double mcc = h->h_scal[4] + h->h_scal[10];
    const double step_norm = std::sqrt(h->h_scal[5] + h->h_scal[11]);
    const double xn_cand = std::sqrt(h->h_scal[6] + h->h_scal[12]);
    bool step_valid = flags == 0 && std::isfinite(mcc) && mcc > 0.0;
    if (!step_valid) {
      invalid_run++;
      if (invalid_run >= o.max_num_consecutive_invalid_steps) return finish(2, 4, invalid_run);
      radius /= decrease_factor; decrease_factor *= 2.0;
      S.num_unsuccessful++;
      speculate = false;
      log_row(cost, 0, gmax, 0, 0, radius);
      if (radius < o.min_trust_region_radius) return finish(0, 5, radius);
      continue;
    }
    invalid_run = 0;
    if (step_norm <= o.parameter_tolerance * (x_norm + o.parameter_tolerance)) return finish(0, 1, step_norm / (x_norm + o.parameter_tolerance));
    const double cost_change = cost - new_cost;
    if (std::fabs(cost_change) <= o.function_tolerance * cost) return finish(0, 2, std::fabs(cost_change) / cost);
    const double ratio = cost_change / mcc;
    if (ratio > o.min_relative_decrease) {
      h->cur = 1 - h->cur;
      x_norm = xn_cand;
      radius = radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * ratio - 1.0, 3));
      radius = std::min(o.max_trust_region_radius, radius);
      decrease_factor = 2.0;
      if (spec) {
        h->blk = 1 - h->blk;                 // the candidate's blocks are already there
        cost = new_cost;
        gmax = std::max(h->h_scal[2], h->h_scal[8]);
      } else {
        t0 = now_s();
        ST(launch_jacobian(h, o.use_loss, h->cur, h->blk));
        ST(read_scalars(h));
        S.t_jacobian += now_s() - t0; S.n_jacobian_evals++;
        cost = h->h_scal[0];
        gmax = std::max(h->h_scal[2], h->h_scal[8]);
      }
      speculate = true;
      S.num_successful++;
      log_row(cost, cost_change, gmax, step_norm, ratio, radius);
      if (gmax <= o.gradient_tolerance) return finish(0, 0, gmax);
    } else {
      radius /= decrease_factor; decrease_factor *= 2.0;
      S.num_unsuccessful++;
      speculate = false;
      log_row(cost, cost_change, 0.0, step_norm, ratio, radius);
    }
    if (radius < o.min_trust_region_radius) return finish(0, 5, radius);
  }
}

}  // namespace


// =================================================================================================
// Several devices behind one handle (mcc_set_devices).  The front handle keeps what the reference's Calibrator keeps
// about the problem as a whole (cameras, point counts, pose guesses, the parameter vector in the reference's layout) and
// deals the frames out to one member handle per device; every compute call runs the members on one host thread each.
// The members are ordinary handles with nranks = number of devices: they shard the solve exactly like the
// one-process-per-GPU mode, but find each other's peer windows inside the process (ensure_peer_window).
// =================================================================================================
namespace {

template <typename F>
mcc_status group_each(mcc_handle* h, F fn) {
  std::vector<mcc_handle*>& K = h->group->kids;
  std::vector<mcc_status> st(K.size(), MCC_OK);
  std::vector<std::thread> th;
  for (size_t r = 0; r < K.size(); r++) th.emplace_back([&, r] { st[r] = fn(K[r], (int)r); });
  for (std::thread& t : th) t.join();
  for (size_t r = 0; r < K.size(); r++)
    if (st[r] != MCC_OK) return fail(h, st[r], "device %d: %s", K[r]->device, K[r]->err.c_str());
  return MCC_OK;
}

mcc_status group_push_cameras(mcc_handle* h) {
  mcc_group* g = h->group;
  if (!g->cams_dirty) return MCC_OK;
  std::vector<mcc_camera> cams(h->cams.size());
  for (size_t k = 0; k < h->cams.size(); k++) {
    const CamHost& c = h->cams[k];
    mcc_camera& m = cams[k];
    m.model = c.model; m.n_intr = c.n_intr; m.n_dist = c.n_dist; m.intr = c.intr; m.dist = c.dist;
    std::memcpy(m.T_cn_cnm1, c.T, sizeof(m.T_cn_cnm1));
    m.fix_intr = c.fix_intr; m.fix_extr = c.fix_extr; m.active = c.active;
  }
  for (mcc_handle* k : g->kids) {
    mcc_status st = mcc_set_cameras(k, (int)cams.size(), cams.data());
    if (st == MCC_OK) st = mcc_set_fix_all_intrinsics(k, h->fix_all ? 1 : 0);
    if (st != MCC_OK) return fail(h, st, "device %d: %s", k->device, k->err.c_str());
  }
  g->cams_dirty = false;
  return MCC_OK;
}

// the front's parameter vector (reference layout, all frames) -> every member's vector (its frames only)
mcc_status group_push_params(mcc_handle* h) {
  mcc_group* g = h->group;
  if (g->params_pushed || !h->params_valid) return MCC_OK;
  const size_t ncp = h->params.size() - 6 * (size_t)h->nframes;
  for (size_t r = 0; r < g->kids.size(); r++) {
    std::vector<double> p(h->params.begin(), h->params.begin() + ncp);
    for (int f : g->kid_frames[r]) p.insert(p.end(), h->params.begin() + ncp + 6 * (size_t)f, h->params.begin() + ncp + 6 * (size_t)f + 6);
    const mcc_status st = mcc_set_params(g->kids[r], p.data(), p.size());
    if (st != MCC_OK) return fail(h, st, "device %d: %s", g->kids[r]->device, g->kids[r]->err.c_str());
  }
  g->params_pushed = true;
  return MCC_OK;
}

// every member's vector -> the front's; the camera part must be the same bits on every member
mcc_status group_pull_params(mcc_handle* h) {
  mcc_group* g = h->group;
  const long long np = mcc_num_params(h);
  const size_t ncp = (size_t)np - 6 * (size_t)h->nframes;
  h->params.assign((size_t)np, 0.0);
  for (size_t r = 0; r < g->kids.size(); r++) {
    const std::vector<double>& p = g->kids[r]->params;
    if (p.size() != ncp + 6 * g->kid_frames[r].size()) return fail(h, MCC_ERR_NUMERICAL, "device %d holds %zu parameters, expected %zu", g->kids[r]->device, p.size(), ncp + 6 * g->kid_frames[r].size());
    if (r == 0) std::memcpy(h->params.data(), p.data(), ncp * sizeof(double));
    else if (std::memcmp(h->params.data(), p.data(), ncp * sizeof(double)) != 0) return fail(h, MCC_ERR_NUMERICAL, "the devices disagree on the camera parameters (device %d)", g->kids[r]->device);
    for (size_t i = 0; i < g->kid_frames[r].size(); i++) std::memcpy(&h->params[ncp + 6 * (size_t)g->kid_frames[r][i]], &p[ncp + 6 * i], 6 * sizeof(double));
  }
  h->params_valid = true;
  g->params_pushed = true;
  return MCC_OK;
}

mcc_status group_add_frames(mcc_handle* h, int nframes, int ncam, const int* npts, const double* world_xyz, const double* image_uv, const double* poses) {
  mcc_group* g = h->group;
  ST(group_push_cameras(h));
  const int n = (int)g->kids.size();
  std::vector<long long> fobs(nframes), off(nframes + 1, 0);
  for (int f = 0; f < nframes; f++) {
    long long c = 0;
    for (int k = 0; k < ncam; k++) { if (npts[(size_t)f * ncam + k] < 0) return fail(h, MCC_ERR_BAD_ARG, "negative point count"); c += npts[(size_t)f * ncam + k]; }
    fobs[f] = c; off[f + 1] = off[f] + c;
  }
  if (off[nframes] && (!world_xyz || !image_uv)) return fail(h, MCC_ERR_BAD_ARG, "null point arrays");
  // who gets which frames of this batch: contiguous pieces that level the members' observation counts; a batch
  // smaller than the member count goes frame by frame to whoever holds the least
  std::vector<int> piece_lo(n, 0), piece_hi(n, 0);
  std::vector<int> single_owner;
  if (nframes >= n) {
    long long total = off[nframes];
    for (long long v : g->kid_obs) total += v;
    int f = 0;
    long long have_all = 0, want_all = 0;
    for (int r = 0; r < n; r++) {
      piece_lo[r] = f;
      want_all += (total + n - 1 - r) / n;                     // cumulative target after member r
      have_all += g->kid_obs[r];
      while (f < nframes && (r == n - 1 || have_all + fobs[f] / 2 < want_all)) { have_all += fobs[f]; f++; }
      piece_hi[r] = f;
    }
  } else {
    std::vector<long long> load = g->kid_obs;
    for (int f = 0; f < nframes; f++) {
      int best = 0;
      for (int r = 1; r < n; r++) if (load[r] < load[best]) best = r;
      single_owner.push_back(best);
      load[best] += fobs[f];
    }
  }
  auto add_piece = [&](int r, int lo, int hi) -> mcc_status {
    if (hi <= lo) return MCC_OK;
    return mcc_add_frames(g->kids[r], hi - lo, ncam, npts + (size_t)lo * ncam, world_xyz ? world_xyz + 3 * off[lo] : nullptr,
                          image_uv ? image_uv + 2 * off[lo] : nullptr, poses + 16 * (size_t)lo);
  };
  if (nframes >= n) {
    // one host thread per member: page-locked caller arrays go to the devices over their own PCIe links at the same time
    ST(group_each(h, [&](mcc_handle*, int r) { return add_piece(r, piece_lo[r], piece_hi[r]); }));
  } else {
    for (int f = 0; f < nframes; f++) {
      const mcc_status st = add_piece(single_owner[f], f, f + 1);
      if (st != MCC_OK) return fail(h, st, "device %d: %s", g->kids[single_owner[f]]->device, g->kids[single_owner[f]]->err.c_str());
    }
  }
  // bookkeeping of the front
  const int f0 = h->nframes;
  g->frame_kid.resize(f0 + nframes); g->frame_local.resize(f0 + nframes);
  auto own = [&](int f, int r) {
    g->frame_kid[f0 + f] = r; g->frame_local[f0 + f] = (int)g->kid_frames[r].size();
    g->kid_frames[r].push_back(f0 + f); g->kid_obs[r] += fobs[f];
  };
  if (nframes >= n) { for (int r = 0; r < n; r++) for (int f = piece_lo[r]; f < piece_hi[r]; f++) own(f, r); }
  else for (int f = 0; f < nframes; f++) own(f, single_owner[f]);
  h->npts.insert(h->npts.end(), npts, npts + (size_t)nframes * ncam);
  h->poses.insert(h->poses.end(), poses, poses + 16 * (size_t)nframes);
  h->nobs += (size_t)off[nframes];
  h->nframes += nframes;
  h->init_valid = false;
  h->params_valid = false;
  g->params_pushed = false;
  return MCC_OK;
}

mcc_status group_run(mcc_handle* h, const mcc_options& o, mcc_summary& S) {
  mcc_group* g = h->group;
  const double t0 = now_s();
  ST(group_push_cameras(h));
  if (o.keep_params) {
    if (!h->params_valid) init_params(h);
    ST(group_push_params(h));
  }
  if (o.verbose) std::printf("Num params: %lld\nNum residuals: %lld\n", mcc_num_params(h), mcc_num_residuals(h));
  std::vector<mcc_summary> sums(g->kids.size());
  ST(group_each(h, [&](mcc_handle* k, int r) {
    mcc_options ok = o;
    if (r != 0) ok.verbose = 0;
    return mcc_run(k, &ok, &sums[r]);
  }));
  ST(group_pull_params(h));
  S = sums[0];
  const int Cc = mcc_num_reduced_cam_cols(h);
  S.num_params = (int)h->params.size();
  S.num_residuals = 0; S.num_reduced_params = Cc; S.n_kernel_launches = 0; S.h2d_bytes = 0; S.d2h_bytes = 0;
  for (const mcc_summary& s : sums) {
    S.num_residuals += s.num_residuals; S.num_reduced_params += s.num_reduced_params - Cc;
    S.n_kernel_launches += s.n_kernel_launches; S.h2d_bytes += s.h2d_bytes; S.d2h_bytes += s.d2h_bytes;
  }
  S.t_total = now_s() - t0;
  return MCC_OK;
}

// per frame of the front: residual doubles it contributes (valid points of active cameras x 2), from the members' masks
mcc_status group_frame_residual_counts(mcc_handle* h, std::vector<long long>& cnt) {
  mcc_group* g = h->group;
  const int ncam = h->ncam();
  cnt.assign(h->nframes, 0);
  for (size_t r = 0; r < g->kids.size(); r++) {
    mcc_handle* k = g->kids[r];
    std::vector<uint8_t> m(k->nobs + 1);
    if (mcc_get_mask(k, m.data(), k->nobs) != MCC_OK) return fail(h, MCC_ERR_BAD_ARG, "mask of device %d", k->device);
    size_t o = 0;
    for (int f : g->kid_frames[r])
      for (int c = 0; c < ncam; c++) {
        const int np = h->npts[(size_t)f * ncam + c];
        if (h->cams[c].active) for (int i = 0; i < np; i++) cnt[f] += 2 * (m[o + i] != 0);
        o += np;
      }
  }
  return MCC_OK;
}

mcc_status group_eval_residuals(mcc_handle* h, double* out_xy, size_t n_out) {
  mcc_group* g = h->group;
  ST(group_push_cameras(h));
  ST(group_push_params(h));
  std::vector<std::vector<double>> buf(g->kids.size());
  ST(group_each(h, [&](mcc_handle* k, int r) {
    const long long n = mcc_num_residuals(k);
    buf[r].assign((size_t)std::max<long long>(n, 1), 0.0);
    return mcc_eval_residuals(k, buf[r].data(), (size_t)n);
  }));
  std::vector<long long> cnt;
  ST(group_frame_residual_counts(h, cnt));
  long long total = 0;
  for (long long c : cnt) total += c;
  if ((long long)n_out < total) return fail(h, MCC_ERR_BAD_ARG, "output too small: need %lld doubles", total);
  std::vector<size_t> cur(g->kids.size(), 0);
  size_t w = 0;
  for (int f = 0; f < h->nframes; f++) {          // frame order of the front; a member's stream is in ITS frame order, which is ascending too
    const int r = g->frame_kid[f];
    std::memcpy(out_xy + w, buf[r].data() + cur[r], (size_t)cnt[f] * sizeof(double));
    w += (size_t)cnt[f]; cur[r] += (size_t)cnt[f];
  }
  return MCC_OK;
}

// mask of the front's observation order <-> the members' (dir 0: gather into out, 1: scatter from in)
mcc_status group_mask(mcc_handle* h, uint8_t* out, const uint8_t* in) {
  mcc_group* g = h->group;
  const int ncam = h->ncam();
  std::vector<size_t> fo(h->nframes + 1, 0);
  for (int f = 0; f < h->nframes; f++) { size_t c = 0; for (int k = 0; k < ncam; k++) c += h->npts[(size_t)f * ncam + k]; fo[f + 1] = fo[f] + c; }
  for (size_t r = 0; r < g->kids.size(); r++) {
    mcc_handle* k = g->kids[r];
    std::vector<uint8_t> m(k->nobs + 1);
    if (out && mcc_get_mask(k, m.data(), k->nobs) != MCC_OK) return fail(h, MCC_ERR_BAD_ARG, "mask of device %d", k->device);
    size_t o = 0;
    for (int f : g->kid_frames[r]) {
      const size_t n = fo[f + 1] - fo[f];
      if (out) std::memcpy(out + fo[f], m.data() + o, n); else std::memcpy(m.data() + o, in + fo[f], n);
      o += n;
    }
    if (in && k->nobs && mcc_set_mask(k, m.data(), k->nobs) != MCC_OK) return fail(h, MCC_ERR_BAD_ARG, "mask of device %d", k->device);
  }
  return MCC_OK;
}

}  // namespace

// =================================================================================================
// C ABI
// =================================================================================================
extern "C" {

const char* mcc_version(void) { return "multicam_calibration_b200 0.1 (sm_100a, fp64)"; }

void mcc_default_options(mcc_options* o) {
  o->max_num_iterations = 200;          // calibrator.cpp:302
  o->function_tolerance = 1e-12;        // calibrator.cpp:304
  o->parameter_tolerance = 1e-12;       // calibrator.cpp:305
  o->gradient_tolerance = 1e-10;
  o->initial_trust_region_radius = 1e4;
  o->max_trust_region_radius = 1e16;
  o->min_trust_region_radius = 1e-32;
  o->min_relative_decrease = 1e-3;
  o->min_lm_diagonal = 1e-6;
  o->max_lm_diagonal = 1e32;
  o->max_num_consecutive_invalid_steps = 5;
  o->jacobi_scaling = 1;
  o->use_loss = 1;
  o->verbose = 0;
  o->keep_params = 0;
}
int mcc_summary_size(void) { return (int)sizeof(mcc_summary); }

mcc_status mcc_create(mcc_handle** out) {
  if (!out) return MCC_ERR_BAD_ARG;
  *out = new (std::nothrow) mcc_handle();
  if (!*out) return MCC_ERR_BAD_ARG;
  return MCC_OK;
}

void mcc_destroy(mcc_handle* h) {
  if (!h) return;
  if (h->group) {
    for (mcc_handle* k : h->group->kids) mcc_destroy(k);
    delete h->group;
    h->world.release(); h->image.release(); h->pose_stage.release();
    delete h;
    return;
  }
  if (h->device >= 0) cudaSetDevice(h->device);
  free_peer_window(h);
  if (h->comm && nccl().has_windows()) free_comm_windows(h);
  if (h->comm && nccl().CommDestroy) nccl().CommDestroy(h->comm);
  DBuf<double>* dd[] = {&h->d_aos_w, &h->d_aos_i, &h->d_X, &h->d_Y, &h->d_Z, &h->d_U, &h->d_V, &h->d_res, &h->d_camp[0], &h->d_camp[1], &h->d_pose[0], &h->d_pose[1],
                        &h->d_setup, &h->d_frot, &h->d_camrec, &h->d_tile_s, &h->d_Hff2[0], &h->d_Hff2[1], &h->d_gf2[0], &h->d_gf2[1], &h->d_Hfc2[0], &h->d_Hfc2[1], &h->d_Lf, &h->d_cost_f, &h->d_gmax_f, &h->d_xn2_f, &h->d_newcost_f,
                        &h->d_mcc_f, &h->d_dx2_f, &h->d_xn2c_f, &h->d_cc_part, &h->d_cc_part2, &h->d_k1_temp, &h->d_cc_sum, &h->d_Hcc2[0], &h->d_Hcc2[1], &h->d_gc2[0], &h->d_gc2[1], &h->d_scale_c, &h->d_scale_f, &h->d_S_part,
                        &h->d_S_sum, &h->d_M, &h->d_yc, &h->d_ysc, &h->d_scal, &h->d_small};
  for (auto* b : dd) b->release();
  h->d_mask.release(); h->d_tiles.release(); h->d_gto.release(); h->d_tab.release(); h->d_col_param.release(); h->d_dist.release(); h->d_cls_li.release(); h->d_gs_tab.release();
  h->d_cams.release(); h->d_cc.release(); h->d_Linv.release(); h->d_zf.release();
  h->world.release(); h->image.release(); h->pose_stage.release();
  if (h->h_scal) cudaFreeHost(h->h_scal);
  drop_iteration_graphs(h);
  if (h->d_state) { cudaFree(h->d_state); cudaFree(h->d_ticket); cudaFreeHost(h->h_rec); for (int i = 0; i < 2; i++) cudaEventDestroy(h->ev_rec[i]); }
  for (cudaEvent_t e : h->ev_pool) cudaEventDestroy(e);
  for (KPending& pe : h->ev_pending) { cudaEventDestroy(pe.a); cudaEventDestroy(pe.b); }
  if (h->tm_a) { cudaEventDestroy(h->tm_a); cudaEventDestroy(h->tm_b); }
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

const char* mcc_last_error(const mcc_handle* h) { return h ? h->err.c_str() : "null handle"; }

mcc_status mcc_set_devices(mcc_handle* h, int ndev, const int* ids) {
  if (!h || ndev < 1 || !ids) return fail(h, MCC_ERR_BAD_ARG, "bad device list");
  if (h->group) return fail(h, MCC_ERR_BAD_ARG, "devices already set for this handle");
  if (h->stream || h->nframes) return fail(h, MCC_ERR_BAD_ARG, "set the devices before frames are added and before the first computation");
  if (ndev == 1) return mcc_set_device(h, ids[0]);
  if (ndev > PEER_MAX) return fail(h, MCC_ERR_BAD_ARG, "at most %d devices per handle", PEER_MAX);
  int have = 0;
  if (cudaGetDeviceCount(&have) != cudaSuccess || have == 0) { cudaGetLastError(); return fail(h, MCC_ERR_NO_DEVICE, "no CUDA device available; this library has no CPU fallback"); }
  for (int i = 0; i < ndev; i++) {
    if (ids[i] < 0 || ids[i] >= have) return fail(h, MCC_ERR_BAD_ARG, "device %d does not exist (%d devices)", ids[i], have);
    for (int j = 0; j < i; j++) if (ids[j] == ids[i]) return fail(h, MCC_ERR_BAD_ARG, "device %d listed twice", ids[i]);
  }
  mcc_group* g = new (std::nothrow) mcc_group();
  if (!g) return MCC_ERR_BAD_ARG;
  for (int r = 0; r < ndev; r++) {
    mcc_handle* k = nullptr;
    if (mcc_create(&k) != MCC_OK) { for (mcc_handle* q : g->kids) mcc_destroy(q); delete g; return fail(h, MCC_ERR_BAD_ARG, "out of memory"); }
    k->device = ids[r]; k->nranks = ndev; k->rank = r; k->member_of = g;
    g->kids.push_back(k);
  }
  g->kid_frames.resize(ndev); g->kid_obs.assign(ndev, 0); g->win.assign(ndev, nullptr);
  h->group = g;
  return MCC_OK;
}

mcc_status mcc_set_device(mcc_handle* h, int device) {
  if (!h) return MCC_ERR_BAD_ARG;
  if (h->group) return fail(h, MCC_ERR_BAD_ARG, "devices already set for this handle");
  if (h->stream) return fail(h, MCC_ERR_BAD_ARG, "device already in use by this handle");
  h->device = device;
  return MCC_OK;
}

mcc_status mcc_set_cameras(mcc_handle* h, int ncam, const mcc_camera* cams) {
  if (!h || ncam < 0 || ncam > MCC_MAX_CAMERAS || (ncam && !cams)) return fail(h, MCC_ERR_BAD_ARG, "bad camera array");
  if (h->nframes && ncam != h->ncam()) return fail(h, MCC_ERR_BAD_ARG, "camera count changed while frames are held");
  std::vector<CamHost> v(ncam);
  for (int k = 0; k < ncam; k++) {
    const mcc_camera& c = cams[k];
    if (c.n_intr != 4) return fail(h, MCC_ERR_BAD_ARG, "camera %d: n_intr must be 4", k);
    if (c.n_dist < 0 || c.n_dist > MCC_MAX_DIST) return fail(h, MCC_ERR_BAD_ARG, "camera %d: n_dist must be 0..%d", k, MCC_MAX_DIST);
    if (c.model != MCC_MODEL_EQUIDISTANT && c.model != MCC_MODEL_RADTAN) return fail(h, MCC_ERR_BAD_MODEL, "unknown distortion model: %d", c.model);
    CamHost& d = v[k];
    d.model = c.model; d.n_intr = 4; d.n_dist = c.n_dist;
    for (int i = 0; i < 4; i++) d.intr[i] = c.intr[i];
    for (int i = 0; i < c.n_dist; i++) d.dist[i] = c.dist[i];
    std::memcpy(d.T, c.T_cn_cnm1, sizeof(d.T));
    d.fix_intr = c.fix_intr; d.fix_extr = c.fix_extr; d.active = c.active;
  }
  // like the reference (calibrator.cpp:225-227) this replaces calibrationData_ only: params_ of a finished run
  // stay valid for testCalibration / removeOutliers as long as the parameter layout is unchanged
  bool same_layout = v.size() == h->cams.size();
  for (size_t k = 0; same_layout && k < v.size(); k++) same_layout = v[k].n_dist == h->cams[k].n_dist;
  h->cams.swap(v);
  h->init_valid = false;
  if (h->group) { h->group->cams_dirty = true; h->group->params_pushed = false; }
  h->structure_dirty = true;
  if (!same_layout) h->params_valid = false;
  return MCC_OK;
}

mcc_status mcc_set_fix_all_intrinsics(mcc_handle* h, int fix) {
  if (!h) return MCC_ERR_BAD_ARG;
  h->fix_all = fix != 0;
  h->structure_dirty = true;
  if (h->group) h->group->cams_dirty = true;
  return MCC_OK;
}

mcc_status mcc_set_camera_flags(mcc_handle* h, int cam, int fix_intr, int fix_extr, int active) {
  if (!h || cam < 0 || cam >= h->ncam()) return fail(h, MCC_ERR_BAD_ARG, "bad camera index");
  h->cams[cam].fix_intr = fix_intr; h->cams[cam].fix_extr = fix_extr; h->cams[cam].active = active;
  h->structure_dirty = true;
  if (h->group)
    for (mcc_handle* k : h->group->kids)
      if (!h->group->cams_dirty) ST(mcc_set_camera_flags(k, cam, fix_intr, fix_extr, active));
  return MCC_OK;
}

mcc_status mcc_add_frames(mcc_handle* h, int nframes, int ncam, const int* npts, const double* world_xyz, const double* image_uv, const double* poses) {
  if (!h) return MCC_ERR_BAD_ARG;
  if (ncam != h->ncam() || ncam == 0) {
    std::fprintf(stderr, "number of cams != width of array!\n");   // calibrator.cpp:238
    return fail(h, MCC_ERR_BAD_ARG, "number of cams != width of array!");
  }
  if (nframes < 0 || !npts || !poses) return fail(h, MCC_ERR_BAD_ARG, "null argument");
  if (h->group) return group_add_frames(h, nframes, ncam, npts, world_xyz, image_uv, poses);
  size_t n = 0;
  for (size_t i = 0; i < (size_t)nframes * ncam; i++) { if (npts[i] < 0) return fail(h, MCC_ERR_BAD_ARG, "negative point count"); n += npts[i]; }
  if (n && (!world_xyz || !image_uv)) return fail(h, MCC_ERR_BAD_ARG, "null point arrays");
  // Where do the points go?  If the caller's arrays are page-locked (cudaHostAlloc / cudaHostRegister) and the
  // device copy is in step, they are DMA'd straight into the device AoS buffers -- no host staging pass.  Otherwise
  // they are copied into the handle's own pinned staging buffers and uploaded by the next run.  Either way the
  // caller's memory is no longer referenced when this returns.
  bool direct = false;
  if (n >= 4096 && h->dev_count == h->nobs) {
    cudaPointerAttributes aw, ai;
    const bool pw = cudaPointerGetAttributes(&aw, world_xyz) == cudaSuccess && aw.type == cudaMemoryTypeHost;
    const bool pi = cudaPointerGetAttributes(&ai, image_uv) == cudaSuccess && ai.type == cudaMemoryTypeHost;
    cudaGetLastError();
    direct = pw && pi && ensure_device(h) == MCC_OK;
  }
  if (direct) {
    const size_t o = h->nobs;
    CU(h->d_aos_w.reserve_keep(3 * (o + n), 3 * o, h->stream));
    CU(h->d_aos_i.reserve_keep(2 * (o + n), 2 * o, h->stream));
    CU(cudaMemcpyAsync(h->d_aos_w.p + 3 * o, world_xyz, 3 * n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CU(cudaMemcpyAsync(h->d_aos_i.p + 2 * o, image_uv, 2 * n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    h->h2d += 5 * n * sizeof(double);
    h->dev_count = o + n;
  } else {
    if (!h->world.reserve(3 * (h->nobs + n), 3 * h->host_count) || !h->image.reserve(2 * (h->nobs + n), 2 * h->host_count)) return fail(h, MCC_ERR_BAD_ARG, "out of host memory");
    if (h->host_count < h->nobs) {
      // earlier frames went to the device only: bring them back so that the staging buffers are whole again
      const size_t o = h->host_count, m = h->nobs - o;
      CU(cudaMemcpyAsync(h->world.p + 3 * o, h->d_aos_w.p + 3 * o, 3 * m * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
      CU(cudaMemcpyAsync(h->image.p + 2 * o, h->d_aos_i.p + 2 * o, 2 * m * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
      CU(cudaStreamSynchronize(h->stream));
      h->d2h += 5 * m * sizeof(double);
    }
    parallel_memcpy(h->world.p + 3 * h->nobs, world_xyz, 3 * n * sizeof(double));
    parallel_memcpy(h->image.p + 2 * h->nobs, image_uv, 2 * n * sizeof(double));
    h->host_count = h->nobs + n;
  }
  h->npts.insert(h->npts.end(), npts, npts + (size_t)nframes * ncam);
  h->poses.insert(h->poses.end(), poses, poses + 16 * (size_t)nframes);
  if (!h->mask_all_one) h->mask.resize(h->nobs + n, 1);
  h->nobs += n;
  h->nframes += nframes;
  h->init_valid = false;
  h->structure_dirty = h->obs_dirty = h->mask_dirty = true;
  h->params_valid = false;
  if (direct) {
    // The copy engine is still busy with the caller's arrays.  For the first bulk load of a data set, use the time:
    // build the tile structure on the host and queue the AoS -> SoA pass behind the copies, so that the next run
    // finds both done.  An error here is not this call's to report: the run rebuilds and reports it.
    if (h->nframes == nframes && nframes >= 64) {
      if (build_structure(h) == MCC_OK) {
        if (upload_obs(h) != MCC_OK) h->obs_dirty = h->mask_dirty = true;
      } else {
        h->structure_dirty = true;
      }
    }
    CU(cudaStreamSynchronize(h->stream));   // from here on the caller's arrays are no longer referenced
  }
  return MCC_OK;
}

mcc_status mcc_add_frame(mcc_handle* h, int ncam, const int* npts, const double* const* world_xyz, const double* const* image_uv, const double* T_c0_w) {
  if (!h) return MCC_ERR_BAD_ARG;
  if (ncam != h->ncam() || ncam == 0) {
    std::fprintf(stderr, "number of cams != width of array!\n");
    return fail(h, MCC_ERR_BAD_ARG, "number of cams != width of array!");
  }
  if (!npts || !T_c0_w) return fail(h, MCC_ERR_BAD_ARG, "null argument");
  std::vector<double> w, im;
  for (int k = 0; k < ncam; k++) {
    if (npts[k] < 0 || (npts[k] && (!world_xyz || !image_uv || !world_xyz[k] || !image_uv[k]))) return fail(h, MCC_ERR_BAD_ARG, "bad point arrays");
    if (npts[k]) { w.insert(w.end(), world_xyz[k], world_xyz[k] + 3 * (size_t)npts[k]); im.insert(im.end(), image_uv[k], image_uv[k] + 2 * (size_t)npts[k]); }
  }
  return mcc_add_frames(h, 1, ncam, npts, w.data(), im.data(), T_c0_w);
}

mcc_status mcc_clear_frames(mcc_handle* h) {
  if (!h) return MCC_ERR_BAD_ARG;
  if (h->group) {
    mcc_group* g = h->group;
    for (mcc_handle* k : g->kids) mcc_clear_frames(k);
    g->frame_kid.clear(); g->frame_local.clear();
    for (auto& v : g->kid_frames) v.clear();
    std::fill(g->kid_obs.begin(), g->kid_obs.end(), 0);
    g->params_pushed = false;
  }
  h->npts.clear(); h->poses.clear(); h->mask.clear(); h->mask_all_one = true;
  h->nobs = 0; h->nframes = 0; h->host_count = 0; h->dev_count = 0;
  h->init_valid = false;
  h->structure_dirty = h->obs_dirty = h->mask_dirty = true;
  h->params_valid = false;
  return MCC_OK;
}

long long mcc_num_frames(const mcc_handle* h) { return h ? h->nframes : 0; }
long long mcc_num_observations(const mcc_handle* h) { return h ? (long long)h->nobs : 0; }

long long mcc_num_params(const mcc_handle* h) {
  if (!h) return 0;
  long long n = 0;
  for (const CamHost& c : h->cams) n += c.nparam();
  n += 6 * std::max(0, h->ncam() - 1);
  return n + 6ll * h->nframes;
}

mcc_status mcc_init_params(mcc_handle* h) {
  if (!h) return MCC_ERR_BAD_ARG;
  if (h->group) h->group->params_pushed = false;
  init_params(h);
  return MCC_OK;
}

mcc_status mcc_get_params(const mcc_handle* h, double* out, size_t n) {
  if (!h || !out) return MCC_ERR_BAD_ARG;
  if (!h->params_valid) return fail(const_cast<mcc_handle*>(h), MCC_ERR_NO_DATA, "parameters not initialised");
  if (n < h->params.size()) return fail(const_cast<mcc_handle*>(h), MCC_ERR_BAD_ARG, "output too small: need %zu", h->params.size());
  std::memcpy(out, h->params.data(), h->params.size() * sizeof(double));
  return MCC_OK;
}

mcc_status mcc_set_params(mcc_handle* h, const double* in, size_t n) {
  if (!h || !in) return MCC_ERR_BAD_ARG;
  if ((long long)n != mcc_num_params(h)) return fail(h, MCC_ERR_BAD_ARG, "expected %lld parameters, got %zu", mcc_num_params(h), n);
  h->params.assign(in, in + n);
  h->params_valid = true;
  h->params_dirty = true;
  if (h->group) h->group->params_pushed = false;
  return MCC_OK;
}

mcc_status mcc_upload(mcc_handle* h) {
  if (!h) return MCC_ERR_BAD_ARG;
  if (h->group) {
    if (h->cams.empty() || h->nframes == 0) return fail(h, MCC_ERR_NO_DATA, "no data to run on!");
    ST(group_push_cameras(h));
    ST(group_push_params(h));
    return group_each(h, [](mcc_handle* k, int) { return mcc_upload(k); });
  }
  ST(prepare(h));
  CU(cudaStreamSynchronize(h->stream));
  return MCC_OK;
}

mcc_status mcc_run(mcc_handle* h, const mcc_options* opts, mcc_summary* out) {
  if (!h) return MCC_ERR_BAD_ARG;
  mcc_options o;
  mcc_default_options(&o);
  if (opts) o = *opts;
  mcc_summary local;
  mcc_summary& S = out ? *out : local;
  std::memset(&S, 0, sizeof(S));
  if (h->cams.empty() || (h->nframes == 0 && h->nranks <= 1)) {
    std::fprintf(stderr, "no data to run on!\n");   // calibrator.cpp:287
    S.termination = 2; S.why = 7;
    return fail(h, MCC_ERR_NO_DATA, "no data to run on!");
  }
  if (h->group) return group_run(h, o, S);
  const double t0 = now_s();
  const long long l0 = h->launches, u0 = h->h2d, d0 = h->d2h;
  if (!o.keep_params) init_params(h);
  ST(prepare(h));
  S.t_upload = now_s() - t0;
  S.num_params = (int)h->params.size();
  S.num_reduced_params = h->C + 6 * h->nslots;
  S.num_residuals = h->nres;
  if (o.verbose && !h->member_of) std::printf("Num params: %d\nNum residuals: %lld\n", S.num_params, S.num_residuals);   // (a multi-device front prints the totals)
  if (h->nslots == 0 && h->nranks <= 1) { S.termination = 2; S.why = 7; return fail(h, MCC_ERR_NO_DATA, "no valid points in any active camera"); }
  mcc_status st = solve(h, o, S);
  if (st != MCC_OK) return st;
  ST(download_params(h));
  S.t_total = now_s() - t0;
  S.n_kernel_launches = h->launches - l0;
  S.h2d_bytes = h->h2d - u0;
  S.d2h_bytes = h->d2h - d0;
  return MCC_OK;
}

mcc_status mcc_get_camera(const mcc_handle* hc, int k, double* intr, double* dist, double* T_cn_cnm1) {
  mcc_handle* h = const_cast<mcc_handle*>(hc);
  if (!h || k < 0 || k >= h->ncam()) return fail(h, MCC_ERR_BAD_ARG, "bad camera index");
  if (!h->params_valid) return fail(h, MCC_ERR_NO_DATA, "parameters not initialised");
  int off = 0, num_intr = 0;
  for (int i = 0; i < h->ncam(); i++) { if (i < k) off += h->cams[i].nparam(); num_intr += h->cams[i].nparam(); }
  const CamHost& c = h->cams[k];
  if (intr) for (int i = 0; i < 4; i++) intr[i] = h->params[off + i];
  if (dist) for (int i = 0; i < c.n_dist; i++) dist[i] = h->params[off + 4 + i];
  if (T_cn_cnm1) {
    double Tk[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}, Tp[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    if (k > 0) {
      h_rvec_tvec_to_transform(&h->params[num_intr + 6 * (k - 1)], Tk);
      if (k > 1) h_rvec_tvec_to_transform(&h->params[num_intr + 6 * (k - 2)], Tp);
      double inv[16];
      h_mat4_inv(Tp, inv);
      h_mat4_mul(Tk, inv, T_cn_cnm1);       // calibrator.cpp:441
    } else {
      std::memcpy(T_cn_cnm1, c.T, sizeof(Tk));
    }
  }
  return MCC_OK;
}

long long mcc_num_residuals(const mcc_handle* hc) {
  mcc_handle* h = const_cast<mcc_handle*>(hc);
  if (!h) return 0;
  if (h->group) { long long t = 0; for (mcc_handle* k : h->group->kids) t += mcc_num_residuals(k); return t; }
  long long n = 0;
  size_t o = 0;
  for (int f = 0; f < h->nframes; f++)
    for (int k = 0; k < h->ncam(); k++) {
      const int np = h->npts[(size_t)f * h->ncam() + k];
      if (h->cams[k].active) { if (h->mask_all_one) n += np; else for (int i = 0; i < np; i++) n += h->mask[o + i] != 0; }
      o += np;
    }
  return 2 * n;
}

mcc_status mcc_eval_residuals(mcc_handle* h, double* out_xy, size_t n_out) {
  if (!h || !out_xy) return MCC_ERR_BAD_ARG;
  if (h->group) return group_eval_residuals(h, out_xy, n_out);
  ST(prepare(h));
  if ((long long)n_out < h->nres) return fail(h, MCC_ERR_BAD_ARG, "output too small: need %lld doubles", h->nres);
  CU(h->d_res.reserve(2 * std::max<size_t>(h->nobs, 1)));
  ST(launch_cost(h, h->cur, 0, nullptr, h->d_res.p));
  std::vector<double> all(2 * h->nobs);
  CU(cudaMemcpyAsync(all.data(), h->d_res.p, all.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  h->d2h += all.size() * sizeof(double);
  // compact: frames with residual blocks -> active cameras -> valid points (calibrator.cpp:336-347)
  size_t w = 0;
  const int ncam = h->ncam();
  for (int s = 0; s < h->nslots; s++) {
    const int f = h->slot_frame[s];
    for (int k = 0; k < ncam; k++) {
      if (!h->cams[k].active) continue;
      const size_t o0 = h->group_obs0[(size_t)f * ncam + k];
      const int np = h->npts[(size_t)f * ncam + k];
      for (int i = 0; i < np; i++)
        if (h->mask_all_one || h->mask[o0 + i]) { out_xy[w++] = all[2 * (o0 + i)]; out_xy[w++] = all[2 * (o0 + i) + 1]; }
    }
  }
  return MCC_OK;
}

mcc_status mcc_eval_cost(mcc_handle* h, int use_loss, double* cost) {
  if (!h || !cost) return MCC_ERR_BAD_ARG;
  if (h->group) {
    ST(group_push_cameras(h));
    ST(group_push_params(h));
    std::vector<double> c(h->group->kids.size(), 0.0);
    ST(group_each(h, [&](mcc_handle* k, int r) { return mcc_eval_cost(k, use_loss, &c[r]); }));   // every member returns the sum over all
    *cost = c[0];
    return MCC_OK;
  }
  ST(prepare(h));
  ST(launch_cost(h, h->cur, use_loss, h->d_newcost_f.p, nullptr));
  ReduceJobs jobs;
  std::memset(&jobs, 0, sizeof(jobs));
  jobs.j[0] = ReduceJob{h->d_newcost_f.p, 0};
  {
    KScope ks(h, KC_REDUCE);
    k_reduce_arrays<<<1, 1024, 0, h->stream>>>(jobs, h->nslots, h->d_scal.p + 3);
  }
  CU(cudaGetLastError());
  ST(allreduce_small(h, h->d_scal.p + 3, 1));
  ST(read_scalars(h));
  *cost = h->h_scal[3];
  return MCC_OK;
}

mcc_status mcc_remove_outliers(mcc_handle* h, double max_err, long long* ndropped) {
  if (!h) return MCC_ERR_BAD_ARG;
  if (h->group) {
    ST(group_push_cameras(h));
    ST(group_push_params(h));
    std::vector<long long> d(h->group->kids.size(), 0);
    ST(group_each(h, [&](mcc_handle* k, int r) { return mcc_remove_outliers(k, max_err, &d[r]); }));
    long long t = 0;
    for (long long v : d) t += v;
    if (ndropped) *ndropped = t;
    return MCC_OK;
  }
  ST(prepare(h));
  CU(h->d_res.reserve(2 * std::max<size_t>(h->nobs, 1)));
  ST(launch_cost(h, h->cur, 0, nullptr, h->d_res.p));
  std::vector<double> all(2 * h->nobs);
  CU(cudaMemcpyAsync(all.data(), h->d_res.p, all.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  h->d2h += all.size() * sizeof(double);
  const double m2 = max_err * max_err;
  long long dropped = 0;
  materialize_mask(h);
  const int ncam = h->ncam();
  for (int s = 0; s < h->nslots; s++) {
    const int f = h->slot_frame[s];
    for (int k = 0; k < ncam; k++) {
      if (!h->cams[k].active) continue;
      const size_t o0 = h->group_obs0[(size_t)f * ncam + k];
      const int np = h->npts[(size_t)f * ncam + k];
      for (int i = 0; i < np; i++) {
        if (!h->mask[o0 + i]) continue;
        const double dx = all[2 * (o0 + i)], dy = all[2 * (o0 + i) + 1];
        if (dx * dx + dy * dy > m2) { h->mask[o0 + i] = 0; dropped++; }
      }
    }
  }
  if (dropped) { h->mask_dirty = true; h->structure_dirty = true; h->mask_all_one = false; }
  if (ndropped) *ndropped = dropped;
  return MCC_OK;
}

mcc_status mcc_get_mask(const mcc_handle* h, uint8_t* out, size_t n) {
  if (!h || !out || n < h->nobs) return MCC_ERR_BAD_ARG;
  if (h->group) return group_mask(const_cast<mcc_handle*>(h), out, nullptr);
  if (h->mask_all_one) std::memset(out, 1, h->nobs);
  else std::memcpy(out, h->mask.data(), h->nobs);
  return MCC_OK;
}
mcc_status mcc_set_mask(mcc_handle* h, const uint8_t* in, size_t n) {
  if (!h || !in || n != h->nobs) return MCC_ERR_BAD_ARG;
  if (h->group) return group_mask(h, nullptr, in);
  materialize_mask(h);
  std::memcpy(h->mask.data(), in, n);
  h->mask_all_one = false;
  h->mask_dirty = true; h->structure_dirty = true;
  return MCC_OK;
}

int mcc_num_reduced_cam_cols(const mcc_handle* hc) {
  mcc_handle* h = const_cast<mcc_handle*>(hc);
  if (!h) return 0;
  int C = 0;
  for (int k = 0; k < h->ncam(); k++) if (!(h->fix_all || h->cams[k].fix_intr)) C += h->cams[k].nparam();
  for (int k = 1; k < h->ncam(); k++) if (!h->cams[k].fix_extr) C += 6;
  return C;
}
long long mcc_num_active_frames(const mcc_handle* h) {
  if (!h) return 0;
  if (h->group) { long long t = 0; for (mcc_handle* k : h->group->kids) t += k->nslots; return t; }
  return h->nslots;
}

mcc_status mcc_eval_normal_blocks(mcc_handle* h, int use_loss, double* cost, double* Hcc, double* gc, double* Hff, double* gf, double* Hfc, int* frame_ids) {
  if (!h) return MCC_ERR_BAD_ARG;
  if (h->group) return fail(h, MCC_ERR_BAD_ARG, "the block test hook works on single-device handles");
  ST(prepare(h));
  ST(launch_jacobian(h, use_loss, h->cur, h->blk));
  ST(read_scalars(h));
  if (cost) *cost = h->h_scal[0];
  const size_t C = h->C, ns = h->nslots;
  if (Hcc && C) CU(cudaMemcpyAsync(Hcc, h->d_Hcc2[h->blk].p, C * C * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  if (gc && C) CU(cudaMemcpyAsync(gc, h->d_gc2[h->blk].p, C * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  if (Hff && ns) CU(cudaMemcpyAsync(Hff, h->d_Hff2[h->blk].p, ns * 36 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  if (gf && ns) CU(cudaMemcpyAsync(gf, h->d_gf2[h->blk].p, ns * 6 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  if (Hfc && ns && C) CU(cudaMemcpyAsync(Hfc, h->d_Hfc2[h->blk].p, ns * 6 * C * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CU(cudaStreamSynchronize(h->stream));
  if (frame_ids) for (size_t s = 0; s < ns; s++) frame_ids[s] = h->slot_frame[s];
  return MCC_OK;
}

mcc_status mcc_comm_unique_id(unsigned char id[MCC_UNIQUE_ID_BYTES]) {
  if (!id || !nccl().ok()) return MCC_ERR_NCCL;
  ncclUniqueId u;
  if (nccl().GetUniqueId(&u) != 0) return MCC_ERR_NCCL;
  std::memcpy(id, u.internal, MCC_UNIQUE_ID_BYTES);
  return MCC_OK;
}

mcc_status mcc_comm_init(mcc_handle* h, int nranks, int rank, const unsigned char id[MCC_UNIQUE_ID_BYTES]) {
  if (!h || !id || nranks < 1 || rank < 0 || rank >= nranks) return fail(h, MCC_ERR_BAD_ARG, "bad communicator arguments");
  if (h->group || h->member_of) return fail(h, MCC_ERR_BAD_ARG, "a multi-device handle shards inside the process; one process per GPU uses single-device handles");
  if (nranks == 1) { h->nranks = 1; h->rank = 0; drop_iteration_graphs(h); return MCC_OK; }
  if (nranks > 64) return fail(h, MCC_ERR_BAD_ARG, "at most 64 ranks (per-rank slots of the merged all-reduce)");
  if (!nccl().ok()) return fail(h, MCC_ERR_NCCL, "libnccl.so.2 not found");
  ST(ensure_device(h));
  ncclUniqueId u;
  std::memcpy(u.internal, id, MCC_UNIQUE_ID_BYTES);
  NC(nccl().CommInitRank(&h->comm, nranks, u, rank));
  h->nranks = nranks; h->rank = rank;
  drop_iteration_graphs(h);   // captured for another communicator layout
  return MCC_OK;
}

int mcc_comm_transport(const mcc_handle* h) {
  if (h && h->group) return mcc_comm_transport(h->group->kids[0]);
  if (!h || h->nranks <= 1) return 0;
  return h->peer_on ? 1 : 2;
}

mcc_status mcc_set_kernel_timing(mcc_handle* h, int enabled) {
  if (!h) return MCC_ERR_BAD_ARG;
  if (h->group) { for (mcc_handle* k : h->group->kids) mcc_set_kernel_timing(k, enabled); return MCC_OK; }
  h->ktiming = enabled != 0;
  for (int i = 0; i < KC_COUNT; i++) { h->kc_ms[i] = 0.0; h->kc_n[i] = 0; }
  return MCC_OK;
}

int mcc_num_kernel_classes(void) { return KC_COUNT; }
const char* mcc_kernel_class_name(int c) {
  static const char* names[KC_COUNT] = {"setup", "frame_normal", "frame_assemble", "schur_syrk", "solve_cam", "backsub", "tile_cost", "reduce", "other", "exchange"};
  return (c >= 0 && c < KC_COUNT) ? names[c] : "";
}

mcc_status mcc_get_kernel_times(mcc_handle* h, double* ms, long long* counts, int n) {
  if (!h || n < KC_COUNT) return MCC_ERR_BAD_ARG;
  if (h->group) return mcc_get_kernel_times(h->group->kids[0], ms, counts, n);      // the first device's
  if (h->stream) { CU(cudaStreamSynchronize(h->stream)); resolve_ktiming(h); }
  for (int i = 0; i < KC_COUNT; i++) { if (ms) ms[i] = h->kc_ms[i]; if (counts) counts[i] = h->kc_n[i]; }
  return MCC_OK;
}

mcc_status mcc_timer_start(mcc_handle* h) {
  if (!h) return MCC_ERR_BAD_ARG;
  if (h->group) return mcc_timer_start(h->group->kids[0]);
  ST(ensure_device(h));
  if (!h->tm_a) { CU(cudaEventCreate(&h->tm_a)); CU(cudaEventCreate(&h->tm_b)); }
  CU(cudaStreamSynchronize(h->stream));
  CU(cudaEventRecord(h->tm_a, h->stream));
  return MCC_OK;
}

mcc_status mcc_timer_stop(mcc_handle* h, double* ms) {
  if (h && h->group) return mcc_timer_stop(h->group->kids[0], ms);
  if (!h || !ms || !h->tm_a) return MCC_ERR_BAD_ARG;
  CU(cudaEventRecord(h->tm_b, h->stream));
  CU(cudaEventSynchronize(h->tm_b));
  float f = 0.f;
  CU(cudaEventElapsedTime(&f, h->tm_a, h->tm_b));
  *ms = f;
  return MCC_OK;
}

int mcc_debug_lane_map(int n_dist, int rowstride, signed char* lane_u, signed char* lane_q, signed char* partner, int* n_units) {
  if (n_dist < 0 || n_dist > MCC_MAX_DIST || !lane_u || !lane_q || !partner || !n_units) return MCC_ERR_BAD_ARG;
  CamMeta m;
  std::memset(&m, 0, sizeof(m));
  build_units(m, n_dist);
  build_lane_map(m, rowstride);
  for (int l = 0; l < 32; l++) { lane_u[l] = m.lane_u[l]; lane_q[l] = m.lane_q[l]; }
  for (int d = 0; d < 3; d++) for (int u = 0; u < 8; u++) partner[d * 8 + u] = m.unit_o[d][u];
  *n_units = m.nunits;
  return m.ksl;
}

mcc_status mcc_measure_fp64_peak(int device, double* tflops) {
  mcc_handle* h = nullptr;
  if (!tflops) return MCC_ERR_BAD_ARG;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); return MCC_ERR_NO_DEVICE; }
  CU(cudaSetDevice(device));
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 16;
  double* out = nullptr;
  CU(cudaMalloc((void**)&out, (size_t)blocks * threads * sizeof(double)));
  cudaEvent_t a, b;
  CU(cudaEventCreate(&a)); CU(cudaEventCreate(&b));
  double best = 0.0;
  for (int rep = 0; rep < 4; rep++) {
    CU(cudaEventRecord(a, 0));
    k_dfma_peak<<<blocks, threads>>>(out, iters);
    CU(cudaEventRecord(b, 0));
    CU(cudaEventSynchronize(b));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, a, b));
    const double tf = 2.0 * 8.0 * iters * (double)blocks * threads / (ms * 1e-3) / 1e12;
    if (rep > 0) best = std::max(best, tf);
  }
  cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(out);
  *tflops = best;
  return MCC_OK;
}

// -------------------------------------------------------------------------------------------------
// mcc_reprojection_cost_noalloc: the residual-only pass (k_tile_cost's arithmetic) for callers that must not
// allocate device memory.  The per-tile partial sums that the handle-based path keeps in a cudaMalloc'ed buffer
// live here in a fixed file-scope device array instead; its size bounds the number of observations per call.
// -------------------------------------------------------------------------------------------------
}  // extern "C"

namespace {

constexpr int NA_TILE = 128;                                          // observations per warp
constexpr int NA_MAX_TILES = MCC_NOALLOC_MAX_OBS / NA_TILE;           // 32768 partial sums = 256 KiB per GPU
static __device__ __align__(256) double g_noalloc_scratch[NA_MAX_TILES];

// In host code `g_noalloc_scratch` is only a stand-in for the symbol; the device address is looked up once per
// device with cudaGetSymbolAddress and cached here.
constexpr int NA_MAX_DEVICES = 64;
double* g_noalloc_ptr[NA_MAX_DEVICES] = {nullptr};
std::mutex g_noalloc_mu;

__global__ void __launch_bounds__(K1_WARPS * 32)
k_cost_static(size_t n, int model, int nd, const double* __restrict__ intr, const double* __restrict__ Rt,
              const double* __restrict__ oX, const double* __restrict__ oY, const double* __restrict__ oZ,
              const double* __restrict__ oU, const double* __restrict__ oV, const uint8_t* __restrict__ omask,
              double* __restrict__ scratch) {
  __shared__ double smem[K1_WARPS][32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const size_t tile = (size_t)blockIdx.x * K1_WARPS + warp;
  const size_t o0 = tile * NA_TILE;
  if (o0 >= n) return;
  double* sS = smem[warp];
  double* sI = sS + 12;
  if (lane < 12) sS[lane] = Rt[lane];                       // R row-major (9) then t (3)
  else if (lane < 16 + nd) sI[lane - 12] = intr[lane - 12];  // fx fy cx cy D...
  __syncwarp();
  const int count = (int)min((size_t)NA_TILE, n - o0);
  double ssum = 0.0;
  for (int base = 0; base < count; base += 32) {
    const int idx = base + lane;
    if (idx < count) {
      const size_t o = o0 + idx;
      if (!omask || omask[o]) {
        const double X = oX[o], Y = oY[o], Z = oZ[o];
        const double x = sS[0] * X + sS[1] * Y + sS[2] * Z + sS[9];
        const double y = sS[3] * X + sS[4] * Y + sS[5] * Z + sS[10];
        const double z = sS[6] * X + sS[7] * Y + sS[8] * Z + sS[11];
        double u, v, d0, d1;
        if (model == 0) project_equidistant<false>(x, y, z, sI, nd, u, v, nullptr, nullptr, d0, d1, nullptr, nullptr);
        else project_radtan<false>(x, y, z, sI, nd, u, v, nullptr, nullptr, d0, d1, nullptr, nullptr);
        const double ru = u - oU[o], rv = v - oV[o];
        ssum += ru * ru + rv * rv;
      }
    }
  }
  ssum = warp_sum(ssum);
  if (lane == 0) scratch[tile] = ssum;    // every slot in [0, ntiles) is overwritten: nothing has to start at zero
}

// out[0] = sum of scratch[0 .. ntiles) in a fixed order (one CTA); slots beyond ntiles keep older calls' values
// and are never read
__global__ void __launch_bounds__(1024)
k_cost_static_reduce(int ntiles, const double* __restrict__ scratch, double* __restrict__ out) {
  __shared__ double red[1024];
  const int tid = threadIdx.x, nt = blockDim.x;
  const int per = (ntiles + nt - 1) / nt;
  const int i0 = tid * per, i1 = min(ntiles, i0 + per);
  double a = 0.0;
  for (int i = i0; i < i1; i++) a += scratch[i];
  red[tid] = a;
  __syncthreads();
  for (int o = nt / 2; o > 0; o >>= 1) { if (tid < o) red[tid] += red[tid + o]; __syncthreads(); }
  if (tid == 0) out[0] = red[0];
}

mcc_status na_fail(char* errbuf, size_t errbuf_len, mcc_status st, const char* fmt, ...) {
  if (errbuf && errbuf_len) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(errbuf, errbuf_len, fmt, ap);
    va_end(ap);
  }
  return st;
}

}  // namespace

extern "C" {

mcc_status mcc_reprojection_cost_noalloc(int model, int n_dist, const double* d_intr, const double* d_Rt, size_t n,
                                         const double* d_X, const double* d_Y, const double* d_Z, const double* d_U,
                                         const double* d_V, const uint8_t* d_mask, double* d_sum_sq, void* stream,
                                         char* errbuf, size_t errbuf_len) {
  if (errbuf && errbuf_len) errbuf[0] = 0;
  if (model != MCC_MODEL_EQUIDISTANT && model != MCC_MODEL_RADTAN) return na_fail(errbuf, errbuf_len, MCC_ERR_BAD_MODEL, "unknown distortion model: %d", model);
  if (n_dist < 0 || n_dist > MCC_MAX_DIST) return na_fail(errbuf, errbuf_len, MCC_ERR_BAD_ARG, "n_dist must be 0..%d, got %d", MCC_MAX_DIST, n_dist);
  if (n > (size_t)MCC_NOALLOC_MAX_OBS)
    return na_fail(errbuf, errbuf_len, MCC_ERR_TOO_LARGE,
                   "%zu observations exceed MCC_NOALLOC_MAX_OBS = %d (the fixed device scratch holds one partial sum per %d observations); "
                   "split the call or use mcc_eval_cost on a handle", n, MCC_NOALLOC_MAX_OBS, NA_TILE);
  if (!d_intr || !d_Rt || !d_sum_sq || (n && (!d_X || !d_Y || !d_Z || !d_U || !d_V))) return na_fail(errbuf, errbuf_len, MCC_ERR_BAD_ARG, "null device pointer");
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { cudaGetLastError(); return na_fail(errbuf, errbuf_len, MCC_ERR_NO_DEVICE, "cudaGetDevice failed: %s (no CPU fallback)", cudaGetErrorString(e)); }
  if (dev < 0 || dev >= NA_MAX_DEVICES) return na_fail(errbuf, errbuf_len, MCC_ERR_BAD_ARG, "device ordinal %d out of range", dev);
  double* scratch = nullptr;
  {
    // once per device, with that device current: resolve the device address of the static array
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
      double* p = nullptr;
      e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
      if (e != cudaSuccess || !p) { cudaGetLastError(); return na_fail(errbuf, errbuf_len, MCC_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch) failed: %s", cudaGetErrorString(e)); }
      g_noalloc_ptr[dev] = p;
    }
    scratch = g_noalloc_ptr[dev];
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int ntiles = (int)((n + NA_TILE - 1) / NA_TILE);
  if (ntiles) {
    k_cost_static<<<(ntiles + K1_WARPS - 1) / K1_WARPS, K1_WARPS * 32, 0, st>>>(n, model, n_dist, d_intr, d_Rt, d_X, d_Y, d_Z, d_U, d_V, d_mask, scratch);
    e = cudaGetLastError();
    if (e != cudaSuccess) return na_fail(errbuf, errbuf_len, MCC_ERR_CUDA, "k_cost_static launch failed: %s", cudaGetErrorString(e));
  }
  k_cost_static_reduce<<<1, 1024, 0, st>>>(ntiles, scratch, d_sum_sq);
  e = cudaGetLastError();
  if (e != cudaSuccess) return na_fail(errbuf, errbuf_len, MCC_ERR_CUDA, "k_cost_static_reduce launch failed: %s", cudaGetErrorString(e));
  return MCC_OK;
}

void mcc_transform_to_rvec_tvec(const double T[16], double out6[6]) { h_transform_to_rvec_tvec(T, out6); }
void mcc_rvec_tvec_to_transform(const double in6[6], double T[16]) { h_rvec_tvec_to_transform(in6, T); }

}  // extern "C"
