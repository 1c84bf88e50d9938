// Host-side handle of libnautilus_csm_b200.so, shared by the C-ABI translation
// units (csm_capi.cu: tables + search, nn_capi.cu: nearest-neighbour matching).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/nautilus_csm.h"
#include "csm_kernels.h"
#include "nn_kernels.h"

namespace csmh {
using namespace csm;

// thread-local error text behind nautilus_csm_last_error() (defined in csm_capi.cu)
std::string& error_text();
inline int fail(int code, const std::string& msg) {
  error_text() = msg;
  return code;
}

#define CU_TRY(expr)                                                              \
  do {                                                                            \
    cudaError_t e_ = (expr);                                                      \
    if (e_ != cudaSuccess) {                                                      \
      int code_ = (e_ == cudaErrorMemoryAllocation) ? NAUTILUS_ENOMEM : NAUTILUS_ECUDA; \
      return fail(code_, std::string(#expr) + ": " + cudaGetErrorString(e_));     \
    }                                                                             \
  } while (0)

// CImg<T>::vanvliet coefficients (CImg.h:49267-49284) + Triggs matrix (49066-49081)
inline void vanvliet_coeffs(double sigma, VanVliet* f) {
  const double nnsigma = sigma < 0.5 ? 0.5 : sigma, m0 = 1.16680, m1 = 1.10783, m2 = 1.40586,
               m1sq = m1 * m1, m2sq = m2 * m2,
               q = (nnsigma < 3.556 ? -0.2568 + 0.5784 * nnsigma + 0.0561 * nnsigma * nnsigma
                                    : 2.5091 + 0.9804 * (nnsigma - 3.556)),
               qsq = q * q, scale = (m0 + q) * (m1sq + m2sq + 2 * m1 * q + qsq),
               b1 = -q * (2 * m0 * m1 + m1sq + m2sq + (2 * m0 + 4 * m1) * q + 3 * qsq) / scale,
               b2 = qsq * (m0 + 2 * m1 + 3 * q) / scale, b3 = -qsq * q / scale,
               B = (m0 * (m1sq + m2sq)) / scale;
  f->B = B; f->a1 = -b1; f->a2 = -b2; f->a3 = -b3;
  const double a1 = f->a1, a2 = f->a2, a3 = f->a3,
               scaleM = 1. / ((1. + a1 - a2 + a3) * (1. - a1 - a2 - a3) *
                              (1. + a2 + (a1 - a3) * a3));
  f->M[0] = scaleM * (-a3 * a1 + 1. - a3 * a3 - a2);
  f->M[1] = scaleM * (a3 + a1) * (a2 + a3 * a1);
  f->M[2] = scaleM * a3 * (a1 + a3 * a2);
  f->M[3] = scaleM * (a1 + a3 * a2);
  f->M[4] = -scaleM * (a2 - 1.) * (a2 + a3 * a1);
  f->M[5] = -scaleM * a3 * (a3 * a1 + a3 * a3 + a2 - 1.);
  f->M[6] = scaleM * (a3 * a1 + a2 + a1 * a1 - a2 * a2);
  f->M[7] = scaleM * (a1 * a2 + a3 * a2 * a2 - a1 * a3 * a3 - a3 * a3 * a3 - a3 * a2 + a3);
  f->M[8] = scaleM * a3 * (a1 + a3 * a2);
}

// Pooling windows of one axis (CorrelativeScanMatcher.cc:37-45 + MaxArea h:91-101).
struct PoolListHost {
  std::vector<int> s, e, dst;
  std::vector<unsigned char> win;
  bool tiles = false;  // windows tile [0, dim) in order without gaps/overlap
};

struct DevPool {
  PoolListDev d{nullptr, nullptr, nullptr, nullptr, 0};
  void release() {
    cudaFree(d.s); cudaFree(d.e); cudaFree(d.dst); cudaFree(d.win);
    d = PoolListDev{nullptr, nullptr, nullptr, nullptr, 0};
  }
  int upload(const PoolListHost& h) {
    release();
    const int n = (int)h.s.size();
    d.n = n;
    if (n == 0) return NAUTILUS_OK;
    CU_TRY(cudaMalloc(&d.s, sizeof(int) * n));
    CU_TRY(cudaMalloc(&d.e, sizeof(int) * n));
    CU_TRY(cudaMalloc(&d.dst, sizeof(int) * n));
    CU_TRY(cudaMalloc(&d.win, n));
    CU_TRY(cudaMemcpy(d.s, h.s.data(), sizeof(int) * n, cudaMemcpyHostToDevice));
    CU_TRY(cudaMemcpy(d.e, h.e.data(), sizeof(int) * n, cudaMemcpyHostToDevice));
    CU_TRY(cudaMemcpy(d.dst, h.dst.data(), sizeof(int) * n, cudaMemcpyHostToDevice));
    CU_TRY(cudaMemcpy(d.win, h.win.data(), n, cudaMemcpyHostToDevice));
    bool ok = h.tiles && n <= 1024;
    for (int k = 0; ok && k < n; ++k) ok = h.e[k] - h.s[k] + 1 >= 8;
    d.two_run_ok = ok ? 1 : 0;
    return NAUTILUS_OK;
  }
};

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t cap = 0;
  int reserve(size_t n) {
    if (n <= cap) return NAUTILUS_OK;
    cudaFree(p);
    p = nullptr; cap = 0;
    const size_t want = std::max(n, (size_t)16);
    CU_TRY(cudaMalloc(&p, sizeof(T) * want));
    cap = want;
    return NAUTILUS_OK;
  }
  void release() { cudaFree(p); p = nullptr; cap = 0; }
};

// Pinned host staging that only grows: the per-call lists go H2D and the results
// come D2H through it, so a steady-state call allocates nothing on the host and
// its copies are true asynchronous DMA.
template <typename T>
struct PinBuf {
  T* p = nullptr;
  size_t cap = 0;
  int reserve(size_t n) {
    if (n <= cap) return NAUTILUS_OK;
    if (p) cudaFreeHost(p);
    p = nullptr; cap = 0;
    const size_t want = std::max(n + n / 4, (size_t)16);
    CU_TRY(cudaHostAlloc((void**)&p, sizeof(T) * want, cudaHostAllocDefault));
    cap = want;
    return NAUTILUS_OK;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

// A set of clouds resident on the device.
struct ScanSet {
  float2* points = nullptr;
  long long* offsets = nullptr;
  std::vector<long long> host_offsets;
  std::vector<float> host_xy;  // host copy of the clouds (KD-tree build of the NN path)
  PinBuf<long long> offsets_pin;  // staging of the offsets for uploads from pinned clouds
  size_t cap_points = 0, cap_offsets = 0;
  int n = 0;
  void release() {
    cudaFree(points); cudaFree(offsets);
    points = nullptr; offsets = nullptr; cap_points = cap_offsets = 0; n = 0;
  }
};

// One queued single-pair call (nautilus_csm_match): see the combining queue there.
struct SingleRequest {
  const float* a = nullptr; int64_t na = 0;
  const float* b = nullptr; int64_t nb = 0;
  double rot_min = 0, rot_max = 0;
  double* score = nullptr; float* x = nullptr; float* y = nullptr; float* theta = nullptr;
  NautilusCsmStats* stats = nullptr;
  int rc = 0;
  std::string error;
  bool done = false;
};

}  // namespace csmh

struct NautilusCsm;
// csm_capi.cu: the two halves of a batch match (see there)
int nautilus_internal_match_launch(NautilusCsm* h, csmh::ScanSet& ss, const int32_t* a_ids,
                                   const int32_t* b_ids, const double* rot_min,
                                   const double* rot_max, int32_t n_pairs, size_t result_capacity);
int nautilus_internal_queue_spec_counts(NautilusCsm* h, int** out);
int nautilus_internal_match_timings(NautilusCsm* h, const int* spec_counts);
// nn_capi.cu: joins and drops a host index build begun by nautilus_nn_build_begin
void nautilus_internal_nn_cancel(NautilusCsm* h);
// csm_comm.cu: releases the handle's communicator
void nautilus_internal_comm_release(NautilusCsm* h);
// frontend.cu: makes ScanSet::host_xy of the user scans valid (D2H when the clouds
// were produced on the device)
int nautilus_internal_host_xy(NautilusCsm* h);

// internal header: the handle is written in terms of the two namespaces above
using namespace csm;
using namespace csmh;

// Entry points that take a handle hold its lock for the whole call: the
// reference object is called concurrently from OpenMP threads
// (CorrelativeScanMatcher.cc:381-386, bag_avg_matcher.cc:104-141); calls on one
// handle are serialised here, calls on different handles run concurrently.
// The CUDA current device is per host thread, so the guard also selects the handle's.
#define NAUTILUS_GUARD(h)                                               \
  std::lock_guard<std::recursive_mutex> nautilus_guard_((h)->mu);       \
  cudaSetDevice((h)->device)

// A launched, not yet collected batch (nautilus_internal_match_launch).
struct PendingBatch {
  int n_pairs = 0, n_waves = 0, launches = 0;
  bool rf_serial = false;
  float ms_host_plan = 0.f;
};

struct HostTimer {
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  float ms() const {
    return std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
  }
};

struct NautilusCsm {
  mutable std::recursive_mutex mu;
  NautilusCsmParams params;
  Geometry geo;
  VanVliet vv;
  double rot_step;
  int device;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  cudaStream_t table_stream = nullptr;  // table builds of the next wave overlap the search
  int sm_count = 148;

  // tables
  PoolListHost pool_host;
  DevPool pool_dev;
  int n_slots = 0;
  std::vector<TableSlot> slots_host;
  DevBuf<TableSlot> slots_dev;
  DevBuf<TableMeta> metas_dev;
  double* slab_values = nullptr;
  unsigned* slab_bitmap = nullptr;
  unsigned char* slab_rowflag = nullptr;
  double* slab_low = nullptr;
  double* slab_lowlog = nullptr;
  double* slab_ck = nullptr;  // y-pass checkpoints of the two-run table build (TableSlot::ck)
  DevBuf<double> tail_dev;  // Geometry::tail
  DevBuf<double> dense_scratch;
  DevBuf<double> scalar_scratch;

  // search grid
  std::vector<float> t_low_host;
  std::vector<int> slot_rank_host, n_h_host;
  std::vector<float> t_high_host;
  DevBuf<float> t_low_dev, t_high_dev;
  DevBuf<int> slot_rank_dev, n_h_dev;
  SearchGrid grid{};
  std::vector<float> thr_low_host, thr_high_host;
  DevBuf<float> thr_low_dev, thr_high_dev;
  CellMap cm_low{}, cm_high{};
  int grid_rc = 0;          // build_search_grid result; matching needs 0
  std::string grid_error;

  // scans + per-batch buffers
  ScanSet user_scans, tmp_scans;
  DevBuf<int> table_scan_dev;
  DevBuf<PairWork> pairs_dev;
  DevBuf<Rotation> rots_dev;
  DevBuf<PairResult> results_dev;
  // host-side planning scratch of match_batch, kept between calls
  PinBuf<int> targets_pin;
  PinBuf<PairWork> sorted_pin;
  PinBuf<PairResult> results_pin;
  PinBuf<int> spec_counts_pin;
  std::vector<int> plan_tindex, plan_first_pair, plan_fill;
  std::vector<PairWork> plan_work;
  DevBuf<double> chunk_best_dev;
  DevBuf<int> chunk_rot_dev;
  DevBuf<SpecPair> spec_pairs_dev;
  DevBuf<SpecItem> spec_items_dev;
  DevBuf<SpecOut> spec_out_dev;
  DevBuf<int> spec_counters_dev;
  std::vector<cudaEvent_t> events;
  NautilusCsmTimings timings{};
  int last_n_pairs = 0;
  PendingBatch pending;

  // multi-GPU (csm_comm.cu): one process per GPU, NCCL bound at run time
  void* nccl_comm = nullptr;   // ncclComm_t
  bool comm_owned = false;     // created by nautilus_csm_comm_init (else attached)
  int world_rank = 0, world_size = 1;
  DevBuf<PairResult> gather_dev;
  PinBuf<PairResult> gather_pin;
  std::vector<int> shard_order, shard_a, shard_b;
  std::vector<long long> shard_bounds;
  std::vector<double> shard_w, shard_rmin, shard_rmax;

  // combining queue of concurrent single-pair callers (nautilus_csm_match)
  mutable std::mutex single_mu;
  std::condition_variable single_cv;
  std::vector<SingleRequest*> single_queue;
  bool single_leader = false;
  long long single_requests = 0, single_batches = 0;
  PinBuf<float> single_pin;

  // residual evaluation (residuals.cu)
  DevBuf<unsigned char> residual_arena;
  float residual_kernel_ms = 0.f;
  float frontend_kernel_ms = 0.f;  // last nautilus_csm_upload_ranges: count + write passes

  // nearest-neighbour matching (nn_capi.cu)
  csmh::NnState nn;
  void* nn_host_build = nullptr;  // NnHostBuild of nautilus_nn_build_begin (nn_capi.cu)
};
