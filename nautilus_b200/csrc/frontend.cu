// Front end of the scan-matching path on sm_100a: a bag's scans go to the device
// once as ranges (4 B per beam) and everything up to (pose, score) stays there.
// Replaces
//   nautilus::LaserScanToPointCloud            src/input/pointcloud_helpers.cc:28-48
//   pointcloud_helpers::LaserScanToPointCloud  third_party/csm/src/pointcloud_helpers.cc:44-66
//   FilterGlancing                             third_party/csm/src/pointcloud_helpers.cc:21-42
//   ComputeMean / ComputeScatterMatrixScore    src/loop_closure/lc_candidate_filter.cc:23-52
//   GetLCCandidates                            src/loop_closure/lc_candidate_filter.cc:66-81 (host)
//
// A beam's angle depends on its index only, so cosf/sinf and the angle window test
// come from a per-sensor table built on the host with the reference's own float
// loops and libm (Rotation2Df evaluates sinf/cosf of the float angle); the device
// does the products of Rotation2Df(angle) * (range, 0) with separately rounded
// operations, the range tests, the ordered compaction and the glancing filter.
// Both kernels are HBM-bound streaming passes: 4 B in, 8 B out per beam.
#include <math.h>

#include <vector>

#include "csm_handle.h"

namespace csm {

struct BeamEntry {
  float c, s;  // cosf / sinf of the beam angle
  int ok;      // angle inside the accepted window (always 1 for the nautilus variant)
  int pad;
};

constexpr int FE_THREADS = 256;
constexpr int FE_WARPS = FE_THREADS / 32;

// Ordered compaction step: position of this thread's element among the kept ones
// of the block's current chunk (base + exclusive prefix); returns the chunk total.
__device__ __forceinline__ int block_rank(bool keep, int* s_warp, int* pos) {
  const unsigned m = __ballot_sync(0xffffffffu, keep);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  __syncthreads();  // s_warp of the previous chunk has been read
  if (lane == 0) s_warp[wid] = __popc(m);
  __syncthreads();
  int before = 0, total = 0;
#pragma unroll
  for (int k = 0; k < FE_WARPS; ++k) {
    const int c = s_warp[k];
    before += k < wid ? c : 0;
    total += c;
  }
  *pos = before + __popc(m & ((1u << lane) - 1u));
  return total;
}

// One block per scan.  WRITE = false: counts[scan] only.  WRITE = true: points to
// out + offsets[scan].  variant 1 adds the csm glancing filter.
template <bool WRITE>
__global__ void __launch_bounds__(FE_THREADS) ranges_kernel(
    const float* __restrict__ ranges, int n_beams, const BeamEntry* __restrict__ beams,
    float range_min, double max_range, int variant, long long* __restrict__ counts,
    const long long* __restrict__ offsets, float2* __restrict__ out) {
  extern __shared__ __align__(8) unsigned char s_raw[];
  float2* tmp = reinterpret_cast<float2*>(s_raw);  // [n_beams] range-valid points, in order
  __shared__ int s_warp[FE_WARPS];
  const float* r = ranges + (size_t)blockIdx.x * n_beams;
  float2* dst = WRITE ? out + offsets[blockIdx.x] : nullptr;

  int n_valid = 0;
  for (int i0 = 0; i0 < n_beams; i0 += FE_THREADS) {
    const int i = i0 + threadIdx.x;
    bool ok = false;
    float2 p = make_float2(0.f, 0.f);
    if (i < n_beams) {
      const float range = r[i];
      const BeamEntry b = beams[i];
      // range >= range_min && range <= max_range (the latter compared in double)
      ok = b.ok && range >= range_min && (double)range <= max_range;
      p.x = __fadd_rn(__fmul_rn(b.c, range), __fmul_rn(-b.s, 0.0f));
      p.y = __fadd_rn(__fmul_rn(b.s, range), __fmul_rn(b.c, 0.0f));
    }
    int pos;
    const int total = block_rank(ok, s_warp, &pos);
    if (ok) {
      if (variant == 0) { if (WRITE) dst[n_valid + pos] = p; }
      else tmp[n_valid + pos] = p;
    }
    n_valid += total;
  }
  if (variant == 0) {
    if (!WRITE && threadIdx.x == 0) counts[blockIdx.x] = n_valid;
    return;
  }
  __syncthreads();
  // FilterGlancing: keep a point when it lies within 0.10 of its predecessor in
  // the range-valid list (the first one is compared with itself)
  int n_kept = 0;
  for (int k0 = 0; k0 < n_valid; k0 += FE_THREADS) {
    const int k = k0 + threadIdx.x;
    bool keep = false;
    float2 p = make_float2(0.f, 0.f);
    if (k < n_valid) {
      p = tmp[k];
      const float2 q = tmp[k > 0 ? k - 1 : 0];
      const float dx = __fsub_rn(q.x, p.x), dy = __fsub_rn(q.y, p.y);
      const float norm = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
      keep = (double)norm <= 0.10;
    }
    int pos;
    const int total = block_rank(keep, s_warp, &pos);
    if (WRITE && keep) dst[n_kept + pos] = p;
    n_kept += total;
  }
  if (!WRITE && threadIdx.x == 0) counts[blockIdx.x] = n_kept;
}

// offsets[0] = 0, offsets[i + 1] = offsets[i] + counts[i]; one block.
__global__ void __launch_bounds__(1024) offsets_kernel(const long long* __restrict__ counts, int n,
                                                       long long* __restrict__ offsets) {
  __shared__ long long s_part[1024];
  const int per = (n + 1023) / 1024;
  const int b = threadIdx.x * per, e = min(n, b + per);
  long long s = 0;
  for (int i = b; i < e; ++i) s += counts[i];
  s_part[threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    long long run = 0;
    for (int k = 0; k < 1024; ++k) { const long long v = s_part[k]; s_part[k] = run; run += v; }
    offsets[0] = 0;
  }
  __syncthreads();
  long long run = s_part[threadIdx.x];
  for (int i = b; i < e; ++i) { run += counts[i]; offsets[i + 1] = run; }
}

// ComputeScatterMatrixScore: one warp per scan.  The float sums run in point
// order (every lane carries the same running sums; a chunk of 32 points is loaded
// coalesced and passed around by shuffle), so they are the reference's sums.
__global__ void __launch_bounds__(128) scatter_kernel(const float2* __restrict__ points,
                                                      const long long* __restrict__ offsets,
                                                      int n_scans, double* __restrict__ score) {
  const int scan = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (scan >= n_scans) return;
  const int lane = threadIdx.x & 31;
  const float2* a = points + offsets[scan];
  const int n = (int)(offsets[scan + 1] - offsets[scan]);
  float sx = 0.f, sy = 0.f;
  for (int p0 = 0; p0 < n; p0 += 32) {
    const float2 v = p0 + lane < n ? a[p0 + lane] : make_float2(0.f, 0.f);
    const int m = min(32, n - p0);
    for (int k = 0; k < m; ++k) {
      sx = __fadd_rn(sx, __shfl_sync(0xffffffffu, v.x, k));
      sy = __fadd_rn(sy, __shfl_sync(0xffffffffu, v.y, k));
    }
  }
  const float inv = (float)(1.0 / (double)n);  // the double factor is narrowed to float first
  const float mx = __fmul_rn(inv, sx), my = __fmul_rn(inv, sy);
  float m00 = 0.f, m01 = 0.f, m10 = 0.f, m11 = 0.f;
  for (int p0 = 0; p0 < n; p0 += 32) {
    const float2 v = p0 + lane < n ? a[p0 + lane] : make_float2(0.f, 0.f);
    const float dx = __fsub_rn(v.x, mx), dy = __fsub_rn(v.y, my);
    const float xx = __fmul_rn(dx, dx), xy = __fmul_rn(dx, dy), yx = __fmul_rn(dy, dx),
                yy = __fmul_rn(dy, dy);
    const int m = min(32, n - p0);
    for (int k = 0; k < m; ++k) {
      m00 = __fadd_rn(m00, __shfl_sync(0xffffffffu, xx, k));
      m01 = __fadd_rn(m01, __shfl_sync(0xffffffffu, xy, k));
      m10 = __fadd_rn(m10, __shfl_sync(0xffffffffu, yx, k));
      m11 = __fadd_rn(m11, __shfl_sync(0xffffffffu, yy, k));
    }
  }
  if (lane == 0) {
    // eigenvalues of the 2x2 scatter matrix (closed form, double), stored as float
    // like EigenSolver<Matrix2f>::eigenvalues(), then min / max in double
    const double A = m00, B = m01, Cc = m10, D = m11;
    const double tr = A + D, det = A * D - B * Cc;
    const double disc = tr * tr / 4.0 - det;
    const double root = disc > 0 ? sqrt(disc) : 0.0;
    const double e1 = (double)(float)(tr / 2.0 + root), e2 = (double)(float)(tr / 2.0 - root);
    score[scan] = n > 0 ? fmin(e1, e2) / fmax(e1, e2) : nan("");
  }
}

}  // namespace csm

namespace {

int ensure_scan_capacity(ScanSet& ss, size_t total, int n_scans) {
  if (total > ss.cap_points) {
    cudaFree(ss.points); ss.points = nullptr; ss.cap_points = 0;
    CU_TRY(cudaMalloc(&ss.points, sizeof(float2) * std::max<size_t>(total, 1024)));
    ss.cap_points = std::max<size_t>(total, 1024);
  }
  if ((size_t)n_scans + 1 > ss.cap_offsets) {
    cudaFree(ss.offsets); ss.offsets = nullptr; ss.cap_offsets = 0;
    CU_TRY(cudaMalloc(&ss.offsets, sizeof(long long) * std::max<size_t>(n_scans + 1, 64)));
    ss.cap_offsets = std::max<size_t>(n_scans + 1, 64);
  }
  return NAUTILUS_OK;
}

}  // namespace

// host copy of the resident clouds (the NN path builds its KD-trees on the host)
int nautilus_internal_host_xy(NautilusCsm* h) {
  ScanSet& ss = h->user_scans;
  const long long total = ss.n > 0 ? ss.host_offsets[ss.n] : 0;
  if ((long long)ss.host_xy.size() == 2 * total) return NAUTILUS_OK;
  ss.host_xy.resize(2 * total);
  if (total > 0) {
    CU_TRY(cudaMemcpyAsync(ss.host_xy.data(), ss.points, sizeof(float2) * total,
                           cudaMemcpyDeviceToHost, h->stream));
    CU_TRY(cudaStreamSynchronize(h->stream));
  }
  return NAUTILUS_OK;
}

extern "C" {

int nautilus_csm_upload_ranges(NautilusCsm* h, const float* ranges, int32_t n_scans,
                               int32_t n_beams, const NautilusLaserSpec* spec) {
  if (!h || !spec) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  if (n_scans < 0 || n_beams < 0 || (n_scans > 0 && n_beams > 0 && !ranges))
    return fail(NAUTILUS_EINVAL, "bad range batch");
  if (spec->variant != 0 && spec->variant != 1) return fail(NAUTILUS_EINVAL, "unknown variant");
  const size_t smem = spec->variant == 1 ? sizeof(float2) * (size_t)n_beams : 0;
  if (smem > 200 * 1024) return fail(NAUTILUS_EINVAL, "too many beams per scan for the glancing filter");
  CU_TRY(cudaSetDevice(h->device));
  h->nn.built = false;  // the NN index belongs to the previous scan set
  nautilus_internal_nn_cancel(h);
  ScanSet& ss = h->user_scans;
  cudaStream_t st = h->stream;

  // per-sensor beam table, the reference's float loops on the host
  std::vector<BeamEntry> beams(n_beams);
  if (spec->variant == 0) {
    for (int i = 0; i < n_beams; ++i) {
      // angle_min + angle_increment * index, index converted size_t -> float (cc:38-40)
      const float a = spec->angle_min + spec->angle_increment * (float)(size_t)i;
      beams[i] = BeamEntry{cosf(a), sinf(a), 1, 0};
    }
  } else {
    float a = spec->angle_min;                                       // csm cc:48
    const float trunc = spec->truncate_ends ? (float)(M_PI / 12.0) : 0.f;  // csm cc:49
    const float lo = spec->angle_min + trunc, hi = spec->angle_max - trunc;
    for (int i = 0; i < n_beams; ++i) {
      beams[i] = BeamEntry{cosf(a), sinf(a), (a > lo && a < hi) ? 1 : 0, 0};
      a += spec->angle_increment;                                    // csm cc:63
    }
  }
  BeamEntry* beams_dev = nullptr;
  float* ranges_dev = nullptr;
  long long* counts_dev = nullptr;
  const size_t n_r = (size_t)n_scans * n_beams;
  int rc = NAUTILUS_OK;
  auto cleanup = [&]() { cudaFree(beams_dev); cudaFree(ranges_dev); cudaFree(counts_dev); };
#define FE_TRY(expr)                                                         \
  do {                                                                       \
    cudaError_t e_ = (expr);                                                 \
    if (e_ != cudaSuccess) {                                                 \
      cleanup();                                                             \
      return fail(e_ == cudaErrorMemoryAllocation ? NAUTILUS_ENOMEM : NAUTILUS_ECUDA, \
                  std::string(#expr) + ": " + cudaGetErrorString(e_));      \
    }                                                                        \
  } while (0)
  FE_TRY(cudaMalloc(&beams_dev, sizeof(BeamEntry) * std::max(n_beams, 1)));
  FE_TRY(cudaMalloc(&ranges_dev, sizeof(float) * std::max<size_t>(n_r, 1)));
  FE_TRY(cudaMalloc(&counts_dev, sizeof(long long) * std::max(n_scans, 1)));
  if ((rc = ensure_scan_capacity(ss, 0, n_scans))) { cleanup(); return rc; }
  if (n_beams > 0) FE_TRY(cudaMemcpyAsync(beams_dev, beams.data(), sizeof(BeamEntry) * n_beams,
                                          cudaMemcpyHostToDevice, st));
  if (n_r > 0) FE_TRY(cudaMemcpyAsync(ranges_dev, ranges, sizeof(float) * n_r, cudaMemcpyHostToDevice, st));
  ss.host_offsets.assign((size_t)n_scans + 1, 0);
  ss.host_xy.clear();
  if (n_scans > 0) {
    if (smem > 48 * 1024) {
      cudaFuncSetAttribute(ranges_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(ranges_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    }
    cudaEvent_t e0, e1, e2, e3;
    cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2); cudaEventCreate(&e3);
    cudaEventRecord(e0, st);
    ranges_kernel<false><<<n_scans, FE_THREADS, smem, st>>>(ranges_dev, n_beams, beams_dev,
                                                            spec->range_min, spec->max_range,
                                                            spec->variant, counts_dev, nullptr, nullptr);
    offsets_kernel<<<1, 1024, 0, st>>>(counts_dev, n_scans, ss.offsets);
    cudaEventRecord(e1, st);
    FE_TRY(cudaMemcpyAsync(ss.host_offsets.data(), ss.offsets, sizeof(long long) * (n_scans + 1),
                           cudaMemcpyDeviceToHost, st));
    FE_TRY(cudaStreamSynchronize(st));
    const long long total = ss.host_offsets[n_scans];
    if ((rc = ensure_scan_capacity(ss, (size_t)total, n_scans))) { cleanup(); return rc; }
    cudaEventRecord(e2, st);
    ranges_kernel<true><<<n_scans, FE_THREADS, smem, st>>>(ranges_dev, n_beams, beams_dev,
                                                           spec->range_min, spec->max_range,
                                                           spec->variant, nullptr, ss.offsets, ss.points);
    cudaEventRecord(e3, st);
    FE_TRY(cudaGetLastError());
    FE_TRY(cudaStreamSynchronize(st));
    float ta = 0.f, tb = 0.f;
    cudaEventElapsedTime(&ta, e0, e1);
    cudaEventElapsedTime(&tb, e2, e3);
    h->frontend_kernel_ms = ta + tb;
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2); cudaEventDestroy(e3);
  } else {
    FE_TRY(cudaMemsetAsync(ss.offsets, 0, sizeof(long long), st));
    FE_TRY(cudaStreamSynchronize(st));
  }
#undef FE_TRY
  cleanup();
  ss.n = n_scans;
  return NAUTILUS_OK;
}

int nautilus_csm_frontend_last_ms(const NautilusCsm* h, float* kernel_ms) {
  if (!h || !kernel_ms) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  *kernel_ms = h->frontend_kernel_ms;
  return NAUTILUS_OK;
}

int nautilus_csm_download_scans(NautilusCsm* h, float* xy, int64_t cap_points, int64_t* offsets) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  ScanSet& ss = h->user_scans;
  if (ss.host_offsets.empty()) return fail(NAUTILUS_ESTATE, "no scans uploaded");
  CU_TRY(cudaSetDevice(h->device));
  const long long total = ss.host_offsets[ss.n];
  if (offsets)
    for (int i = 0; i <= ss.n; ++i) offsets[i] = ss.host_offsets[i];
  if (xy) {
    if (cap_points < total) return fail(NAUTILUS_EINVAL, "xy buffer too small");
    if (total > 0) {
      CU_TRY(cudaMemcpyAsync(xy, ss.points, sizeof(float2) * total, cudaMemcpyDeviceToHost, h->stream));
      CU_TRY(cudaStreamSynchronize(h->stream));
    }
  }
  return NAUTILUS_OK;
}

int nautilus_lc_scatter_scores(NautilusCsm* h, double* score) {
  if (!h || !score) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  ScanSet& ss = h->user_scans;
  if (ss.n == 0) return fail(NAUTILUS_ESTATE, "no scans uploaded");
  CU_TRY(cudaSetDevice(h->device));
  int rc = h->scalar_scratch.reserve(ss.n);
  if (rc) return rc;
  scatter_kernel<<<(ss.n + 3) / 4, 128, 0, h->stream>>>(ss.points, ss.offsets, ss.n, h->scalar_scratch.p);
  CU_TRY(cudaGetLastError());
  CU_TRY(cudaMemcpyAsync(score, h->scalar_scratch.p, sizeof(double) * ss.n, cudaMemcpyDeviceToHost, h->stream));
  CU_TRY(cudaStreamSynchronize(h->stream));
  return NAUTILUS_OK;
}

int nautilus_lc_candidates(const double* score, const float* pose_xy, int32_t n, double min_dist,
                           double min_score, int64_t* out, int32_t* n_out) {
  if (n < 0 || !n_out || (n > 0 && (!score || !pose_xy || !out)))
    return fail(NAUTILUS_EINVAL, "bad argument");
  int m = 0;
  for (int i = 0; i < n; ++i) {
    if (m > 0) {  // DistantFromLastScan (lc_candidate_filter.cc:54-64): float norm, double compare
      const int64_t l = out[m - 1];
      const float dx = pose_xy[2 * i] - pose_xy[2 * l], dy = pose_xy[2 * i + 1] - pose_xy[2 * l + 1];
      const float t0 = dx * dx, t1 = dy * dy;
      if (!((double)sqrtf(t0 + t1) >= min_dist)) continue;
    }
    if (score[i] >= min_score) out[m++] = i;
  }
  *n_out = m;
  return NAUTILUS_OK;
}

}  // extern "C"
