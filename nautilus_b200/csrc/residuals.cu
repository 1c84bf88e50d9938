// Batched evaluation of the lidar residual blocks that follow the point-to-point
// matching, with analytic Jacobians, on sm_100a.  Replaces
//   LIDARNormalResidual::operator()  src/optimization/slam_residuals.h:64-90
//   LIDARPointResidual::operator()   src/optimization/slam_residuals.h:123-145
//   PoseArrayToAffine                src/util/slam_util.h:20-28
// evaluated by ceres::AutoDiffCostFunction<..., DYNAMIC, 3, 3> (slam_residuals.h:104-113,
// 159-168): residuals[2k], residuals[2k+1] per correspondence k, jacobians[0] /
// jacobians[1] row-major [2 n][3] with respect to the source / target pose (x, y, theta).
//
//   source_to_target = (T(t_t) R(th_t))^-1 * (T(t_s) R(th_s))
//   p' = R(th_t)^T (R(th_s) p + t_s - t_t)
//   dp'/dt_s = R(th_t)^T      dp'/dth_s = R(th_t)^T J R(th_s) p      (J = [[0,-1],[1,0]])
//   dp'/dt_t = -R(th_t)^T     dp'/dth_t = -J p'
// One thread per correspondence; HBM-bound streaming (32 B in, 16 + 96 B out).
#include "csm_handle.h"

namespace csm {

struct ResidualArgs {
  const double* src_pose;   // [n_blocks][3]
  const double* tgt_pose;   // [n_blocks][3]
  const long long* offsets; // [n_blocks + 1]
  int n_blocks;
  const float2 *sp, *tp, *sn, *tn;  // [n_corr]
  double* res;              // [n_corr][2]
  double* jac_s;            // [n_corr][2][3] or null
  double* jac_t;
  int kind;                 // 0: point residual, 1: normal residual
  const double4* trig;      // [n_blocks] sin/cos of the source and target headings
};

// sin / cos of a block's two headings, once per block instead of once per
// correspondence (the same device sincos on the same arguments: same bits).
__global__ void pose_trig_kernel(const double* __restrict__ src_pose,
                                 const double* __restrict__ tgt_pose, int n_blocks,
                                 double4* __restrict__ trig) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_blocks) return;
  double ss, cs, st, ct;
  sincos(src_pose[3 * b + 2], &ss, &cs);
  sincos(tgt_pose[3 * b + 2], &st, &ct);
  trig[b] = make_double4(ss, cs, st, ct);
}

// Residuals and Jacobians of correspondence k of residual block `blk`.
__device__ __forceinline__ void residual_math(const ResidualArgs& a, long long k, int blk, float2 pf,
                                              float2 qf, float2 nsf, float2 ntf) {
  const double xs = a.src_pose[3 * blk], ys = a.src_pose[3 * blk + 1];
  const double xt = a.tgt_pose[3 * blk], yt = a.tgt_pose[3 * blk + 1];
  const double4 tr = a.trig[blk];
  const double ss = tr.x, cs = tr.y, st = tr.z, ct = tr.w;
  const double px = pf.x, py = pf.y, qx = qf.x, qy = qf.y;
  // w = R(th_s) p + t_s - t_t (world frame, relative to the target origin)
  const double rx = cs * px - ss * py, ry = ss * px + cs * py;
  const double wx = rx + xs - xt, wy = ry + ys - yt;
  // p' = R(th_t)^T w
  const double ppx = ct * wx + st * wy, ppy = -st * wx + ct * wy;
  // d p' / d (x_s, y_s, th_s, x_t, y_t, th_t), rows = p'_x, p'_y
  const double jrx = -ry, jry = rx;  // J R(th_s) p
  double dx[6], dy[6];
  dx[0] = ct;  dy[0] = -st;
  dx[1] = st;  dy[1] = ct;
  dx[2] = ct * jrx + st * jry;  dy[2] = -st * jrx + ct * jry;
  dx[3] = -ct; dy[3] = st;
  dx[4] = -st; dy[4] = -ct;
  dx[5] = ppy; dy[5] = -ppx;      // -J p'
  double r0, r1, g0[6], g1[6];
  if (a.kind == 0) {              // difference_in_target = target - source'
    r0 = qx - ppx; r1 = qy - ppy;
#pragma unroll
    for (int c = 0; c < 6; ++c) { g0[c] = -dx[c]; g1[c] = -dy[c]; }
  } else {                        // n_t . (p' - q),  n_s . (q - p')
    const double nsx = nsf.x, nsy = nsf.y, ntx = ntf.x, nty = ntf.y;
    r0 = ntx * (ppx - qx) + nty * (ppy - qy);
    r1 = nsx * (qx - ppx) + nsy * (qy - ppy);
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      g0[c] = ntx * dx[c] + nty * dy[c];
      g1[c] = -(nsx * dx[c] + nsy * dy[c]);
    }
  }
  reinterpret_cast<double2*>(a.res)[k] = make_double2(r0, r1);
  // 48 bytes per correspondence and array, 16-byte aligned: three 16-byte stores
  if (a.jac_s) {
    double2* j = reinterpret_cast<double2*>(a.jac_s + 6 * k);
    j[0] = make_double2(g0[0], g0[1]); j[1] = make_double2(g0[2], g1[0]); j[2] = make_double2(g1[1], g1[2]);
  }
  if (a.jac_t) {
    double2* j = reinterpret_cast<double2*>(a.jac_t + 6 * k);
    j[0] = make_double2(g0[3], g0[4]); j[1] = make_double2(g0[5], g1[3]); j[2] = make_double2(g1[4], g1[5]);
  }
}

// residual block of correspondence k: last b with offsets[b] <= k.  One binary
// search per CUDA block (for its first correspondence), then every thread walks
// forward from there: consecutive correspondences sit in the same or the next
// few residual blocks.
__device__ __forceinline__ int block_of(const long long* __restrict__ offsets, int n_blocks,
                                        long long k, long long k_first, int* s_first) {
  if (threadIdx.x == 0) {
    int lo = 0, hi = n_blocks;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (offsets[mid] <= k_first) lo = mid; else hi = mid;
    }
    *s_first = lo;
  }
  __syncthreads();
  int b = *s_first;
  while (b + 1 < n_blocks && offsets[b + 1] <= k) ++b;
  return b;
}

__global__ void __launch_bounds__(256) residual_kernel(ResidualArgs a, long long n_corr) {
  __shared__ int s_first;
  const long long k0 = (long long)blockIdx.x * blockDim.x, k = k0 + threadIdx.x;
  const int blk = block_of(a.offsets, a.n_blocks, min(k, n_corr - 1), k0, &s_first);
  if (k >= n_corr) return;
  const float2 z = make_float2(0.f, 0.f);
  residual_math(a, k, blk, a.sp[k], a.tp[k], a.kind ? a.sn[k] : z, a.kind ? a.tn[k] : z);
}

// The same over the resident correspondence set of nautilus_nn_match_resident:
// points and normals are gathered from the scan set on the device.
__global__ void __launch_bounds__(256) residual_resident_kernel(
    ResidualArgs a, long long n_corr, const int4* __restrict__ corr,
    const float2* __restrict__ points, const float2* __restrict__ normals) {
  __shared__ int s_first;
  const long long k0 = (long long)blockIdx.x * blockDim.x, k = k0 + threadIdx.x;
  const int blk = block_of(a.offsets, a.n_blocks, min(k, n_corr - 1), k0, &s_first);
  if (k >= n_corr) return;
  const int4 c = corr[k];
  const float2 z = make_float2(0.f, 0.f);
  residual_math(a, k, blk, points[c.x], points[c.y], a.kind ? normals[c.z] : z,
                a.kind ? normals[c.w] : z);
}

}  // namespace csm

extern "C" int nautilus_residuals_eval(NautilusCsm* h, int32_t kind, const double* source_poses,
                                       const double* target_poses, const int64_t* offsets,
                                       int32_t n_blocks, const float* source_points,
                                       const float* target_points, const float* source_normals,
                                       const float* target_normals, double* residuals,
                                       double* jac_source, double* jac_target) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  if (kind != 0 && kind != 1) return fail(NAUTILUS_EINVAL, "kind must be 0 (point) or 1 (normal)");
  if (n_blocks < 0) return fail(NAUTILUS_EINVAL, "n_blocks < 0");
  if (n_blocks == 0) return NAUTILUS_OK;
  if (!source_poses || !target_poses || !offsets || !residuals)
    return fail(NAUTILUS_EINVAL, "null argument");
  if (offsets[0] != 0) return fail(NAUTILUS_EINVAL, "offsets[0] must be 0");
  for (int b = 0; b < n_blocks; ++b)
    if (offsets[b + 1] < offsets[b]) return fail(NAUTILUS_EINVAL, "offsets must be non-decreasing");
  const long long n = offsets[n_blocks];
  if (n == 0) return NAUTILUS_OK;
  if (!source_points || !target_points || (kind == 1 && (!source_normals || !target_normals)))
    return fail(NAUTILUS_EINVAL, "null correspondence array");
  CU_TRY(cudaSetDevice(h->device));
  cudaStream_t st = h->stream;
  // one device arena per call: poses | offsets | 4 float2 arrays | residuals | jacobians
  const size_t b_pose = sizeof(double) * 3 * (size_t)n_blocks;
  const size_t b_off = sizeof(long long) * ((size_t)n_blocks + 1);
  const size_t b_pts = sizeof(float2) * (size_t)n;
  const size_t b_res = sizeof(double) * 2 * (size_t)n;
  const size_t b_jac = sizeof(double) * 6 * (size_t)n;
  auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
  const size_t b_trig = sizeof(double4) * (size_t)n_blocks;
  const size_t total = 2 * up(b_pose) + up(b_off) + 4 * up(b_pts) + up(b_res) + 2 * up(b_jac) + up(b_trig);
  int rc = h->residual_arena.reserve(total);
  if (rc) return rc;
  unsigned char* base = h->residual_arena.p;
  size_t at = 0;
  auto take = [&](size_t bytes) { unsigned char* p = base + at; at += up(bytes); return p; };
  csm::ResidualArgs a;
  double* d_sp = (double*)take(b_pose); double* d_tp = (double*)take(b_pose);
  long long* d_off = (long long*)take(b_off);
  float2* d_p = (float2*)take(b_pts); float2* d_q = (float2*)take(b_pts);
  float2* d_ns = (float2*)take(b_pts); float2* d_nt = (float2*)take(b_pts);
  double* d_res = (double*)take(b_res);
  double* d_js = (double*)take(b_jac); double* d_jt = (double*)take(b_jac);
  double4* d_trig = (double4*)take(b_trig);
  CU_TRY(cudaMemcpyAsync(d_sp, source_poses, b_pose, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(d_tp, target_poses, b_pose, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(d_off, offsets, b_off, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(d_p, source_points, b_pts, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(d_q, target_points, b_pts, cudaMemcpyHostToDevice, st));
  if (kind == 1) {
    CU_TRY(cudaMemcpyAsync(d_ns, source_normals, b_pts, cudaMemcpyHostToDevice, st));
    CU_TRY(cudaMemcpyAsync(d_nt, target_normals, b_pts, cudaMemcpyHostToDevice, st));
  }
  a.src_pose = d_sp; a.tgt_pose = d_tp; a.offsets = d_off; a.n_blocks = n_blocks;
  a.sp = d_p; a.tp = d_q; a.sn = d_ns; a.tn = d_nt;
  a.res = d_res; a.jac_s = jac_source ? d_js : nullptr; a.jac_t = jac_target ? d_jt : nullptr;
  a.kind = kind;
  a.trig = d_trig;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, st);
  csm::pose_trig_kernel<<<(n_blocks + 255) / 256, 256, 0, st>>>(d_sp, d_tp, n_blocks, d_trig);
  csm::residual_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a, n);
  cudaEventRecord(e1, st);
  CU_TRY(cudaGetLastError());
  CU_TRY(cudaMemcpyAsync(residuals, d_res, b_res, cudaMemcpyDeviceToHost, st));
  if (jac_source) CU_TRY(cudaMemcpyAsync(jac_source, d_js, b_jac, cudaMemcpyDeviceToHost, st));
  if (jac_target) CU_TRY(cudaMemcpyAsync(jac_target, d_jt, b_jac, cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  cudaEventElapsedTime(&h->residual_kernel_ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return NAUTILUS_OK;
}

extern "C" int nautilus_residuals_eval_resident(NautilusCsm* h, int32_t kind,
                                                const double* source_poses,
                                                const double* target_poses, double* residuals,
                                                double* jac_source, double* jac_target) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  if (kind != 0 && kind != 1) return fail(NAUTILUS_EINVAL, "kind must be 0 (point) or 1 (normal)");
  NnState& nn = h->nn;
  if (!nn.built || nn.corr_pairs <= 0)
    return fail(NAUTILUS_ESTATE, "no resident correspondence set: call nautilus_nn_match_resident");
  if (kind == 1 && !nn.normals)
    return fail(NAUTILUS_ESTATE, "normal residuals need nautilus_nn_upload_normals");
  const int n_blocks = nn.corr_pairs;
  const long long n = nn.corr_off_host[n_blocks];
  if (n == 0) return NAUTILUS_OK;
  if (!source_poses || !target_poses || !residuals) return fail(NAUTILUS_EINVAL, "null argument");
  cudaStream_t st = h->stream;
  const size_t b_pose = sizeof(double) * 3 * (size_t)n_blocks;
  const size_t b_res = sizeof(double) * 2 * (size_t)n;
  const size_t b_jac = sizeof(double) * 6 * (size_t)n;
  auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
  const size_t b_trig = sizeof(double4) * (size_t)n_blocks;
  const size_t total = 2 * up(b_pose) + up(b_res) + 2 * up(b_jac) + up(b_trig);
  int rc = h->residual_arena.reserve(total);
  if (rc) return rc;
  unsigned char* base = h->residual_arena.p;
  size_t at = 0;
  auto take = [&](size_t bytes) { unsigned char* p = base + at; at += up(bytes); return p; };
  double* d_sp = (double*)take(b_pose); double* d_tp = (double*)take(b_pose);
  double* d_res = (double*)take(b_res);
  double* d_js = (double*)take(b_jac); double* d_jt = (double*)take(b_jac);
  double4* d_trig = (double4*)take(b_trig);
  CU_TRY(cudaMemcpyAsync(d_sp, source_poses, b_pose, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(d_tp, target_poses, b_pose, cudaMemcpyHostToDevice, st));
  csm::ResidualArgs a;
  a.src_pose = d_sp; a.tgt_pose = d_tp; a.offsets = nn.corr_off; a.n_blocks = n_blocks;
  a.sp = a.tp = a.sn = a.tn = nullptr;
  a.res = d_res; a.jac_s = jac_source ? d_js : nullptr; a.jac_t = jac_target ? d_jt : nullptr;
  a.kind = kind;
  a.trig = d_trig;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, st);
  csm::pose_trig_kernel<<<(n_blocks + 255) / 256, 256, 0, st>>>(d_sp, d_tp, n_blocks, d_trig);
  csm::residual_resident_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(
      a, n, nn.corr, h->user_scans.points, nn.normals);
  cudaEventRecord(e1, st);
  CU_TRY(cudaGetLastError());
  CU_TRY(cudaMemcpyAsync(residuals, d_res, b_res, cudaMemcpyDeviceToHost, st));
  if (jac_source) CU_TRY(cudaMemcpyAsync(jac_source, d_js, b_jac, cudaMemcpyDeviceToHost, st));
  if (jac_target) CU_TRY(cudaMemcpyAsync(jac_target, d_jt, b_jac, cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  cudaEventElapsedTime(&h->residual_kernel_ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return NAUTILUS_OK;
}

extern "C" int nautilus_residuals_last_ms(const NautilusCsm* h, float* kernel_ms) {
  if (!h || !kernel_ms) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  *kernel_ms = h->residual_kernel_ms;
  return NAUTILUS_OK;
}
