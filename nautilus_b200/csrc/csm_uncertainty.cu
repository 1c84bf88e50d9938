// Olson's covariance estimate on sm_100a.  Replaces
//   CorrelativeScanMatcher::GetUncertaintyMatrix (3x3)  CorrelativeScanMatcher.cc:267-319
//   CorrelativeScanMatcher::GetUncertaintyMatrix (2x2)  CorrelativeScanMatcher.cc:322-373
// Both walk rotations x low-grid translations, take the low-res memo value when
// it is negligible (<= log 1e-4, cc:12) and the cost on the HIGH-res table at the
// low-grid translation otherwise, and accumulate exp(cost)-weighted moments in
// float matrices.  The costs are computed in parallel (unc_cost_kernel); the
// moments are accumulated by one thread in the reference's loop order
// (unc_moments_kernel), because float accumulation is order-dependent.
#include "csm_common.cuh"
#include "csm_kernels.h"

namespace csm {

__device__ __forceinline__ float2 unc_rotate(float2 p, float c, float s) {
  float2 q;  // Rotation2Df(theta).toRotationMatrix() * p, Eigen's evaluation order
  q.x = __fadd_rn(__fmul_rn(c, p.x), __fmul_rn(-s, p.y));
  q.y = __fadd_rn(__fmul_rn(s, p.x), __fmul_rn(c, p.y));
  return q;
}

// grid.x = rotation; one thread per low-grid translation tr = i * n_t + j.
__global__ void __launch_bounds__(256) unc_cost_kernel(
    SearchGrid sg, CellMap cm_high, const float2* __restrict__ points,
    const long long* __restrict__ offsets, const TableSlot* __restrict__ slots,
    const TableMeta* __restrict__ metas, PairWork pw, const Rotation* __restrict__ rots,
    SearchBuffers sb, double useless_threshold, double* __restrict__ costs) {
  const int r = blockIdx.x;
  const int n_t = sg.n_t, n_tr = n_t * n_t;
  const float2* a = points + offsets[pw.a_scan];
  const int n_pts = (int)(offsets[pw.a_scan + 1] - offsets[pw.a_scan]);
  const TableMeta tm = metas[pw.table];
  const double* __restrict__ hi = slots[pw.table].values;  // lookup_log of every cell
  const Rotation rot = rots[pw.rot_begin + r];
  const int n_chunks = (pw.n_rot + sb.rot_per_chunk - 1) / sb.rot_per_chunk;
  for (int tr = threadIdx.x; tr < n_tr; tr += blockDim.x) {
    // low_res_costs[AbsCoords(x, y)].first: the memo, best over all rotations (cc:296-297)
    double memo = -INFINITY;
    for (int ch = 0; ch < n_chunks; ++ch) {
      const double s = sb.chunk_best[(size_t)ch * n_tr + tr];
      if (s > memo) memo = s;
    }
    double cost = memo;
    if (!(memo <= useless_threshold)) {
      // CalculatePointcloudCost(rotated a, x_trans, y_trans, high-res table) (cc:301-302)
      const int i = tr / n_t, j = tr - i * n_t;
      const float tx = sg.t_low[i], ty = sg.t_low[j];
      double sum = 0.0;
      for (int p = 0; p < n_pts; ++p) {
        const float2 q = unc_rotate(a[p], rot.c, rot.s);
        const int cx = cell_fast(cm_high, __fadd_rn(q.x, tx));
        const int cy = cell_fast(cm_high, __fadd_rn(q.y, ty));
        double v = CSM_FLOOR_LOG;
        if (cx >= tm.ax0 && cx <= tm.ax1 && cy >= tm.ay0 && cy <= tm.ay1)
          v = __ldg(hi + (size_t)(cy - tm.ay0) * tm.stride + (cx - tm.ax0));
        sum = __dadd_rn(sum, v);
      }
      cost = __ddiv_rn(sum, (double)n_pts);
    }
    costs[(size_t)r * n_tr + tr] = cost;
  }
}

// One thread, the reference's loop order: rotation outer, x, then y (cc:286-312 /
// 346-366).  Float matrices, double s; the scalar cost is narrowed to float before
// it multiplies a float matrix (Eigen's scalar promotion).
__global__ void unc_moments_kernel(SearchGrid sg, const Rotation* __restrict__ rots, int n_rot,
                                   const double* __restrict__ costs, int dims,
                                   float* __restrict__ out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const int n_t = sg.n_t, n_tr = n_t * n_t;
  float K[9], u[3];
  for (int k = 0; k < 9; ++k) K[k] = 0.f;
  for (int k = 0; k < 3; ++k) u[k] = 0.f;
  double s = 0.0;
  for (int r = 0; r < n_rot; ++r) {
    for (int i = 0; i < n_t; ++i) {
      for (int j = 0; j < n_t; ++j) {
        const double cost = exp(costs[(size_t)r * n_tr + i * n_t + j]);
        const float cf = (float)cost;
        float x[3] = {sg.t_low[i], sg.t_low[j], rots[r].theta};
        for (int aa = 0; aa < dims; ++aa) {
          for (int bb = 0; bb < dims; ++bb)
            K[aa * 3 + bb] = __fadd_rn(K[aa * 3 + bb], __fmul_rn(__fmul_rn(x[aa], x[bb]), cf));
          u[aa] = __fadd_rn(u[aa], __fmul_rn(x[aa], cf));
        }
        s = __dadd_rn(s, cost);
      }
    }
  }
  // (1.0 / s) * K - (1.0 / (s * s)) * u * u.transpose()
  const float inv_s = (float)__ddiv_rn(1.0, s);
  const float inv_s2 = (float)__ddiv_rn(1.0, __dmul_rn(s, s));
  for (int aa = 0; aa < dims; ++aa)
    for (int bb = 0; bb < dims; ++bb)
      out[aa * dims + bb] = __fsub_rn(__fmul_rn(K[aa * 3 + bb], inv_s),
                                      __fmul_rn(__fmul_rn(u[aa], inv_s2), u[bb]));
}

// ---- batched 2x2 form: GetLocalUncertaintyStats' inner call for many pairs at once
// (bag_avg_matcher.cc:104-141: 200 base clouds x up to 10 comparison clouds).  Block =
// pair; every pair has ONE rotation (MemoizeLowRes(rotation, rotation + EPSILON),
// cc:341-343), so the memo of a translation is the pair's single chunk value.
__global__ void __launch_bounds__(256) unc_cost_batch_kernel(
    SearchGrid sg, CellMap cm_high, const float2* __restrict__ points,
    const long long* __restrict__ offsets, const TableSlot* __restrict__ slots,
    const TableMeta* __restrict__ metas, const PairWork* __restrict__ pairs,
    const Rotation* __restrict__ rots, SearchBuffers sb, double useless_threshold,
    double* __restrict__ costs) {
  const PairWork pw = pairs[blockIdx.x];
  const int n_t = sg.n_t, n_tr = n_t * n_t;
  const float2* a = points + offsets[pw.a_scan];
  const int n_pts = (int)(offsets[pw.a_scan + 1] - offsets[pw.a_scan]);
  const TableMeta tm = metas[pw.table];
  const double* __restrict__ hi = slots[pw.table].values;  // lookup_log of every cell
  const Rotation rot = rots[pw.rot_begin];
  for (int tr = threadIdx.x; tr < n_tr; tr += blockDim.x) {
    const double memo = sb.chunk_best[(size_t)blockIdx.x * sb.chunks_per_pair * n_tr + tr];
    double cost = memo;
    if (!(memo <= useless_threshold)) {
      const int i = tr / n_t, j = tr - i * n_t;
      const float tx = sg.t_low[i], ty = sg.t_low[j];
      double sum = 0.0;
      for (int p = 0; p < n_pts; ++p) {
        const float2 q = unc_rotate(a[p], rot.c, rot.s);
        const int cx = cell_fast(cm_high, __fadd_rn(q.x, tx));
        const int cy = cell_fast(cm_high, __fadd_rn(q.y, ty));
        double v = CSM_FLOOR_LOG;
        if (cx >= tm.ax0 && cx <= tm.ax1 && cy >= tm.ay0 && cy <= tm.ay1)
          v = __ldg(hi + (size_t)(cy - tm.ay0) * tm.stride + (cx - tm.ax0));
        sum = __dadd_rn(sum, v);
      }
      cost = __ddiv_rn(sum, (double)n_pts);
    }
    costs[(size_t)blockIdx.x * n_tr + tr] = cost;
  }
}

// One thread per pair, the reference's loop order (x outer, y inner, cc:346-366).
__global__ void __launch_bounds__(128) unc_moments_batch_kernel(
    SearchGrid sg, const PairWork* __restrict__ pairs, int n_pairs,
    const double* __restrict__ costs, float* __restrict__ out) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_pairs) return;
  const int n_t = sg.n_t, n_tr = n_t * n_t;
  float K[4] = {0.f, 0.f, 0.f, 0.f}, u[2] = {0.f, 0.f};
  double s = 0.0;
  for (int i = 0; i < n_t; ++i)
    for (int j = 0; j < n_t; ++j) {
      const double cost = exp(costs[(size_t)k * n_tr + i * n_t + j]);
      const float cf = (float)cost;
      const float x[2] = {sg.t_low[i], sg.t_low[j]};
      for (int aa = 0; aa < 2; ++aa) {
        for (int bb = 0; bb < 2; ++bb)
          K[aa * 2 + bb] = __fadd_rn(K[aa * 2 + bb], __fmul_rn(__fmul_rn(x[aa], x[bb]), cf));
        u[aa] = __fadd_rn(u[aa], __fmul_rn(x[aa], cf));
      }
      s = __dadd_rn(s, cost);
    }
  const float inv_s = (float)__ddiv_rn(1.0, s);
  const float inv_s2 = (float)__ddiv_rn(1.0, __dmul_rn(s, s));
  float* o = out + 4 * (size_t)pairs[k].out_index;
  for (int aa = 0; aa < 2; ++aa)
    for (int bb = 0; bb < 2; ++bb)
      o[aa * 2 + bb] = __fsub_rn(__fmul_rn(K[aa * 2 + bb], inv_s),
                                 __fmul_rn(__fmul_rn(u[aa], inv_s2), u[bb]));
}

void launch_uncertainty_batch(const SearchGrid& sg, const CellMap& cm_high, const float2* points,
                              const long long* offsets, const TableSlot* slots,
                              const TableMeta* metas, const PairWork* pairs, const Rotation* rots,
                              int n_pairs, const SearchBuffers& sb, double useless_threshold,
                              double* costs, float* out, cudaStream_t stream) {
  if (n_pairs <= 0) return;
  unc_cost_batch_kernel<<<n_pairs, 256, 0, stream>>>(sg, cm_high, points, offsets, slots, metas, pairs,
                                                     rots, sb, useless_threshold, costs);
  unc_moments_batch_kernel<<<(n_pairs + 127) / 128, 128, 0, stream>>>(sg, pairs, n_pairs, costs, out);
}

void launch_uncertainty(const SearchGrid& sg, const CellMap& cm_high, const float2* points,
                        const long long* offsets, const TableSlot* slots, const TableMeta* metas,
                        const PairWork& pw, const Rotation* rots, const SearchBuffers& sb,
                        double useless_threshold, double* costs, int dims, float* out,
                        cudaStream_t stream) {
  unc_cost_kernel<<<pw.n_rot, 256, 0, stream>>>(sg, cm_high, points, offsets, slots, metas, pw, rots,
                                                sb, useless_threshold, costs);
  unc_moments_kernel<<<1, 32, 0, stream>>>(sg, rots + pw.rot_begin, pw.n_rot, costs, dims, out);
}

}  // namespace csm
