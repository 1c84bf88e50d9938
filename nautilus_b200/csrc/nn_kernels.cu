// Point-to-point nearest-neighbour correspondences on sm_100a (K7).  Replaces
//   FindClosestPoint / GetPointNormal        (src/optimization/solver.cc:67-89)
//   the query loop of GetPointToPointMatching (src/optimization/solver.cc:146-170)
//   KDTree<float,2>::FindNearestPoint         (src/util/kdtree.cc:253-305)
//
// The reference answers every query by walking a pointer KD-tree.  That walk is
// NOT an exact nearest-neighbour search: a node with a single child (always a
// left leaf, kdtree.cc:131-147 with two values) never looks at that child when
// the query lies on its right (pd > 0 fails cc:270, right_tree_ == NULL fails
// cc:280, other_tree stays NULL for cc:293).  Apart from that the pruning of
// cc:292-293 never discards a closer point.  So the tree's answer is
//     argmin over { target points that are not such a skipped leaf for q }
// -- a per-(query, point) filter that needs only the leaf's parent (split
// dimension + split value, "skip" word of GridPoint), not the traversal.  The
// search itself therefore needs no tree: every target cloud is bucketed once
// into a hashed uniform grid whose cell is at least the threshold, and a query
// scans the 3 x 3 cells around it.  One block matches one (source, target)
// pair; the target's grid (bucket offsets + bucket-ordered points) is staged in
// shared memory with bulk asynchronous copies (TMA, cp.async.bulk) completing
// on an mbarrier.
//
// Distances are the reference's floats: sqrtf(dx*dx + dy*dy) with separately
// rounded products (Eigen's norm(); the library is built with --fmad=false) and
// strict '<' comparisons.  The argument above is exact except where a float
// comparison of the walk can come out equal; those queries are recognised
// conservatively and re-run the reference's own traversal over a device copy of
// the same tree (nn_walk), so the result is the tree's by construction:
//  (a) two different candidates at exactly the same float distance (the tree
//      keeps whichever it visits first);
//  (b) a nearest distance equal to (float)threshold when that float is below the
//      double threshold of solver.cc:85;
//  (c) a nearest distance that equals |dx| or |dy| in float: only then can
//      fabs(point_distance) < min(best, threshold) (cc:293) fail for a plane
//      that separates the query from its neighbour.
#include <float.h>

#include "csm_common.cuh"
#include "nn_kernels.h"

namespace csm {

// Grid cell of a coordinate: floor(v / cell) evaluated in double so that points
// closer than one cell always land in adjacent cells; clamped so +-1 is safe.
__device__ __forceinline__ int nn_cell(float v, double inv_cell) {
  const double c = floor(__dmul_rn((double)v, inv_cell));
  if (!(c > -1.0e9)) return -1000000000;  // also NaN
  if (c > 1.0e9) return 1000000000;
  return (int)c;
}

__device__ __forceinline__ int nn_bucket(int cx, int cy, int mask) {
  return (int)(((unsigned)cx * 73856093u) ^ ((unsigned)cy * 19349663u)) & mask;
}

// (value_.point - point).norm() (kdtree.cc:257): Eigen evaluates
// sqrt(d0*d0 + d1*d1) in float.
__device__ __forceinline__ float nn_dist(float tx, float ty, float2 q, float* adx, float* ady) {
  const float dx = __fsub_rn(tx, q.x), dy = __fsub_rn(ty, q.y);
  *adx = fabsf(dx); *ady = fabsf(dy);
  return __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
}
__device__ __forceinline__ float nn_dist(float2 t, float2 q) {
  float ax, ay;
  return nn_dist(t.x, t.y, q, &ax, &ay);
}

// ---------------------------------------------------------------------------
// Grid build: one block per scan.  Counting sort of the scan's points by hash
// bucket: histogram in shared memory, exclusive scan, scatter.
// ---------------------------------------------------------------------------
constexpr int NB_THREADS = 256;

__global__ void __launch_bounds__(NB_THREADS) nn_build_kernel(
    const float2* __restrict__ points, const long long* __restrict__ offsets,
    const long long* __restrict__ pad_off, const float* __restrict__ skip_val,
    const int* __restrict__ skip_dim, int n_buckets, double inv_cell,
    GridPoint* __restrict__ sorted_pts, int* __restrict__ bucket_start) {
  extern __shared__ int s_count[];  // [n_buckets] counts, then cursors
  __shared__ int s_scan[NB_THREADS];
  const int scan = blockIdx.x;
  const long long p0 = offsets[scan];
  const int n = (int)(offsets[scan + 1] - p0);
  const float2* pts = points + p0;
  const int mask = n_buckets - 1;
  for (int b = threadIdx.x; b < n_buckets; b += NB_THREADS) s_count[b] = 0;
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += NB_THREADS) {
    const float2 p = pts[i];
    atomicAdd(&s_count[nn_bucket(nn_cell(p.x, inv_cell), nn_cell(p.y, inv_cell), mask)], 1);
  }
  __syncthreads();
  // exclusive scan: each thread owns a contiguous run of buckets
  const int per = (n_buckets + NB_THREADS - 1) / NB_THREADS;
  const int b0 = threadIdx.x * per, b1 = min(n_buckets, b0 + per);
  int local = 0;
  for (int b = b0; b < b1; ++b) local += s_count[b];
  s_scan[threadIdx.x] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0;
    for (int t = 0; t < NB_THREADS; ++t) { const int v = s_scan[t]; s_scan[t] = run; run += v; }
  }
  __syncthreads();
  int* bs = bucket_start + (size_t)scan * (n_buckets + 4);
  int run = s_scan[threadIdx.x];
  for (int b = b0; b < b1; ++b) {
    const int c = s_count[b];
    bs[b] = run;
    s_count[b] = run;  // becomes the scatter cursor
    run += c;
  }
  if (threadIdx.x == 0) { bs[n_buckets] = n; bs[n_buckets + 1] = n; bs[n_buckets + 2] = n; bs[n_buckets + 3] = n; }
  __syncthreads();
  GridPoint* sp = sorted_pts + pad_off[scan];
  for (int i = threadIdx.x; i < n; i += NB_THREADS) {
    const float2 p = pts[i];
    const int b = nn_bucket(nn_cell(p.x, inv_cell), nn_cell(p.y, inv_cell), mask);
    const int at = atomicAdd(&s_count[b], 1);
    GridPoint g;
    g.x = p.x; g.y = p.y;
    g.skip_val = skip_val[p0 + i];
    g.idx_dim = i | (skip_dim[p0 + i] << 30);  // dim code: 0 none, 1 = x, 2 = y
    sp[at] = g;
  }
}

void launch_nn_build(const float2* points, const long long* offsets, const long long* pad_off,
                     const float* skip_val, const int* skip_dim, int n_scans, int n_buckets,
                     double inv_cell, GridPoint* sorted_pts, int* bucket_start,
                     cudaStream_t stream) {
  if (n_scans <= 0) return;
  const size_t smem = sizeof(int) * (size_t)n_buckets;
  static SmemOptIn opt;
  opt.allow(nn_build_kernel, smem);
  nn_build_kernel<<<n_scans, NB_THREADS, smem, stream>>>(points, offsets, pad_off, skip_val,
                                                         skip_dim, n_buckets, inv_cell,
                                                         sorted_pts, bucket_start);
}

// ---------------------------------------------------------------------------
// KDTree<float,2>::FindNearestPoint (kdtree.cc:253-305) as an explicit-stack
// state machine over the device copy of the tree.  Returns the distance; *out
// is the index (within the scan) of neighbor_node.
// ---------------------------------------------------------------------------
constexpr int NN_MAX_DEPTH = 48;

struct WalkFrame {
  int node, other, best_idx, stage;
  float thr, best, pd;
};

__device__ __noinline__ float nn_walk(const KdNode* __restrict__ nodes, const float2* __restrict__ pts,
                         float2 q, float threshold, int* out) {
  WalkFrame st[NN_MAX_DEPTH];
  int sp = 0;
  st[0].node = 0; st[0].thr = threshold; st[0].stage = 0;
  float ret_d = 0.f;
  int ret_i = 0;
  while (sp >= 0) {
    WalkFrame& f = st[sp];
    const KdNode nd = nodes[f.node];
    if (f.stage == 0) {
      const float2 v = pts[nd.point];
      f.best = nn_dist(v, q);             // cc:257
      f.best_idx = nd.point;              // cc:258
      if (f.best < FLT_MIN) {             // cc:259-261
        ret_d = 0.f; ret_i = nd.point; --sp;
        continue;
      }
      f.pd = __fsub_rn(nd.dim == 0 ? q.x : q.y, nd.dim == 0 ? v.x : v.y);  // cc:264-265
      f.other = -1;
      f.stage = 1;
      if (f.pd <= 0.0f && nd.left >= 0 && sp + 1 < NN_MAX_DEPTH) {  // cc:270-279
        f.other = nd.right;
        f.stage = 2;
        WalkFrame& c = st[++sp];
        c.node = nd.left; c.thr = fminf(f.best, f.thr); c.stage = 0;
        continue;
      }
    }
    if (f.stage == 2) {  // back from the left subtree
      if (ret_d < f.best) { f.best = ret_d; f.best_idx = ret_i; }
      f.stage = 1;
    }
    if (f.stage == 1) {
      f.stage = 3;
      if (f.pd >= 0.0f && nd.right >= 0 && sp + 1 < NN_MAX_DEPTH) {  // cc:280-289
        f.other = nd.left;
        f.stage = 4;
        WalkFrame& c = st[++sp];
        c.node = nd.right; c.thr = fminf(f.best, f.thr); c.stage = 0;
        continue;
      }
    }
    if (f.stage == 4) {  // back from the right subtree
      if (ret_d < f.best) { f.best = ret_d; f.best_idx = ret_i; }
      f.stage = 3;
    }
    if (f.stage == 3) {
      f.stage = 5;
      if (f.other >= 0 && f.pd != 0.0f && fabsf(f.pd) < fminf(f.best, f.thr) &&
          sp + 1 < NN_MAX_DEPTH) {  // cc:293-302
        f.stage = 6;
        WalkFrame& c = st[++sp];
        c.node = f.other; c.thr = fminf(f.best, f.thr); c.stage = 0;
        continue;
      }
    }
    if (f.stage == 6) {  // back from the far subtree
      if (ret_d < f.best) { f.best = ret_d; f.best_idx = ret_i; }
      f.stage = 5;
    }
    // stage 5: return current_best_dist (cc:304)
    ret_d = f.best; ret_i = f.best_idx; --sp;
  }
  *out = ret_i;
  return ret_d;
}

// ---------------------------------------------------------------------------
// K7  match: one block per pair.
// ---------------------------------------------------------------------------
constexpr int NM_THREADS = 128;

__device__ __forceinline__ unsigned smem_u32(const void* p) {
  return (unsigned)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src_gmem, unsigned bytes,
                                          unsigned long long* mbar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
      :: "r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(mbar)) : "memory");
}

size_t nn_match_smem_bytes(int n_buckets, int max_points) {
  return 16 + sizeof(int) * ((size_t)n_buckets + 4) + (size_t)max_points * sizeof(GridPoint);
}

template <bool STAGED>
__global__ void __launch_bounds__(NM_THREADS) nn_match_kernel(
    NnIndex ix, const NnPair* __restrict__ pairs, float thr_f, double thr_d,
    int* __restrict__ match_idx, float* __restrict__ match_dist,
    unsigned long long* __restrict__ counters) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  const NnPair pr = pairs[blockIdx.x];
  const long long s0 = ix.offsets[pr.src];
  const int n_src = (int)(ix.offsets[pr.src + 1] - s0);
  const long long t0 = ix.offsets[pr.tgt];
  const int n_tgt = (int)(ix.offsets[pr.tgt + 1] - t0);
  const int nb = ix.n_buckets, mask = nb - 1;
  const int* g_bs = ix.bucket_start + (size_t)pr.tgt * (nb + 4);
  const GridPoint* g_pts = ix.sorted_pts + ix.pad_off[pr.tgt];

  const int* bs = g_bs;
  const GridPoint* tp = g_pts;
  if (STAGED && n_tgt > 0) {
    unsigned long long* mbar = reinterpret_cast<unsigned long long*>(s_raw);
    int* s_bs = reinterpret_cast<int*>(s_raw + 16);
    GridPoint* s_pts = reinterpret_cast<GridPoint*>(s_bs + nb + 4);
    if (threadIdx.x == 0) {
      const unsigned b_bs = sizeof(int) * (nb + 4), b_pts = sizeof(GridPoint) * n_tgt;
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" :: "r"(smem_u32(mbar)));
      asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n"
                   :: "r"(smem_u32(mbar)), "r"(b_bs + b_pts) : "memory");
      bulk_load(s_bs, g_bs, b_bs, mbar);
      bulk_load(s_pts, g_pts, b_pts, mbar);
    }
    __syncthreads();  // the barrier is initialised before anyone polls it
    unsigned done = 0;
    while (!done) {
      asm volatile(
          "{\n.reg .pred p;\n"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n"
          "selp.u32 %0, 1, 0, p;\n}\n"
          : "=r"(done) : "r"(smem_u32(mbar)) : "memory");
    }
    bs = s_bs; tp = s_pts;
  }

  const float2* src = ix.points + s0;
  const float2* tgt_raw = ix.points + t0;
  const KdNode* nodes = ix.nodes + t0;
  unsigned long long walked = 0;
  for (int i = threadIdx.x; i < n_src; i += NM_THREADS) {
    const float2 p = src[i];
    // Affine2f * Vector2f: linear * p + translation, row by row
    float2 q;
    q.x = __fadd_rn(__fadd_rn(__fmul_rn(pr.m[0], p.x), __fmul_rn(pr.m[1], p.y)), pr.m[2]);
    q.y = __fadd_rn(__fadd_rn(__fmul_rn(pr.m[3], p.x), __fmul_rn(pr.m[4], p.y)), pr.m[5]);
    int best_i = -1;
    float best2 = FLT_MAX, second2 = FLT_MAX, best_ax = 0.f, best_ay = 0.f;
    if (n_tgt > 0) {
      const int cx = nn_cell(q.x, ix.inv_cell), cy = nn_cell(q.y, ix.inv_cell);
      // Squared distances in the loop (sqrt is monotone: the smallest square has the
      // smallest float distance; only the runner-up can tie with it after rounding).
      // Cells are probed centre first; a neighbour cell whose nearest edge is farther
      // than min(best, threshold) cannot hold the answer or a tie and is skipped
      // (margin 3e-6 relative on the squares: float rounding of dx, dy and the sum is
      // below 3e-7).  r2 = that radius squared, in double.
      const double cell = 1.0 / ix.inv_cell;
      const double gx0 = (double)q.x - (double)cx * cell, gx1 = (double)(cx + 1) * cell - (double)q.x;
      const double gy0 = (double)q.y - (double)cy * cell, gy1 = (double)(cy + 1) * cell - (double)q.y;
      double r2 = (double)thr_f * (double)thr_f;
      // (rolled on purpose: the unrolled 3 x 3 probe overflows the instruction cache)
      // probe order: centre, edge neighbours, corners; 2-bit fields hold ox + 1 / oy + 1
      constexpr unsigned OX = 1u | (0u << 2) | (2u << 4) | (1u << 6) | (1u << 8) | (0u << 10) | (2u << 12) | (0u << 14) | (2u << 16);
      constexpr unsigned OY = 1u | (1u << 2) | (1u << 4) | (0u << 6) | (2u << 8) | (0u << 10) | (0u << 12) | (2u << 14) | (2u << 16);
#pragma unroll 1
      for (int pk = 0; pk < 9; ++pk) {
        const int ox = (int)((OX >> (2 * pk)) & 3u) - 1, oy = (int)((OY >> (2 * pk)) & 3u) - 1;
        if (pk > 0) {
          const double gx = ox < 0 ? gx0 : ox > 0 ? gx1 : 0.0;
          const double gy = oy < 0 ? gy0 : oy > 0 ? gy1 : 0.0;
          const double g = fmax(gx, gy);
          if (g > 0.0 && g * g > r2 * 1.000003) continue;
        }
        const int b = nn_bucket(cx + ox, cy + oy, mask);
        const int k1 = bs[b + 1];
        for (int k = bs[b]; k < k1; ++k) {
          const GridPoint g = tp[k];
          const int dim = (unsigned)g.idx_dim >> 30, id = g.idx_dim & 0x3fffffff;
          // a leaf the walk never reaches from this side of its parent
          if (dim != 0 && (dim == 1 ? q.x : q.y) > g.skip_val) continue;
          const float dx = __fsub_rn(g.x, q.x), dy = __fsub_rn(g.y, q.y);
          const float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
          if (d2 < best2) {
            second2 = best2; best2 = d2; best_i = id;
            best_ax = fabsf(dx); best_ay = fabsf(dy);
            r2 = fmin(r2, (double)d2);
          } else if (id != best_i && d2 < second2) {
            second2 = d2;
          }
        }
      }
    }
    // the reference's float distance of the winner, and whether the runner-up ties with it
    const float best = best_i >= 0 ? __fsqrt_rn(best2) : FLT_MAX;
    const bool tie = best_i >= 0 && second2 < FLT_MAX && __fsqrt_rn(second2) == best;
    int out = -1;
    float out_d = best;
    if (best_i >= 0) {
      const bool below = (double)best < thr_d;   // solver.cc:85 compares in double
      const bool edge = (best == thr_f) && ((double)thr_f < thr_d);
      // (a zero distance makes every separating plane pass through the query: both sides are walked)
      const bool on_axis = below && best > 0.f && (best == best_ax || best == best_ay);
      if ((below && tie) || edge || on_axis) {
        int wi;
        const float wd = nn_walk(nodes, tgt_raw, q, thr_f, &wi);
        ++walked;
        if ((double)wd < thr_d) { out = wi; out_d = wd; }
      } else if (below) {
        out = best_i;
      }
    }
    match_idx[pr.out + i] = out;
    if (match_dist) match_dist[pr.out + i] = out_d;
  }
  if (walked && counters) atomicAdd(counters, walked);
}

// ---------------------------------------------------------------------------
// Resident correspondences: the matched source points of a pair, compacted in
// source order (the order Solver::GetPointToPointMatching appends them,
// solver.cc:146-170).  One block per pair; ballot + running count keep the order.
// ---------------------------------------------------------------------------
constexpr int CC_THREADS = 128;

__global__ void __launch_bounds__(CC_THREADS) corr_count_kernel(
    NnIndex ix, const NnPair* __restrict__ pairs, const int* __restrict__ match_idx,
    int* __restrict__ cnt) {
  const NnPair pr = pairs[blockIdx.x];
  const int n = (int)(ix.offsets[pr.src + 1] - ix.offsets[pr.src]);
  const int* m = match_idx + pr.out;
  int c = 0;
  for (int i = threadIdx.x; i < n; i += CC_THREADS) c += m[i] >= 0;
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  __shared__ int s_c[CC_THREADS / 32];
  if ((threadIdx.x & 31) == 0) s_c[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < CC_THREADS / 32; ++w) t += s_c[w];
    cnt[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(CC_THREADS) corr_write_kernel(
    NnIndex ix, const NnPair* __restrict__ pairs, const int* __restrict__ match_idx,
    const int* __restrict__ self_idx, const long long* __restrict__ corr_off,
    int4* __restrict__ corr) {
  const NnPair pr = pairs[blockIdx.x];
  const long long so = ix.offsets[pr.src], to = ix.offsets[pr.tgt];
  const int n = (int)(ix.offsets[pr.src + 1] - so);
  const int* m = match_idx + pr.out;
  int4* out = corr + corr_off[blockIdx.x];
  __shared__ int s_w[CC_THREADS / 32];
  __shared__ int s_base;
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i0 = 0; i0 < n; i0 += CC_THREADS) {
    const int i = i0 + threadIdx.x;
    const int t = i < n ? m[i] : -1;
    const unsigned b = __ballot_sync(0xffffffffu, t >= 0);
    if (lane == 0) s_w[warp] = __popc(b);
    __syncthreads();
    int before = s_base;
    for (int w = 0; w < warp; ++w) before += s_w[w];
    if (t >= 0) {
      const int k = before + __popc(b & ((1u << lane) - 1u));
      const long long si = so + i, ti = to + t;
      // the normal of a point is the one stored with the tree node GetPointNormal
      // finds for it (solver.cc:67-77): self_idx
      out[k] = make_int4((int)si, (int)ti, (int)(so + self_idx[si]), (int)(to + self_idx[ti]));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int t_all = 0;
      for (int w = 0; w < CC_THREADS / 32; ++w) t_all += s_w[w];
      s_base += t_all;
    }
    __syncthreads();
  }
}

void launch_corr_count(const NnIndex& ix, const NnPair* pairs, int n_pairs, const int* match_idx,
                       int* cnt, cudaStream_t stream) {
  if (n_pairs > 0) corr_count_kernel<<<n_pairs, CC_THREADS, 0, stream>>>(ix, pairs, match_idx, cnt);
}

void launch_corr_write(const NnIndex& ix, const NnPair* pairs, int n_pairs, const int* match_idx,
                       const int* self_idx, const long long* corr_off, int4* corr,
                       cudaStream_t stream) {
  if (n_pairs > 0)
    corr_write_kernel<<<n_pairs, CC_THREADS, 0, stream>>>(ix, pairs, match_idx, self_idx, corr_off,
                                                          corr);
}

void launch_nn_match(const NnIndex& ix, const NnPair* pairs, int n_pairs, int max_tgt_points,
                     float thr_f, double thr_d, int* match_idx, float* match_dist,
                     unsigned long long* counters, cudaStream_t stream) {
  if (n_pairs <= 0) return;
  const size_t smem = nn_match_smem_bytes(ix.n_buckets, max_tgt_points);
  if (smem <= 200 * 1024) {
    static SmemOptIn opt;
    opt.allow(nn_match_kernel<true>, smem);
    nn_match_kernel<true><<<n_pairs, NM_THREADS, smem, stream>>>(ix, pairs, thr_f, thr_d, match_idx,
                                                                 match_dist, counters);
  } else {  // clouds too large to stage: read the grid through L1/L2
    nn_match_kernel<false><<<n_pairs, NM_THREADS, 16, stream>>>(ix, pairs, thr_f, thr_d, match_idx,
                                                                match_dist, counters);
  }
}

}  // namespace csm
