// Multi-resolution (theta, x, y) search on sm_100a.  Replaces
//   RotatePointcloud / CalculatePointcloudCost   (CorrelativeScanMatcher.cc:55-77)
//   MemoizeLowRes                                (CorrelativeScanMatcher.cc:84-115)
//   GetProbAndTransformation                     (CorrelativeScanMatcher.cc:117-143)
//   the best-first loop of GetTransformation     (CorrelativeScanMatcher.cc:170-243)
//
// Two facts shape the kernels:
//  * the cell a translated point falls in is separable: the x cell depends on
//    (point, theta, x-translation) only, the y cell on (point, theta,
//    y-translation) only.  Each block therefore converts a chunk of rotated
//    points once per axis (n_t exact double divisions per point and axis, the
//    reference's floor(x / res)) into shared-memory index lists, and every one
//    of the n_t*n_t translations then costs one table gather + one add per point.
//  * log(max(v, 1e-10)) is a function of the table cell, so it is taken once
//    per low-res cell when the table is pooled (lowlog) instead of once per
//    (candidate, point).
// Scores are summed over points in point order, one thread per candidate, like
// the reference's sequential loop; identical gathered values give identical sums.
#include "csm_common.cuh"
#include "csm_kernels.h"

namespace csm {

// Rotation2Df(theta).toRotationMatrix() * p with Eigen's evaluation order
// (c*x + (-s)*y ; s*x + c*y), each operation rounded separately.
__device__ __forceinline__ float2 rotate_point(float2 p, float c, float s) {
  float2 q;
  q.x = __fadd_rn(__fmul_rn(c, p.x), __fmul_rn(-s, p.y));
  q.y = __fadd_rn(__fmul_rn(s, p.x), __fmul_rn(c, p.y));
  return q;
}

__host__ __device__ inline int padded_chunk(int point_chunk) {
  int pc = (point_chunk + 1) & ~1;      // even number of uint16 per row
  if (((pc >> 1) & 1) == 0) pc += 2;    // odd number of 32-bit words per row
  return pc;
}

size_t low_score_smem_bytes(int n_t, int point_chunk) {
  const int pcp = padded_chunk(point_chunk);
  size_t idx = (size_t)2 * n_t * pcp * sizeof(unsigned short);
  idx = (idx + 7) & ~(size_t)7;
  return idx + (size_t)n_t * n_t * (2 * sizeof(double) + sizeof(int));
}

// ---------------------------------------------------------------------------
// K5  low-res exhaustive scoring + per-translation max over a theta chunk.
// grid = (theta chunks, pairs).
// ---------------------------------------------------------------------------
constexpr int LS_THREADS = 256;

__global__ void __launch_bounds__(LS_THREADS) low_score_kernel(
    Geometry g, SearchGrid sg, const float2* __restrict__ points,
    const long long* __restrict__ offsets, const TableSlot* __restrict__ slots,
    const PairWork* __restrict__ pairs, const Rotation* __restrict__ rots,
    SearchBuffers sb, int point_chunk) {
  const PairWork pw = pairs[blockIdx.y];
  const int rot0 = blockIdx.x * sb.rot_per_chunk;
  if (rot0 >= pw.n_rot) return;
  const int rot1 = min(pw.n_rot, rot0 + sb.rot_per_chunk);
  const int n_t = sg.n_t, n_tr = n_t * n_t;
  const int pcp = padded_chunk(point_chunk);

  extern __shared__ __align__(16) unsigned char s_raw[];
  unsigned short* cxs = reinterpret_cast<unsigned short*>(s_raw);  // [n_t][pcp]
  unsigned short* cys = cxs + (size_t)n_t * pcp;                   // [n_t][pcp]
  size_t off = ((size_t)2 * n_t * pcp * sizeof(unsigned short) + 7) & ~(size_t)7;
  double* acc = reinterpret_cast<double*>(s_raw + off);            // [n_tr]
  double* best = acc + n_tr;                                       // [n_tr]
  int* best_rot = reinterpret_cast<int*>(best + n_tr);             // [n_tr]

  const float2* a = points + offsets[pw.a_scan];
  const int n_pts = (int)(offsets[pw.a_scan + 1] - offsets[pw.a_scan]);
  const double* __restrict__ lowlog = slots[pw.table].lowlog;
  const int wl = g.wl, lstride = g.wl + 1;

  for (int tr = threadIdx.x; tr < n_tr; tr += LS_THREADS) {
    best[tr] = -INFINITY;
    best_rot[tr] = 0;
  }

  for (int r = rot0; r < rot1; ++r) {
    const Rotation rot = rots[pw.rot_begin + r];
    for (int tr = threadIdx.x; tr < n_tr; tr += LS_THREADS) acc[tr] = 0.0;
    for (int p0 = 0; p0 < n_pts; p0 += point_chunk) {
      const int pc = min(point_chunk, n_pts - p0);
      __syncthreads();
      // phase A: cell lists of this point chunk
      for (int pl = threadIdx.x; pl < pc; pl += LS_THREADS) {
        const float2 q = rotate_point(a[p0 + pl], rot.c, rot.s);
        for (int i = 0; i < n_t; ++i) {
          const float t = sg.t_low[i];  // Vector2f(x_trans, y_trans): narrowed to float
          const long long cx = cell_of(__fadd_rn(q.x, t), g.res_l, wl);
          const long long cy = cell_of(__fadd_rn(q.y, t), g.res_l, wl);
          cxs[i * pcp + pl] = (unsigned short)((cx >= 0 && cx < wl) ? cx : wl);
          cys[i * pcp + pl] = (unsigned short)((cy >= 0 && cy < wl) ? cy : wl);
        }
      }
      __syncthreads();
      // phase B: one thread per translation, x index fastest across lanes so a
      // warp's gathers fall on a few table rows
      for (int tq = threadIdx.x; tq < n_tr; tq += LS_THREADS) {
        const int j = tq / n_t, i = tq - j * n_t;
        const unsigned short* px = cxs + i * pcp;
        const unsigned short* py = cys + j * pcp;
        double s = acc[i * n_t + j];
        for (int pl = 0; pl < pc; ++pl) {
          // an out-of-grid x or y must read the floor: row wl and column wl of
          // lowlog both hold log(1e-10)
          const int cy = py[pl], cx = px[pl];
          s = __dadd_rn(s, __ldg(lowlog + cy * lstride + cx));
        }
        acc[i * n_t + j] = s;
      }
    }
    __syncthreads();
    for (int tr = threadIdx.x; tr < n_tr; tr += LS_THREADS) {
      const double score = __ddiv_rn(acc[tr], (double)n_pts);  // probability / size (cc:76)
      if (score > best[tr]) {  // std::max keeps the earlier entry on ties (cc:109-110)
        best[tr] = score;
        best_rot[tr] = r;
      }
    }
  }
  __syncthreads();
  const size_t base = ((size_t)blockIdx.y * sb.chunks_per_pair + blockIdx.x) * n_tr;
  for (int tr = threadIdx.x; tr < n_tr; tr += LS_THREADS) {
    sb.chunk_best[base + tr] = best[tr];
    sb.chunk_rot[base + tr] = best_rot[tr];
  }
}

void launch_low_score(const Geometry& g, const SearchGrid& sg, const float2* points,
                      const long long* offsets, const TableSlot* slots, const PairWork* pairs,
                      const Rotation* rots, int n_pairs, const SearchBuffers& sb,
                      int point_chunk, size_t smem_bytes, cudaStream_t stream) {
  if (n_pairs <= 0) return;
  static size_t configured = 0;
  if (smem_bytes > 48 * 1024 && smem_bytes > configured) {
    cudaFuncSetAttribute(low_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int)smem_bytes);
    configured = smem_bytes;
  }
  dim3 grid(sb.chunks_per_pair, n_pairs);
  low_score_kernel<<<grid, LS_THREADS, smem_bytes, stream>>>(g, sg, points, offsets, slots,
                                                              pairs, rots, sb, point_chunk);
}

// ---------------------------------------------------------------------------
// K6  memo reduction + best-first hi-res refinement, one block per pair.
// Follows GetTransformation's loop (CorrelativeScanMatcher.cc:170-243):
//   pop the first maximum of the memo in slot (AbsCoords) order, invalidate
//   it, stop when it is below the best hi-res score so far, otherwise score
//   the hi-res window it opens and keep a strictly better result.
// ---------------------------------------------------------------------------
constexpr int RF_THREADS = 128;
constexpr int RF_POINT_CHUNK = 128;

struct ArgBest {
  double score;
  int key;   // tie-break: smaller wins
  int idx;
};

__device__ __forceinline__ ArgBest arg_better(ArgBest a, ArgBest b) {
  if (b.score > a.score || (b.score == a.score && b.key < a.key)) return b;
  return a;
}

__device__ ArgBest block_argmax(ArgBest v, ArgBest* s_red) {
  for (int o = 16; o > 0; o >>= 1) {
    ArgBest w;
    w.score = __shfl_xor_sync(0xffffffffu, v.score, o);
    w.key = __shfl_xor_sync(0xffffffffu, v.key, o);
    w.idx = __shfl_xor_sync(0xffffffffu, v.idx, o);
    v = arg_better(v, w);
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
  __syncthreads();
  ArgBest r = s_red[0];
  for (int k = 1; k < (int)(blockDim.x >> 5); ++k) r = arg_better(r, s_red[k]);
  return r;
}

__global__ void __launch_bounds__(RF_THREADS) refine_kernel(
    Geometry g, SearchGrid sg, const float2* __restrict__ points,
    const long long* __restrict__ offsets, const TableSlot* __restrict__ slots,
    const TableMeta* __restrict__ metas, const PairWork* __restrict__ pairs,
    const Rotation* __restrict__ rots, SearchBuffers sb, PairResult* __restrict__ results) {
  const PairWork pw = pairs[blockIdx.x];
  const int n_t = sg.n_t, n_tr = n_t * n_t;
  const int nh_max = sg.n_h_max;

  extern __shared__ __align__(16) unsigned char s_raw[];
  double* memo = reinterpret_cast<double*>(s_raw);                 // [n_tr]
  double* hacc = memo + n_tr;                                      // [nh_max^2]
  int* memo_rot = reinterpret_cast<int*>(hacc + nh_max * nh_max);  // [n_tr]
  int* cxh = memo_rot + n_tr;                                      // [nh_max][RF_POINT_CHUNK]
  int* cyh = cxh + nh_max * RF_POINT_CHUNK;                        // row offset or -1
  __shared__ ArgBest s_red[RF_THREADS / 32];
  __shared__ int s_ctl[4];

  const float2* a = points + offsets[pw.a_scan];
  const int n_pts = (int)(offsets[pw.a_scan + 1] - offsets[pw.a_scan]);
  const TableMeta tm = metas[pw.table];
  const double* __restrict__ hi = slots[pw.table].values;
  const int n_chunks = (pw.n_rot + sb.rot_per_chunk - 1) / sb.rot_per_chunk;

  // memo[slot] = max over theta, earliest theta on ties (chunks ascend in theta)
  for (int tr = threadIdx.x; tr < n_tr; tr += RF_THREADS) {
    double b = -INFINITY;
    int br = 0;
    for (int ch = 0; ch < n_chunks; ++ch) {
      const size_t at = ((size_t)blockIdx.x * sb.chunks_per_pair + ch) * n_tr + tr;
      const double s = sb.chunk_best[at];
      if (s > b) { b = s; br = sb.chunk_rot[at]; }
    }
    memo[tr] = b;
    memo_rot[tr] = br;
  }
  __syncthreads();

  double best_prob = -INFINITY;   // meaningful in thread 0 only
  float best_x = 0.f, best_y = 0.f, best_theta = 0.f;
  long long n_high = 0, n_pops = 0;
  int status = 0;

  for (;;) {
    // std::max_element over the memo: first maximum in slot order (cc:172-174)
    ArgBest mine = {-INFINITY, INT_MAX, -1};
    for (int tr = threadIdx.x; tr < n_tr; tr += RF_THREADS) {
      ArgBest c = {memo[tr], sg.slot_rank[tr], tr};
      mine = arg_better(mine, c);
    }
    const ArgBest top = block_argmax(mine, s_red);
    if (threadIdx.x == 0) {
      int stop = 0;
      if (top.idx < 0 || top.score < best_prob) stop = 1;
      else if (top.score == -INFINITY && best_prob == -INFINITY) { stop = 1; status = 1; }
      if (!stop) { memo[top.idx] = -INFINITY; ++n_pops; }
      s_ctl[0] = stop;
    }
    __syncthreads();
    if (s_ctl[0]) break;

    const int i = top.idx / n_t, j = top.idx - i * n_t;  // tr = i*n_t + j, x outer
    const Rotation rot = rots[pw.rot_begin + memo_rot[top.idx]];
    const int nhx = sg.n_h[i], nhy = sg.n_h[j], n_cand = nhx * nhy;
    const float* thx = sg.t_high + i * nh_max;
    const float* thy = sg.t_high + j * nh_max;

    for (int c = threadIdx.x; c < n_cand; c += RF_THREADS) hacc[c] = 0.0;
    for (int p0 = 0; p0 < n_pts; p0 += RF_POINT_CHUNK) {
      const int pc = min(RF_POINT_CHUNK, n_pts - p0);
      __syncthreads();
      for (int pl = threadIdx.x; pl < pc; pl += RF_THREADS) {
        const float2 q = rotate_point(a[p0 + pl], rot.c, rot.s);
        for (int k = 0; k < nhx; ++k) {
          const long long cx = cell_of(__fadd_rn(q.x, thx[k]), g.res_h, g.w);
          cxh[k * RF_POINT_CHUNK + pl] = (cx >= tm.ax0 && cx <= tm.ax1) ? (int)cx - tm.ax0 : -1;
        }
        for (int k = 0; k < nhy; ++k) {
          const long long cy = cell_of(__fadd_rn(q.y, thy[k]), g.res_h, g.w);
          cyh[k * RF_POINT_CHUNK + pl] =
              (cy >= tm.ay0 && cy <= tm.ay1) ? ((int)cy - tm.ay0) * tm.stride : -1;
        }
      }
      __syncthreads();
      for (int c = threadIdx.x; c < n_cand; c += RF_THREADS) {
        const int ky = c % nhy, kx = c / nhy;   // x outer, y inner (cc:126-129)
        const int* px = cxh + kx * RF_POINT_CHUNK;
        const int* py = cyh + ky * RF_POINT_CHUNK;
        double s = hacc[c];
        for (int pl = 0; pl < pc; ++pl) {
          const int ox = px[pl], oy = py[pl];
          double v = CSM_MIN_VALUE_FOR_LOOKUP;
          if ((ox | oy) >= 0) {
            const double t = __ldg(hi + oy + ox);
            if (t > CSM_MIN_VALUE_FOR_LOOKUP) v = t;
          }
          s = __dadd_rn(s, log(v));
        }
        hacc[c] = s;
      }
    }
    __syncthreads();
    // strict '>' in x-outer / y-inner order == max score, smallest index on ties
    ArgBest hb = {-INFINITY, INT_MAX, -1};
    for (int c = threadIdx.x; c < n_cand; c += RF_THREADS) {
      ArgBest cnd = {__ddiv_rn(hacc[c], (double)n_pts), c, c};
      hb = arg_better(hb, cnd);
    }
    const ArgBest htop = block_argmax(hb, s_red);
    if (threadIdx.x == 0) {
      n_high += n_cand;
      if (htop.idx >= 0 && htop.score > best_prob) {
        best_prob = htop.score;
        best_x = thx[htop.idx / nhy];
        best_y = thy[htop.idx % nhy];
        best_theta = rot.theta;
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    PairResult r;
    r.score = best_prob; r.x = best_x; r.y = best_y; r.theta = best_theta;
    r.pad = status;
    r.n_low = (long long)pw.n_rot * n_tr;
    r.n_high = n_high; r.n_pops = n_pops;
    results[pw.out_index] = r;
  }
}

static size_t refine_smem_bytes(int n_t, int nh_max) {
  return (size_t)n_t * n_t * (sizeof(double) + sizeof(int)) +
         (size_t)nh_max * nh_max * sizeof(double) +
         (size_t)2 * nh_max * RF_POINT_CHUNK * sizeof(int);
}

void launch_refine(const Geometry& g, const SearchGrid& sg, double trans_range,
                   const float2* points, const long long* offsets, const TableSlot* slots,
                   const TableMeta* metas, const PairWork* pairs, const Rotation* rots,
                   int n_pairs, const SearchBuffers& sb, PairResult* results,
                   cudaStream_t stream) {
  (void)trans_range;
  if (n_pairs <= 0) return;
  const size_t smem = refine_smem_bytes(sg.n_t, sg.n_h_max);
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    cudaFuncSetAttribute(refine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = smem;
  }
  refine_kernel<<<n_pairs, RF_THREADS, smem, stream>>>(g, sg, points, offsets, slots, metas,
                                                       pairs, rots, sb, results);
}

}  // namespace csm
