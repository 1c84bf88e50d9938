// Multi-resolution (theta, x, y) search on sm_100a.  Replaces
//   RotatePointcloud / CalculatePointcloudCost   (CorrelativeScanMatcher.cc:55-77)
//   MemoizeLowRes                                (CorrelativeScanMatcher.cc:84-115)
//   GetProbAndTransformation                     (CorrelativeScanMatcher.cc:117-143)
//   the best-first loop of GetTransformation     (CorrelativeScanMatcher.cc:170-243)
//
// Two facts shape the kernels:
//  * the cell a translated point falls in is separable: the x cell depends on
//    (point, theta, x-translation) only, the y cell on (point, theta,
//    y-translation) only.  Rotated points are therefore converted once per axis
//    into shared-memory cell lists (exact cells through float thresholds, see
//    CellMap), and every one of the n_t*n_t translations then costs one table
//    gather + one add per point.
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

// ---------------------------------------------------------------------------
// K5  low-res exhaustive scoring + per-translation max over a theta chunk.
// grid = (theta chunks, pairs), 8 warps per block.
//
// A warp scores one (rotation, translation tile) work item: a tile is 16
// x-translations (lane & 15) by 14 y-translations (7 per lane, lane >> 4 picks
// the half), so the reference's 14 x 14 grid is exactly one tile and every lane
// keeps 7 running sums in registers.  Per chunk of 32 points the warp first
// turns the rotated points into per-axis cell lists (lane = point; exact cells
// through the float threshold map, no division), then every lane adds its 7
// table values per point, in point order.  The part of the pooled log table
// that lies above the lookup floor (TableMeta::lx0..ly1) is staged in shared
// memory once per block; a half-warp reads 14 consecutive doubles of one table
// row, which is conflict-free.  Tables whose box does not fit fall back to
// gathering from global memory through L1.
// ---------------------------------------------------------------------------
constexpr int LS_WARPS = 8;
constexpr int LS_THREADS = LS_WARPS * 32;
constexpr int LS_PC = 32;            // points per chunk = one per lane in phase A
constexpr int LS_PCP = LS_PC + 1;    // odd row stride: rows start on different banks
constexpr int LS_TI = 16;            // x translations per tile
constexpr int LS_JT = 7;             // y translations per lane
constexpr int LS_TJ = 2 * LS_JT;     // y translations per tile
constexpr int LS_TILE = LS_TI * LS_TJ;
constexpr int LS_TABLE_CELLS = 7680; // shared-memory budget of the staged table (60 KB)
// per-warp cell lists: LS_TI rows of ints + 2 * LS_PC row records (32 B each)
constexpr int LS_LIST_INTS = LS_TI * LS_PCP + 2 * LS_PC * 8;

size_t low_score_smem_bytes(int n_t, int wl) {
  size_t b = sizeof(double) * LS_TABLE_CELLS;                       // staged table
  b += sizeof(double) * LS_WARPS * LS_TILE;                          // per-round results
  b += sizeof(double) * (size_t)n_t * n_t;                           // best score
  b += sizeof(int) * (size_t)n_t * n_t;                              // best rotation
  b += sizeof(int) * LS_WARPS * LS_LIST_INTS;                        // cell lists
  b += sizeof(float) * ((size_t)wl + 2 + n_t);                       // thresholds, translations
  return b;
}

// Row offsets (in table cells) of a point for the 7 y-translations of one
// half-tile; two 16-byte shared loads fetch them.
struct __align__(16) RowList {
  int r[8];
};

template <bool SMEM>
__device__ __forceinline__ void low_score_rounds(
    const Geometry& g, const SearchGrid& sg, const CellMap& cm, const float* s_tlow,
    const double* __restrict__ tab, int ox, int oy, int tw, int th, int ts,
    const float2* __restrict__ a, int n_pts, const Rotation* __restrict__ rots, int rot0,
    int rot1, int* s_lists, double* s_res, double* s_best, int* s_best_rot) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_t = sg.n_t;
  const int tiles_x = (n_t + LS_TI - 1) / LS_TI, tiles_y = (n_t + LS_TJ - 1) / LS_TJ;
  const int n_tiles = tiles_x * tiles_y;
  const int n_items = (rot1 - rot0) * n_tiles;
  int* lx = s_lists + warp * LS_LIST_INTS;                       // [LS_TI][LS_PCP] box column
  RowList* ly = reinterpret_cast<RowList*>(lx + LS_TI * LS_PCP);  // [2][LS_PC] row offsets
  const int li = lane & 15, half = lane >> 4, lj = half * LS_JT;
  const bool single = (n_tiles == 1);  // the reference's 14 x 14 grid: best stays in registers
  double rbest[LS_JT];
  int rbest_rot[LS_JT];
#pragma unroll
  for (int u = 0; u < LS_JT; ++u) { rbest[u] = -INFINITY; rbest_rot[u] = 0; }

  for (int round0 = 0; round0 < n_items; round0 += LS_WARPS) {
    const int item = round0 + warp;
    if (item < n_items) {
      const int r = rot0 + item / n_tiles, tile = item % n_tiles;
      const int ti = tile % tiles_x, tj = tile / tiles_x;
      const bool lane_on = ti * LS_TI + li < n_t;  // lanes past the grid do not gather
      const Rotation rot = rots[r];
      double acc[LS_JT];
#pragma unroll
      for (int u = 0; u < LS_JT; ++u) acc[u] = 0.0;
      for (int p0 = 0; p0 < n_pts; p0 += LS_PC) {
        const int pc = min(LS_PC, n_pts - p0);
        if (lane < pc) {  // phase A: this lane's point -> cells for the tile's translations
          const float2 q = rotate_point(a[p0 + lane], rot.c, rot.s);
#pragma unroll 4
          for (int k = 0; k < LS_TI; ++k) {
            const int ii = ti * LS_TI + k;
            int v = tw;  // sentinel column: reads the floor
            if (ii < n_t) {
              const int cx = cell_fast(cm, __fadd_rn(q.x, s_tlow[ii])) - ox;  // Vector2f(x_trans, .)
              if ((unsigned)cx < (unsigned)tw) v = cx;
            }
            lx[k * LS_PCP + lane] = v;
          }
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            RowList pr;
#pragma unroll
            for (int u = 0; u < LS_JT; ++u) {
              const int jj = tj * LS_TJ + hh * LS_JT + u;
              int v = th;  // sentinel row
              if (jj < n_t) {
                const int cy = cell_fast(cm, __fadd_rn(q.y, s_tlow[jj])) - oy;
                if ((unsigned)cy < (unsigned)th) v = cy;
              }
              pr.r[u] = v * ts;
            }
            pr.r[7] = 0;
            ly[hh * LS_PC + lane] = pr;
          }
        }
        __syncwarp();
        if (lane_on) {
          const int* px = lx + li * LS_PCP;
          const RowList* py = ly + half * LS_PC;
#pragma unroll 2
          for (int pl = 0; pl < pc; ++pl) {  // phase B: 7 gathers + adds per lane, point order
            const int cxv = px[pl];
            const RowList pr = py[pl];
#pragma unroll
            for (int u = 0; u < LS_JT; ++u) {
              const int at = pr.r[u] + cxv;
              double t;
              if (SMEM) {
                extern __shared__ __align__(16) unsigned char s_raw[];  // staged table at offset 0
                t = reinterpret_cast<const double*>(s_raw)[at];
              } else {
                t = __ldg(tab + at);
              }
              acc[u] = __dadd_rn(acc[u], t);
            }
          }
        }
        __syncwarp();
      }
      if (single) {
#pragma unroll
        for (int u = 0; u < LS_JT; ++u) {  // this warp's rotations ascend: '>' keeps the earliest
          const double sc = __ddiv_rn(acc[u], (double)n_pts);
          if (sc > rbest[u]) { rbest[u] = sc; rbest_rot[u] = r; }
        }
      } else {
#pragma unroll
        for (int u = 0; u < LS_JT; ++u)  // probability / pointcloud.size() (cc:76)
          s_res[warp * LS_TILE + (lj + u) * LS_TI + li] = __ddiv_rn(acc[u], (double)n_pts);
      }
    }
    if (single) continue;
    __syncthreads();
    // merge the round in work-item (= ascending rotation) order: std::max keeps
    // the earlier entry on ties (cc:109-110)
    if (threadIdx.x < LS_TILE) {
      const int t = threadIdx.x;
      for (int w = 0; w < LS_WARPS && round0 + w < n_items; ++w) {
        const int it = round0 + w, r = rot0 + it / n_tiles, tile = it % n_tiles;
        const int i = (tile % tiles_x) * LS_TI + (t & 15), j = (tile / tiles_x) * LS_TJ + (t >> 4);
        if (i < n_t && j < n_t) {
          const double sc = s_res[w * LS_TILE + t];
          const int tr = i * n_t + j;
          if (sc > s_best[tr]) { s_best[tr] = sc; s_best_rot[tr] = r; }
        }
      }
    }
    __syncthreads();
  }
  if (single) {
    // merge the warps: higher score wins, equal scores keep the smaller rotation
    for (int w = 0; w < LS_WARPS; ++w) {
      if (warp == w && li < n_t) {
#pragma unroll
        for (int u = 0; u < LS_JT; ++u) {
          const int j = lj + u;
          if (j < n_t) {
            const int tr = li * n_t + j;
            if (rbest[u] > s_best[tr] || (rbest[u] == s_best[tr] && rbest_rot[u] < s_best_rot[tr])) {
              s_best[tr] = rbest[u];
              s_best_rot[tr] = rbest_rot[u];
            }
          }
        }
      }
      __syncthreads();
    }
  }
}

__global__ void __launch_bounds__(LS_THREADS) low_score_kernel(
    Geometry g, SearchGrid sg, CellMap cm_global, const float2* __restrict__ points,
    const long long* __restrict__ offsets, const TableSlot* __restrict__ slots,
    const TableMeta* __restrict__ metas, const PairWork* __restrict__ pairs,
    const Rotation* __restrict__ rots, SearchBuffers sb) {
  const PairWork pw = pairs[blockIdx.y];
  const int rot0 = blockIdx.x * sb.rot_per_chunk;
  if (rot0 >= pw.n_rot) return;
  const int rot1 = min(pw.n_rot, rot0 + sb.rot_per_chunk);
  const int n_t = sg.n_t, n_tr = n_t * n_t;

  extern __shared__ __align__(16) unsigned char s_raw[];
  double* s_tab = reinterpret_cast<double*>(s_raw);                  // [LS_TABLE_CELLS]
  double* s_res = s_tab + LS_TABLE_CELLS;                            // [LS_WARPS][LS_TILE]
  int* s_lists = reinterpret_cast<int*>(s_res + LS_WARPS * LS_TILE); // cell lists (16 B aligned)
  double* s_best = reinterpret_cast<double*>(s_lists + LS_WARPS * LS_LIST_INTS);  // [n_tr]
  int* s_best_rot = reinterpret_cast<int*>(s_best + n_tr);           // [n_tr]
  float* s_thr = reinterpret_cast<float*>(s_best_rot + n_tr);
  float* s_tlow = s_thr + (g.wl + 2);                                // [n_t]

  const float2* a = points + offsets[pw.a_scan];
  const int n_pts = (int)(offsets[pw.a_scan + 1] - offsets[pw.a_scan]);
  const TableMeta tm = metas[pw.table];
  const double* __restrict__ lowlog = slots[pw.table].lowlog;
  const int lstride = g.wl + 1;

  for (int k = threadIdx.x; k < g.wl + 2; k += LS_THREADS) s_thr[k] = cm_global.lowf[k];
  for (int k = threadIdx.x; k < n_t; k += LS_THREADS) s_tlow[k] = sg.t_low[k];
  for (int tr = threadIdx.x; tr < n_tr; tr += LS_THREADS) {
    s_best[tr] = -INFINITY;
    s_best_rot[tr] = 0;
  }
  CellMap cm = cm_global;
  cm.lowf = s_thr;

  // box of low-res cells above the floor; everything outside reads log(1e-10)
  const bool empty = tm.lx1 < tm.lx0 || tm.ly1 < tm.ly0;
  const int tw = empty ? 0 : tm.lx1 - tm.lx0 + 1, th = empty ? 0 : tm.ly1 - tm.ly0 + 1;
  const int ox = empty ? 0 : tm.lx0, oy = empty ? 0 : tm.ly0;
  const bool fits = (tw + 1) * (th + 1) <= LS_TABLE_CELLS;
  if (fits) {
    const int ts = tw + 1;
    const double floor_log = lowlog[(size_t)g.wl * lstride + g.wl];
    for (int k = threadIdx.x; k < ts * (th + 1); k += LS_THREADS) {
      const int r = k / ts, c = k - r * ts;
      s_tab[k] = (r < th && c < tw) ? lowlog[(size_t)(oy + r) * lstride + ox + c] : floor_log;
    }
    __syncthreads();
    low_score_rounds<true>(g, sg, cm, s_tlow, s_tab, ox, oy, tw, th, ts, a, n_pts,
                           rots + pw.rot_begin, rot0, rot1, s_lists, s_res, s_best, s_best_rot);
  } else {
    __syncthreads();
    // whole table from global memory: row wl / column wl of lowlog hold the floor
    low_score_rounds<false>(g, sg, cm, s_tlow, lowlog, 0, 0, g.wl, g.wl, lstride, a, n_pts,
                            rots + pw.rot_begin, rot0, rot1, s_lists, s_res, s_best, s_best_rot);
  }
  const size_t base = ((size_t)blockIdx.y * sb.chunks_per_pair + blockIdx.x) * n_tr;
  for (int tr = threadIdx.x; tr < n_tr; tr += LS_THREADS) {
    sb.chunk_best[base + tr] = s_best[tr];
    sb.chunk_rot[base + tr] = s_best_rot[tr];
  }
}

void launch_low_score(const Geometry& g, const SearchGrid& sg, const CellMap& cm_low,
                      const float2* points, const long long* offsets, const TableSlot* slots,
                      const TableMeta* metas, const PairWork* pairs, const Rotation* rots,
                      int n_pairs, const SearchBuffers& sb, cudaStream_t stream) {
  if (n_pairs <= 0) return;
  const size_t smem = low_score_smem_bytes(sg.n_t, g.wl);
  static size_t configured = 0;
  if (smem > configured) {
    cudaFuncSetAttribute(low_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = smem;
  }
  dim3 grid(sb.chunks_per_pair, n_pairs);
  low_score_kernel<<<grid, LS_THREADS, smem, stream>>>(g, sg, cm_low, points, offsets, slots,
                                                       metas, pairs, rots, sb);
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
constexpr int RF_STRIDE = RF_POINT_CHUNK + 1;  // odd: list rows start on different banks

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
    Geometry g, SearchGrid sg, CellMap cm, const float2* __restrict__ points,
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
  int* cyh = cxh + nh_max * RF_STRIDE;                        // row offset or -1
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
          const int cx = cell_fast(cm, __fadd_rn(q.x, thx[k]));
          cxh[k * RF_STRIDE + pl] = (cx >= tm.ax0 && cx <= tm.ax1) ? cx - tm.ax0 : -1;
        }
        for (int k = 0; k < nhy; ++k) {
          const int cy = cell_fast(cm, __fadd_rn(q.y, thy[k]));
          cyh[k * RF_STRIDE + pl] =
              (cy >= tm.ay0 && cy <= tm.ay1) ? (cy - tm.ay0) * tm.stride : -1;
        }
      }
      __syncthreads();
      // lanes run along x (consecutive cells of one table row => coalesced
      // gathers); the candidate's rank in the reference's loop is kx*nhy + ky
      for (int w = threadIdx.x; w < n_cand; w += RF_THREADS) {
        const int kx = w % nhx, ky = w / nhx;
        const int c = kx * nhy + ky;            // x outer, y inner (cc:126-129)
        const int* px = cxh + kx * RF_STRIDE;
        const int* py = cyh + ky * RF_STRIDE;
        double s = hacc[c];
        int pl = 0;
        // the table holds lookup_log of every cell (normalize_pool_kernel); eight
        // independent gathers in flight, added in point order
        for (; pl + 8 <= pc; pl += 8) {
          double t[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int ox = px[pl + u], oy = py[pl + u];
            t[u] = ((ox | oy) >= 0) ? __ldg(hi + oy + ox) : CSM_FLOOR_LOG;
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) s = __dadd_rn(s, t[u]);
        }
        for (; pl < pc; ++pl) {
          const int ox = px[pl], oy = py[pl];
          s = __dadd_rn(s, ((ox | oy) >= 0) ? __ldg(hi + oy + ox) : CSM_FLOOR_LOG);
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
         (size_t)2 * nh_max * RF_STRIDE * sizeof(int);
}

void launch_refine(const Geometry& g, const SearchGrid& sg, const CellMap& cm_high,
                   const float2* points, const long long* offsets, const TableSlot* slots,
                   const TableMeta* metas, const PairWork* pairs, const Rotation* rots,
                   int n_pairs, const SearchBuffers& sb, PairResult* results,
                   cudaStream_t stream) {
  if (n_pairs <= 0) return;
  const size_t smem = refine_smem_bytes(sg.n_t, sg.n_h_max);
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    cudaFuncSetAttribute(refine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = smem;
  }
  refine_kernel<<<n_pairs, RF_THREADS, smem, stream>>>(g, sg, cm_high, points, offsets, slots, metas,
                                                       pairs, rots, sb, results);
}

}  // namespace csm
