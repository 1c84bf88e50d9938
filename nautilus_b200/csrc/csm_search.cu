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

// floor(v / res) + dim/2 (possibly outside the grid) when the float estimate of
// the quotient lies more than eps from a cell edge and below qmax; false when
// the caller has to use cell_fast.  See SearchGrid::reg_eps.
__device__ __forceinline__ bool cell_clear(const CellMap& m, float v, float eps, float qmax,
                                           int* cell) {
  const float e = __fmul_rn(v, m.inv_res);
  const float fl = floorf(e);
  const float fr = e - fl;
  *cell = (int)fl + m.half;
  return fabsf(e) < qmax && fr > eps && fr < 1.0f - eps;
}

// ---------------------------------------------------------------------------
// K5  low-res exhaustive scoring + per-translation max over a theta chunk.
// grid = (theta chunks, pairs), 8 warps per block.
//
// A warp scores one (rotation, translation tile) work item.  A tile is 14
// x-translations by 16 y-translations: lane & 15 picks the y-translation, lane
// >> 4 the group of 7 consecutive x-translations whose running sums the lane
// keeps in registers (the reference's 14 x 14 grid is exactly one tile).
// Because the translation step equals the cell size (both are low_res_,
// cc:96-99), the 7 cells a lane needs for one point are 7 CONSECUTIVE doubles
// of one table row: per point the lane loads its row offset (per-lane list)
// and the half-warp's first column (one broadcast word), then issues 7 loads at
// compile-time offsets -- no index arithmetic per lookup.  The rare point whose
// x cells are not consecutive (a sum that lands within float rounding of a cell
// edge) is flagged in phase A and recomputed cell by cell.
//
// The part of the pooled log table that lies above the lookup floor
// (TableMeta::lx0..ly1) is staged in shared memory once per block with 7 floor
// columns on either side and one floor row, so windows that leave the box need
// no special case.  The row stride is odd: the 16 lanes of a half-warp read 16
// different rows at the same column and hit 16 different bank pairs.  Tables
// whose box does not fit the budget are gathered from global memory through L1.
// Sums run over points in point order, one lane per candidate, like cc:66-77.
// ---------------------------------------------------------------------------
constexpr int LS_WARPS = 12;
constexpr int LS_THREADS = LS_WARPS * 32;
constexpr int LS_PC = 32;            // points per chunk = one per lane in phase A
constexpr int LS_TJ = 16;            // y translations per tile
constexpr int LS_IU = 7;             // x translations per lane
constexpr int LS_TI = 2 * LS_IU;     // x translations per tile
constexpr int LS_TILE = LS_TI * LS_TJ;
constexpr int LS_XPAD = CSM_LOW_PAD;  // floor columns on either side of the staged box
static_assert(CSM_LOW_PAD == LS_IU, "a lane's window must fit the floor padding");
// per-warp lists, one entry per point of the chunk: first row, first column of
// either half, rotated x and y
constexpr int LS_LIST_BYTES = 5 * LS_PC * 4;

size_t low_score_smem_bytes(int n_t, int wl, int table_cells) {
  const bool single = n_t <= LS_TI;
  size_t b = sizeof(double) * (size_t)table_cells;                   // staged table
  if (!single) b += sizeof(double) * LS_WARPS * LS_TILE;             // per-round results
  b += sizeof(double) * (size_t)n_t * n_t;                           // best score
  b += (size_t)LS_WARPS * LS_LIST_BYTES;                             // lists
  b += sizeof(int) * (size_t)n_t * n_t;                              // best rotation
  b += sizeof(float) * ((size_t)wl + 2 + n_t);                       // thresholds, translations
  return b;
}

// Geometry of the table a block scores against (staged box or the whole table).
struct LowTab {
  const double* tab;  // shared (staged) or global (lowlog)
  int ox, oy;         // low-res cell of box column / row 0
  int tw, th;         // box width / height in cells
  int ts;             // row stride in doubles
  int xpad;           // floor columns left of the box (column 0 always reads the floor)
  int row0;           // list value of box row 0 (1 when a floor row is staged above the box)
  int row_above;      // list value for cells above the box (cy < 0): a floor row
  int row_below;      // list value for cells below the box or outside the grid
};

template <bool SMEM>
__device__ __forceinline__ double tab_load(const double* p) {
  if (SMEM) return *p;
  return __ldg(p);
}

template <bool SMEM>
__device__ __forceinline__ void low_score_rounds(
    const SearchGrid& sg, const CellMap& cm, const float* s_tlow, const LowTab lt,
    const float2* __restrict__ a, int n_pts, const Rotation* __restrict__ rots, int rot0,
    int rot1, unsigned char* s_lists, double* s_res, double* s_best, int* s_best_rot) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_t = sg.n_t;
  const int tiles_x = (n_t + LS_TI - 1) / LS_TI, tiles_y = (n_t + LS_TJ - 1) / LS_TJ;
  const int n_tiles = tiles_x * tiles_y;
  const int n_items = (rot1 - rot0) * n_tiles;
  int* lyb = reinterpret_cast<int*>(s_lists + warp * LS_LIST_BYTES);  // [LS_PC]
  int* lxb = lyb + LS_PC;                                     // [2][LS_PC]
  float* lqx = reinterpret_cast<float*>(lxb + 2 * LS_PC);    // [LS_PC]
  float* lqy = lqx + LS_PC;                                   // [LS_PC]
  const int lj = lane & 15, half = lane >> 4;
  const bool single = (n_tiles == 1);  // the reference's 14 x 14 grid: best stays in registers
  double rbest[LS_IU];
  int rbest_rot[LS_IU];
#pragma unroll
  for (int u = 0; u < LS_IU; ++u) { rbest[u] = -INFINITY; rbest_rot[u] = 0; }

  for (int round0 = 0; round0 < n_items; round0 += LS_WARPS) {
    const int item = round0 + warp;
    if (item < n_items) {
      const int r = rot0 + item / n_tiles, tile = item % n_tiles;
      const int ti = tile % tiles_x, tj = tile / tiles_x;
      const int i0 = ti * LS_TI + half * LS_IU;   // this lane's first x translation
      const Rotation rot = rots[r];
      double acc[LS_IU];
#pragma unroll
      for (int u = 0; u < LS_IU; ++u) acc[u] = 0.0;
      for (int p0 = 0; p0 < n_pts; p0 += LS_PC) {
        const int pc = min(LS_PC, n_pts - p0);
        // phase A (lane = point): first cell of the tile's 16 y translations and of
        // each group of 7 x translations.  The translations step by exactly one
        // cell, so when the first sum lies clear of a cell edge (cell_clear, margin
        // SearchGrid::reg_eps proven on the host) the others are the next cells.
        bool odd = false;  // this lane's point needs every cell looked up on its own
        if (lane < pc) {
          const float2 q = rotate_point(a[p0 + lane], rot.c, rot.s);
          int cy0, wy = INT_MIN;
          if (cell_clear(cm, __fadd_rn(q.y, s_tlow[tj * LS_TJ]), sg.reg_eps, sg.reg_qmax, &cy0))
            wy = cy0 - lt.oy;  // Vector2f(., y_trans): box row of the first y translation
          odd = wy == INT_MIN;
          lyb[lane] = wy;
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const int ii0 = ti * LS_TI + hh * LS_IU;
            int word = 0;  // a group past the grid reads the floor padding
            if (ii0 < n_t) {
              int c0;
              // columns c0 .. c0+6 relative to the padded box; a window entirely
              // outside the box is moved onto the floor padding
              if (cell_clear(cm, __fadd_rn(q.x, s_tlow[ii0]), sg.reg_eps, sg.reg_qmax, &c0))
                word = max(0, min(c0 - lt.ox + lt.xpad, lt.tw + lt.xpad));
              else
                odd = true;
            }
            lxb[hh * LS_PC + lane] = word;
          }
          lqx[lane] = q.x;
          lqy[lane] = q.y;
        }
        const bool any_odd = __any_sync(0xffffffffu, odd);  // also orders the list writes
        const int* px = lxb + half * LS_PC;
        if (!any_odd) {
#pragma unroll 4
          for (int pl = 0; pl < pc; ++pl) {  // phase B: 7 consecutive cells per lane, point order
            // rows of the 16 lanes are consecutive, clamped to the floor rows on
            // either side of the box: one span of 16 rows -> 16 different bank pairs
            const int cy = lyb[pl] + lj;
            const int row = cy < 0 ? lt.row_above : (cy >= lt.th ? lt.row_below : cy + lt.row0);
            const double* t = lt.tab + (row * lt.ts + px[pl]);
#pragma unroll
            for (int u = 0; u < LS_IU; ++u) acc[u] = __dadd_rn(acc[u], tab_load<SMEM>(t + u));
          }
        } else {
          const int jj = tj * LS_TJ + lj;
          for (int pl = 0; pl < pc; ++pl) {  // same sums, every cell looked up on its own
            int row = lt.row_below;
            if (jj < n_t) {
              const int cg = cell_fast(cm, __fadd_rn(lqy[pl], s_tlow[jj]));
              const int cy = cg - lt.oy;
              row = cg < 0 ? lt.row_below
                           : (cy < 0 ? lt.row_above : (cy >= lt.th ? lt.row_below : cy + lt.row0));
            }
            const int ro = row * lt.ts;
            const float qx = lqx[pl];
#pragma unroll
            for (int u = 0; u < LS_IU; ++u) {
              int col = 0;
              if (i0 + u < n_t) {
                const int cx = cell_fast(cm, __fadd_rn(qx, s_tlow[i0 + u])) - lt.ox;
                if ((unsigned)cx < (unsigned)lt.tw) col = cx + lt.xpad;
              }
              acc[u] = __dadd_rn(acc[u], tab_load<SMEM>(lt.tab + ro + col));
            }
          }
        }
        __syncwarp();
      }
      if (single) {
#pragma unroll
        for (int u = 0; u < LS_IU; ++u) {  // this warp's rotations ascend: '>' keeps the earliest
          const double sc = __ddiv_rn(acc[u], (double)n_pts);
          if (sc > rbest[u]) { rbest[u] = sc; rbest_rot[u] = r; }
        }
      } else {
#pragma unroll
        for (int u = 0; u < LS_IU; ++u)  // probability / pointcloud.size() (cc:76)
          s_res[warp * LS_TILE + (half * LS_IU + u) * LS_TJ + lj] = __ddiv_rn(acc[u], (double)n_pts);
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
        const int i = (tile % tiles_x) * LS_TI + t / LS_TJ, j = (tile / tiles_x) * LS_TJ + t % LS_TJ;
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
      if (warp == w && lj < n_t) {
#pragma unroll
        for (int u = 0; u < LS_IU; ++u) {
          const int i = half * LS_IU + u;
          if (i < n_t) {
            const int tr = i * n_t + lj;
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
    const Rotation* __restrict__ rots, SearchBuffers sb, int table_cells) {
  const PairWork pw = pairs[blockIdx.y];
  const int rot0 = blockIdx.x * sb.rot_per_chunk;
  if (rot0 >= pw.n_rot) return;
  const int rot1 = min(pw.n_rot, rot0 + sb.rot_per_chunk);
  const int n_t = sg.n_t, n_tr = n_t * n_t;
  const bool single = n_t <= LS_TI;

  extern __shared__ __align__(16) unsigned char s_raw[];
  double* s_tab = reinterpret_cast<double*>(s_raw);                  // [table_cells]
  double* s_res = s_tab + table_cells;                               // [LS_WARPS][LS_TILE] or none
  double* s_best = s_res + (single ? 0 : LS_WARPS * LS_TILE);        // [n_tr]
  unsigned char* s_lists = reinterpret_cast<unsigned char*>(s_best + n_tr);
  int* s_best_rot = reinterpret_cast<int*>(s_lists + LS_WARPS * LS_LIST_BYTES);  // [n_tr]
  float* s_thr = reinterpret_cast<float*>(s_best_rot + n_tr);
  float* s_tlow = s_thr + (g.wl + 2);                                // [n_t]

  const float2* a = points + offsets[pw.a_scan];
  const int n_pts = (int)(offsets[pw.a_scan + 1] - offsets[pw.a_scan]);
  const TableMeta tm = metas[pw.table];
  const double* __restrict__ lowlog = slots[pw.table].lowlog;
  const int lstride = lowlog_stride(g.wl);

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
  LowTab lt;
  lt.tw = empty ? 0 : tm.lx1 - tm.lx0 + 1;
  lt.th = empty ? 0 : tm.ly1 - tm.ly0 + 1;
  lt.ox = empty ? 0 : tm.lx0;
  lt.oy = empty ? 0 : tm.ly0;
  const int ts = (lt.tw + 2 * LS_XPAD) | 1;  // odd stride: see the bank note above
  const bool fits = ts * (lt.th + 2) <= table_cells && lt.th < 65533;
  if (fits) {
    // staged rows: 0 = floor, 1 .. th = box, th + 1 = floor
    lt.tab = s_tab; lt.ts = ts; lt.xpad = LS_XPAD;
    lt.row0 = 1; lt.row_above = 0; lt.row_below = lt.th + 1;
    for (int k = threadIdx.x; k < ts * (lt.th + 2); k += LS_THREADS) {
      const int r = k / ts - 1, c = k - (r + 1) * ts - LS_XPAD;
      s_tab[k] = ((unsigned)r < (unsigned)lt.th && (unsigned)c < (unsigned)lt.tw)
                     ? lowlog[(size_t)(lt.oy + r) * lstride + lt.ox + c + CSM_LOW_PAD] : CSM_FLOOR_LOG;
    }
    __syncthreads();
    low_score_rounds<true>(sg, cm, s_tlow, lt, a, n_pts, rots + pw.rot_begin, rot0, rot1, s_lists,
                           s_res, s_best, s_best_rot);
  } else {
    // whole table from global memory: lowlog has the same padded layout
    lt.tab = lowlog; lt.ox = 0; lt.oy = 0; lt.tw = g.wl; lt.th = g.wl; lt.ts = lstride;
    lt.xpad = CSM_LOW_PAD; lt.row0 = 0; lt.row_above = g.wl; lt.row_below = g.wl;
    __syncthreads();
    low_score_rounds<false>(sg, cm, s_tlow, lt, a, n_pts, rots + pw.rot_begin, rot0, rot1, s_lists,
                            s_res, s_best, s_best_rot);
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
                      int n_pairs, const SearchBuffers& sb, int table_cells, cudaStream_t stream) {
  if (n_pairs <= 0) return;
  const size_t smem = low_score_smem_bytes(sg.n_t, g.wl, table_cells);
  static size_t configured = 0;
  if (smem > configured) {
    cudaFuncSetAttribute(low_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    configured = smem;
  }
  dim3 grid(sb.chunks_per_pair, n_pairs);
  low_score_kernel<<<grid, LS_THREADS, smem, stream>>>(g, sg, cm_low, points, offsets, slots,
                                                       metas, pairs, rots, sb, table_cells);
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
