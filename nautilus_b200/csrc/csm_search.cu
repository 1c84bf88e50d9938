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
// grid = (theta chunks, pairs), LS_WARPS warps per block.
//
// Lanes.  A tile is 14 x-translations by 16 y-translations: lane & 15 picks the
// y-translation, lane >> 4 the group of 7 consecutive x-translations whose
// running sums the lane keeps in registers (the reference's 14 x 14 grid is one
// tile).  Because the translation step equals the cell size (both are low_res_,
// cc:96-99), the 7 cells a lane needs for one point are 7 CONSECUTIVE doubles of
// one table row and the 16 lanes of a half-warp read 16 consecutive rows: a
// point's 14 x 16 patch is addressed by ONE cell index (its first row and column),
// computed once per (rotation, point) in phase A (lane = point) from one float
// quotient per axis; the host proves the margin (SearchGrid::reg_eps) within which
// "translation k lands k cells further" is exact for all 16 translations of a tile.
// The rare sum that lands within that margin of a cell edge is flagged and looked up
// cell by cell (cell_fast) instead.
//
// Table.  The scorer reads the pooled log table through a copy that is padded with a
// whole tile of lookup-floor cells on every side (CSM_LOW_PADX / CSM_LOW_PADY): a patch
// that straddles the table edge reads real cells and floor cells from the same
// addresses as an inner one, and a patch entirely outside is clamped onto the padding.
// Every regular point is therefore the same case - no edge / floor / inner kinds, no
// clamps in the inner loop.  The copy is the above-floor box of the table
// (TableMeta::lx0..ly1, everything outside it reads the floor anyway) staged in shared
// memory once per block, odd row stride so that the 16 rows of a half-warp hit 16
// different bank pairs; tables whose box does not fit the budget are read from the
// global array, which has the same padded layout, through L1.
//
// Runs.  Consecutive beams of a scan land 1-5 cm apart, far less than a 0.3 m
// cell, so consecutive points mostly address the SAME patch (82 % on the bench
// scans).  Phase A run-length encodes the patch indices of a 32-point chunk
// (shuffle + ballot); phase B loads the 7 values of a run once and adds them
// once per point of the run.  Every sum still takes its terms in point order,
// one lane per candidate, like cc:66-77 -- only the loads are shared -- and the
// kernel is bound by the fp64 add pipe instead of shared-memory bandwidth.
// A warp owns a rotation; the warps of a block take rotations round-robin.
// ---------------------------------------------------------------------------
// 16 warps x 2 blocks per SM = 64 registers per thread, which the kernel just fits without
// spilling (18 warps: 56 registers, 200 B of spills).  With the rotations handed out by a
// counter: 12 warps 11.87 ms, 16 warps 11.60 ms on the bench batch; wide search 899 / 790 ms.
#ifndef LS_WARPS_N
#define LS_WARPS_N 16
#endif
#ifndef LS_MINB_N      // resident blocks per SM the register budget is set for
#define LS_MINB_N 2
#endif
constexpr int LS_WARPS = LS_WARPS_N;
constexpr int LS_THREADS = LS_WARPS * 32;
constexpr int LS_PC = 32;            // points per chunk = one per lane in phase A
constexpr int LS_TJ = 16;            // y translations per tile
constexpr int LS_IU = 7;             // x translations per lane
constexpr int LS_TI = 2 * LS_IU;     // x translations per tile
static_assert(CSM_LOW_PADX == LS_TI && CSM_LOW_PADY == LS_TJ, "the padding is one tile");
// run descriptor: bits 0-23 the patch, 24-29 points in the run, 31 slow.  The patch is
// the index of its first cell in the padded array, or - when the staged copy keeps only
// ONE floor row above and below its box to fit the shared-memory budget (ROWCLAMP) -
// its first row + LS_TJ in bits 0-9 and its first column in bits 10-20; the lanes then
// clamp their own row onto the two floor rows.
constexpr unsigned LS_SLOW = 0x80000000u;   // look every cell up on its own (always a run of one)
constexpr unsigned LS_CELL_MASK = 0x00ffffffu;

int low_score_warps() { return LS_WARPS; }

size_t low_score_smem_bytes(int n_t, int wl, int table_cells) {
  size_t b = sizeof(double) * (size_t)table_cells;                   // staged table
  b += sizeof(double) * (((size_t)n_t * n_t + 1) & ~(size_t)1);      // best score
  b += sizeof(int) * (size_t)n_t * n_t;                              // best rotation
  b += sizeof(float) * ((size_t)wl + 2 + n_t);                       // thresholds, translations
  return b;
}

// Geometry of the padded table a block scores against (staged box or the global array).
struct LowTab {
  const double* tab;  // shared (staged) or global (lowlog)
  unsigned tab_s;     // shared-window address of tab when it is staged
  int ox, oy;         // low-res cell of box column / row 0
  int tw, th;         // box width / height in cells
  int ts;             // row stride in doubles
  int pady;           // floor rows above / below the box: CSM_LOW_PADY, or 1 with ROWCLAMP
  // padded rows: [0, pady) floor, [pady, pady + th) box, [pady + th, th + 2 pady) floor;
  // columns likewise with CSM_LOW_PADX
};

template <bool SMEM>
__device__ __forceinline__ double tab_load(const double* p) {
  if (SMEM) return *p;
  return __ldg(p);
}

// Shared memory is addressed by its 32-bit window address, which spares the generic ->
// shared conversion per access.
template <int OFF>
__device__ __forceinline__ double lds_f64(unsigned addr) {
  double v;
  asm("ld.shared.f64 %0, [%1+%2];" : "=d"(v) : "r"(addr), "n"(OFF));
  return v;
}

// The 7 consecutive values a lane adds for one run: `ad` = shared-window byte address
// of the first (staged table), `gp` = its global address otherwise.
template <bool SMEM>
__device__ __forceinline__ void tab_load7(const double* gp, unsigned ad, double* t) {
  static_assert(LS_IU == 7, "seven loads below");
  if (SMEM) {
    t[0] = lds_f64<0>(ad);  t[1] = lds_f64<8>(ad);  t[2] = lds_f64<16>(ad); t[3] = lds_f64<24>(ad);
    t[4] = lds_f64<32>(ad); t[5] = lds_f64<40>(ad); t[6] = lds_f64<48>(ad);
  } else {
#pragma unroll
    for (int u = 0; u < LS_IU; ++u) t[u] = __ldg(gp + u);
  }
}

// Keeps a loop-invariant value in its register: without this the compiler
// re-derives shared-window addresses (3 uniform instructions) at every use.
__device__ __forceinline__ unsigned pinned(unsigned v) {
  asm volatile("mov.u32 %0, %0;" : "+r"(v));
  return v;
}

// The 7 table values of one lane for one (rotation, point), every cell looked up
// on its own (cell_fast): the path of a point flagged LS_SLOW.
template <bool SMEM>
__device__ __noinline__ void low_patch_slow(const CellMap& cm, const float* s_tlow, const LowTab& lt,
                                            float2 q, int n_t, int jj, int i0, double* t) {
  int row = 0;  // padded row 0 reads the floor
  if (jj < n_t) {
    const int cg = cell_fast(cm, __fadd_rn(q.y, s_tlow[jj]));  // Vector2f(., y_trans)
    const int cy = cg - lt.oy;
    if (cg >= 0 && (unsigned)cy < (unsigned)lt.th) row = cy + lt.pady;
  }
  const int ro = row * lt.ts;
#pragma unroll
  for (int u = 0; u < LS_IU; ++u) {
    int col = 0;  // padded column 0 reads the floor
    if (i0 + u < n_t) {
      const int cg = cell_fast(cm, __fadd_rn(q.x, s_tlow[i0 + u]));  // Vector2f(x_trans, .)
      const int cx = cg - lt.ox;
      if (cg >= 0 && (unsigned)cx < (unsigned)lt.tw) col = cx + CSM_LOW_PADX;
    }
    t[u] = tab_load<SMEM>(lt.tab + ro + col);
  }
}

template <bool SMEM, bool ROWCLAMP>
__device__ __forceinline__ void low_score_rounds(
    const SearchGrid& sg, const CellMap& cm, const float* s_tlow, const LowTab lt,
    const float2* __restrict__ a, int n_pts, const Rotation* __restrict__ rots, int rot0,
    int rot1, double* s_best, int* s_best_rot, int* s_next, bool packable) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_t = sg.n_t;
  const int tiles_x = (n_t + LS_TI - 1) / LS_TI, tiles_y = (n_t + LS_TJ - 1) / LS_TJ;
  const int lj = lane & 15, half = lane >> 4;
  const unsigned above = lane == 31 ? 0u : ~((2u << lane) - 1u);     // lanes after this one
  // this lane's cell within a patch: row lj, columns half * 7 .. half * 7 + 6
  const int lane_cell = (ROWCLAMP ? 0 : lj * lt.ts) + half * LS_IU;
  const unsigned lane_s = pinned(lt.tab_s + 8u * (unsigned)lane_cell);
  const double* lane_g = lt.tab + lane_cell;
  // clamps of a patch's first row / column: 0 and the maxima lie entirely in the padding
  // (ROWCLAMP: rows are kept relative to LS_TJ rows above the box and clamped per lane)
  const int r_max = lt.th + CSM_LOW_PADY, c_max = lt.tw + CSM_LOW_PADX;
  const int r_off = CSM_LOW_PADY - lt.oy, c_off = CSM_LOW_PADX - lt.ox;
  const int lane_row = lj - (CSM_LOW_PADY - 1);  // ROWCLAMP: staged row = clamp(first row + lane_row, 0, th + 1)
  const int row_last = lt.th + 1;

  for (int tile = 0; tile < tiles_x * tiles_y; ++tile) {
    const int ti = tile % tiles_x, tj = tile / tiles_x;
    const int i0 = ti * LS_TI + half * LS_IU;   // this lane's first x translation
    const int jj = tj * LS_TJ + lj;             // this lane's y translation
    const float ty0 = s_tlow[tj * LS_TJ], tx0 = s_tlow[ti * LS_TI];
    if (threadIdx.x == 0) *s_next = rot0;  // (the merge loop below ends on a barrier)
    __syncthreads();
    double rbest[LS_IU];  // best of this warp's rotations, which ascend: '>' keeps the earliest
    int rbest_rot[LS_IU];
#pragma unroll
    for (int u = 0; u < LS_IU; ++u) { rbest[u] = -INFINITY; rbest_rot[u] = 0; }

    // Rotations are handed out by a shared counter instead of round-robin: a warp that
    // falls behind (its scheduler's other warps sit in their add loops) takes fewer, and
    // each warp's rotations still ascend, which is all the '>' below relies on.
#pragma unroll 1
    for (;;) {
      int r = 0;
      if (lane == 0) r = atomicAdd(s_next, 1);
      r = __shfl_sync(0xffffffffu, r, 0);
      if (r >= rot1) break;
      const Rotation rot = rots[r];
      double acc[LS_IU];
#pragma unroll
      for (int u = 0; u < LS_IU; ++u) acc[u] = 0.0;

#pragma unroll 1
      for (int p0 = 0; p0 < n_pts; p0 += LS_PC) {
        const int pc = min(LS_PC, n_pts - p0);
        // phase A (lane = point): first cell of the point's patch.  The translations
        // step by exactly one cell, so when the first sum of an axis lies clear of a
        // cell edge (cell_clear, margin SearchGrid::reg_eps proven on the host for the 16
        // translations that follow) translation k hits cell + k on that axis.
        unsigned info = LS_SLOW | (unsigned)lane;  // distinct per lane: a slow point is its own run
        if (lane < pc) {
          const float2 q = rotate_point(a[p0 + lane], rot.c, rot.s);
          int cy0, cx0;
          bool ok = packable;
          ok = cell_clear(cm, __fadd_rn(q.y, ty0), sg.reg_eps, sg.reg_qmax, &cy0) && ok;
          ok = cell_clear(cm, __fadd_rn(q.x, tx0), sg.reg_eps, sg.reg_qmax, &cx0) && ok;
          if (ok) {
            const int r0 = max(0, min(cy0 + r_off, r_max));
            const int c0 = max(0, min(cx0 + c_off, c_max));
            info = ROWCLAMP ? (unsigned)r0 | ((unsigned)c0 << 10) : (unsigned)(r0 * lt.ts + c0);
          }
        }
        // run-length encode: a point whose patch equals the previous point's adds
        // the values already in registers.  Run starts come from a ballot; each start
        // lane works out the length of its run; phase B fetches {patch, length} of a run
        // from its first lane by shuffle and steps from start to start.
        const unsigned prev = __shfl_up_sync(0xffffffffu, info, 1);
        const unsigned starts = __ballot_sync(0xffffffffu, lane < pc && (lane == 0 || info != prev));
        {
          const unsigned later = starts & above;
          info |= (unsigned)((later ? __ffs(later) - 1 : pc) - lane) << 24;
        }
        // phase B (lane = candidate column): per run 7 consecutive cells, added once
        // per point of the run -- each sum still takes its terms in point order
#pragma unroll 1
        for (int pl = 0; pl < pc;) {
          const unsigned inf = __shfl_sync(0xffffffffu, info, pl);
          double t[LS_IU];
          int c = 1;
          if ((int)inf >= 0) {
            unsigned cell8;
            if (ROWCLAMP) {
              const int row = min(max((int)(inf & 1023u) + lane_row, 0), row_last);
              cell8 = (unsigned)(row * lt.ts + (int)((inf >> 10) & 2047u)) << 3;
            } else {
              cell8 = (inf & LS_CELL_MASK) << 3;
            }
            tab_load7<SMEM>(reinterpret_cast<const double*>(reinterpret_cast<const char*>(lane_g) + cell8),
                            lane_s + cell8, t);
            c = (int)(inf >> 24);
          } else {
            double ts[LS_IU];  // (a local array: keeps t itself in registers)
            low_patch_slow<SMEM>(cm, s_tlow, lt, rotate_point(a[p0 + pl], rot.c, rot.s), n_t, jj, i0, ts);
#pragma unroll
            for (int u = 0; u < LS_IU; ++u) t[u] = ts[u];
          }
          pl += c;
#pragma unroll 1
          for (; c >= 2; c -= 2) {
#pragma unroll
            for (int k = 0; k < 2; ++k)
#pragma unroll
              for (int u = 0; u < LS_IU; ++u) acc[u] = __dadd_rn(acc[u], t[u]);
          }
          if (c) {
#pragma unroll
            for (int u = 0; u < LS_IU; ++u) acc[u] = __dadd_rn(acc[u], t[u]);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < LS_IU; ++u) {
        const double sc = __ddiv_rn(acc[u], (double)n_pts);  // probability / size (cc:76)
        if (sc > rbest[u]) { rbest[u] = sc; rbest_rot[u] = r; }
      }
    }
    // merge the warps: higher score wins, equal scores keep the smaller rotation
    // (std::max keeps the earlier entry on ties, cc:109-110)
    for (int w = 0; w < LS_WARPS; ++w) {
      if (warp == w && jj < n_t) {
#pragma unroll
        for (int u = 0; u < LS_IU; ++u) {
          const int i = i0 + u;
          if (i < n_t && i < ti * LS_TI + LS_TI) {
            const int tr = i * n_t + jj;
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

__global__ void __launch_bounds__(LS_THREADS, LS_MINB_N) low_score_kernel(
    Geometry g, SearchGrid sg, CellMap cm_global, const float2* __restrict__ points,
    const long long* __restrict__ offsets, const TableSlot* __restrict__ slots,
    const TableMeta* __restrict__ metas, const PairWork* __restrict__ pairs,
    const Rotation* __restrict__ rots, SearchBuffers sb, int table_cells) {
  const PairWork pw = pairs[blockIdx.y];
  const int rot0 = blockIdx.x * sb.rot_per_chunk;
  if (rot0 >= pw.n_rot) return;
  const int rot1 = min(pw.n_rot, rot0 + sb.rot_per_chunk);
  const int n_t = sg.n_t, n_tr = n_t * n_t;

  extern __shared__ __align__(16) unsigned char s_raw[];
  __shared__ int s_next;  // next rotation of the chunk to hand out
  double* s_tab = reinterpret_cast<double*>(s_raw);                  // [table_cells]
  double* s_best = s_tab + table_cells;                              // [n_tr]
  int* s_best_rot = reinterpret_cast<int*>(s_best + ((n_tr + 1) & ~1));  // [n_tr]
  float* s_thr = reinterpret_cast<float*>(s_best_rot + n_tr);
  float* s_tlow = s_thr + (g.wl + 2);                                // [n_t]

  const float2* a = points + offsets[pw.a_scan];
  const int n_pts = (int)(offsets[pw.a_scan + 1] - offsets[pw.a_scan]);
  const TableMeta tm = metas[pw.table];
  const double* __restrict__ lowlog = slots[pw.table].lowlog;

  for (int k = threadIdx.x; k < g.wl + 1; k += LS_THREADS) s_thr[k] = cm_global.lowf[k];  // make_thresholds: dim + 1 entries
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
  const int ts = (lt.tw + 2 * CSM_LOW_PADX) | 1;  // odd stride: see the bank note above
  const bool full_pad = (long long)ts * (lt.th + 2 * CSM_LOW_PADY) <= table_cells;
  const bool one_pad = (long long)ts * (lt.th + 2) <= table_cells;
  if (full_pad || one_pad) {
    lt.tab = s_tab; lt.ts = ts; lt.pady = full_pad ? CSM_LOW_PADY : 1;
    lt.tab_s = (unsigned)__cvta_generic_to_shared(s_tab);
    const int rows = lt.th + 2 * lt.pady;
    const double* src = lowlog + lowlog_index(g.wl, lt.ox, lt.oy);  // box cell (0, 0)
    const int lstride = lowlog_stride(g.wl);
    for (int k = threadIdx.x; k < ts * rows; k += LS_THREADS) {
      const int r = k / ts - lt.pady, c = k - (r + lt.pady) * ts - CSM_LOW_PADX;
      s_tab[k] = ((unsigned)r < (unsigned)lt.th && (unsigned)c < (unsigned)lt.tw)
                     ? src[(size_t)r * lstride + c] : CSM_FLOOR_LOG;
    }
    __syncthreads();
    if (full_pad)
      low_score_rounds<true, false>(sg, cm, s_tlow, lt, a, n_pts, rots + pw.rot_begin, rot0, rot1,
                                    s_best, s_best_rot, &s_next, (long long)ts * rows <= (long long)LS_CELL_MASK);
    else  // one floor row above and below: rows packed in 10 bits, columns in 11
      low_score_rounds<true, true>(sg, cm, s_tlow, lt, a, n_pts, rots + pw.rot_begin, rot0, rot1,
                                   s_best, s_best_rot, &s_next,
                                   lt.th + 2 * CSM_LOW_PADY < 1024 && lt.tw + 2 * CSM_LOW_PADX < 2048);
  } else {
    // whole table from global memory: lowlog has the fully padded layout
    lt.tab = lowlog; lt.tab_s = 0; lt.ox = 0; lt.oy = 0; lt.tw = g.wl; lt.th = g.wl;
    lt.ts = lowlog_stride(g.wl); lt.pady = CSM_LOW_PADY;
    __syncthreads();
    low_score_rounds<false, false>(sg, cm, s_tlow, lt, a, n_pts, rots + pw.rot_begin, rot0, rot1,
                                   s_best, s_best_rot, &s_next, lowlog_cells(g.wl) <= (size_t)LS_CELL_MASK);
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
  static SmemOptIn opt;
  opt.allow(low_score_kernel, smem);  // a refusal surfaces as the launch error
  dim3 grid(sb.chunks_per_pair, n_pairs);
  low_score_kernel<<<grid, LS_THREADS, smem, stream>>>(g, sg, cm_low, points, offsets, slots,
                                                       metas, pairs, rots, sb, table_cells);
}

// ---------------------------------------------------------------------------
// K6  memo reduction + best-first hi-res refinement.
// Follows GetTransformation's loop (CorrelativeScanMatcher.cc:170-243):
//   pop the first maximum of the memo in slot (AbsCoords) order, invalidate
//   it, stop when it is below the best hi-res score so far, otherwise score
//   the hi-res window it opens and keep a strictly better result.
//
// The loop is a serial chain per pair (mean 3.6 windows, maximum 76 on the bench
// batch), so one block per pair leaves the GPU waiting for the longest chain.
// The default path therefore runs it in three kernels:
//   refine_first_kernel   block per pair: memo, first pop -> best1, and the list
//                         of slots the reference can still pop: every later pop
//                         needs bound >= best >= best1, and pops go in (bound
//                         descending, slot order) -- so they are a PREFIX of the
//                         slots with bound >= best1 in that order;
//   refine_spec_kernel    persistent blocks fetch (pair, slot) windows from one
//                         queue and score them independently (early exit against
//                         best1, which is never above the reference's bound);
//   refine_finish_kernel  block per pair: replays the reference's decisions over
//                         the list in order (stop / count / strictly better) and
//                         writes the result; a pair whose list was cut at
//                         SearchBuffers::spec_cap carries on with the serial loop.
// Windows scored beyond the reference's stopping point are discarded by the
// replay and not counted, so pose, score and the n_high / n_pops counters are
// those of the sequential loop.  launch_refine(serial=true) keeps the
// one-kernel form.
// ---------------------------------------------------------------------------
constexpr int RF_THREADS = 128;
// Resident blocks per SM the refine kernels are compiled for (0: left to the compiler:
// 77 / 80 registers, 6 blocks).  Measured refine stage, bench batch / 0.01 m configuration:
// first 8: 3.85 / 11.65 ms; spec 8: 3.80 / 11.98; both 8: 3.70 / 11.96; neither: 3.95 / 11.69.
// The 30 x 30 windows of the 0.01 m grid want the registers, the 10 x 10 ones the warps.
#ifndef RF_MINB_FIRST
#define RF_MINB_FIRST 8
#endif
#if RF_MINB_FIRST > 0
#define RF_BOUNDS_FIRST __launch_bounds__(RF_THREADS, RF_MINB_FIRST)
#else
#define RF_BOUNDS_FIRST __launch_bounds__(RF_THREADS)
#endif
constexpr int RF_POINT_CHUNK = 1024;  // points whose window words are staged at a time
#ifndef RF_DEPTH_N
#define RF_DEPTH_N 2
#endif
constexpr int RF_DEPTH = RF_DEPTH_N;  // groups of 8 gathers in flight per thread
constexpr int RF_SLOW = -1;           // window word of a point that is looked up cell by cell
static_assert(RF_POINT_CHUNK % 64 == 0, "flags end on an 8-byte boundary");

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

// Shared-memory carve-up of the refine kernels (dynamic part).
struct RefineShared {
  int* word;            // [RF_POINT_CHUNK], 16 B aligned
  unsigned char* flag;  // [RF_POINT_CHUNK / 8]
  double* hacc;         // [nh_max^2]
  double* memo;         // [n_tr]   (first / finish / serial kernels)
  int* memo_rot;        // [n_tr]
};

__device__ __forceinline__ RefineShared carve_refine(unsigned char* raw, int nh_max, int n_tr) {
  RefineShared sh;
  sh.word = reinterpret_cast<int*>(raw);
  sh.flag = reinterpret_cast<unsigned char*>(sh.word + RF_POINT_CHUNK);
  sh.hacc = reinterpret_cast<double*>(sh.flag + RF_POINT_CHUNK / 8);
  sh.memo = sh.hacc + nh_max * nh_max;
  sh.memo_rot = reinterpret_cast<int*>(sh.memo + n_tr);
  return sh;
}

// Every table value is log(v <= 1) <= 0, so a candidate's running sum only
// falls: once it is below best_prob * n_pts (taken a few ulps low, so that the
// rounded quotient cannot come out above best_prob) the candidate cannot win
// the strict '>' of cc:225 and its remaining points are skipped.
__device__ __forceinline__ double dead_bound(double best_prob, int n_pts) {
  return best_prob == -INFINITY ? -INFINITY
                                : __dmul_rn(__dmul_rn(best_prob, (double)n_pts), 1.0 + 0x1p-50);
}

// memo[slot] = max over theta, earliest theta on ties (chunks ascend in theta)
__device__ __forceinline__ void reduce_memo(const SearchBuffers& sb, int pair, int n_rot, int n_tr,
                                            double* memo, int* memo_rot) {
  const int n_chunks = (n_rot + sb.rot_per_chunk - 1) / sb.rot_per_chunk;
  for (int tr = threadIdx.x; tr < n_tr; tr += RF_THREADS) {
    double b = -INFINITY;
    int br = 0;
    for (int ch = 0; ch < n_chunks; ++ch) {
      const size_t at = ((size_t)pair * sb.chunks_per_pair + ch) * n_tr + tr;
      const double s = sb.chunk_best[at];
      if (s > b) { b = s; br = sb.chunk_rot[at]; }
    }
    memo[tr] = b;
    memo_rot[tr] = br;
  }
}

// std::max_element over the memo: first maximum in slot order (cc:172-174)
__device__ __forceinline__ ArgBest memo_top(const SearchGrid& sg, const double* memo, int n_tr,
                                            ArgBest* s_red) {
  ArgBest mine = {-INFINITY, INT_MAX, -1};
  for (int tr = threadIdx.x; tr < n_tr; tr += RF_THREADS) {
    ArgBest c = {memo[tr], sg.slot_rank[tr], tr};
    mine = arg_better(mine, c);
  }
  return block_argmax(mine, s_red);
}

// GetProbAndTransformation (cc:117-143) over the hi-res window of low-res slot
// (i, j) at one rotation, by the whole block: the best candidate (score already
// divided by n_pts; idx = kx * nhy + ky, the reference's x-outer / y-inner rank;
// idx < 0 when no candidate can beat `bound`).  Candidates whose running sum
// falls below `bound` (see dead_bound) are dropped.
__device__ ArgBest score_window(const SearchGrid& sg, const CellMap& cm, const float2* __restrict__ a,
                                int n_pts, const TableMeta& tm, const double* __restrict__ hi,
                                const Rotation rot, int i, int j, double bound,
                                const RefineShared& sh, ArgBest* s_red) {
  const int nh_max = sg.n_h_max;
  const int nhx = sg.n_h[i], nhy = sg.n_h[j], n_cand = nhx * nhy;
  const float* thx = sg.t_high + i * nh_max;
  const float* thy = sg.t_high + j * nh_max;
  int* s_word = sh.word;
  unsigned char* s_flag = sh.flag;
  double* hacc = sh.hacc;

  for (int c = threadIdx.x; c < n_cand; c += RF_THREADS) hacc[c] = 0.0;
  const float tx0 = thx[0], ty0 = thy[0];
  for (int p0 = 0; p0 < n_pts; p0 += RF_POINT_CHUNK) {
    const int pc = min(RF_POINT_CHUNK, n_pts - p0);
    __syncthreads();
    // Window words (thread = point).  The hi-res steps advance by exactly one
    // cell, so when the first sum of an axis lies clear of a cell edge (margin
    // SearchGrid::reg_eps_h, proven on the host) step k hits cell + k: a point
    // whose whole nhx x nhy window lies inside the table's rectangle is one
    // word = index of the window's first cell.  Other points are flagged.
    for (int pl = threadIdx.x; pl < ((pc + 31) & ~31); pl += RF_THREADS) {
      int word = RF_SLOW;
      if (pl < pc) {
        const float2 q = rotate_point(a[p0 + pl], rot.c, rot.s);
        int cx0, cy0;
        bool ok = cell_clear(cm, __fadd_rn(q.x, tx0), sg.reg_eps_h, sg.reg_qmax_h, &cx0);
        ok = cell_clear(cm, __fadd_rn(q.y, ty0), sg.reg_eps_h, sg.reg_qmax_h, &cy0) && ok;
        ok = ok && cx0 >= tm.ax0 && cx0 + nhx - 1 <= tm.ax1 && cy0 >= tm.ay0 && cy0 + nhy - 1 <= tm.ay1;
        if (ok) word = (cy0 - tm.ay0) * tm.stride + (cx0 - tm.ax0);
        s_word[pl] = word;
      }
      // one flag per group of 8 points: does any of them need the slow path
      const unsigned slow = __ballot_sync(0xffffffffu, pl < pc && word == RF_SLOW);
      if ((threadIdx.x & 7) == 0) s_flag[pl >> 3] = (unsigned char)((slow >> (threadIdx.x & 31)) & 0xffu);
    }
    __syncthreads();
    // lanes run along x (consecutive cells of one table row => coalesced
    // gathers); the candidate's rank in the reference's loop is kx*nhy + ky
    for (int w = threadIdx.x; w < n_cand; w += RF_THREADS) {
      const int kx = w % nhx, ky = w / nhx;
      const int c = kx * nhy + ky;            // x outer, y inner (cc:126-129)
      const double* __restrict__ hp = hi + ((size_t)ky * tm.stride + kx);
      double s = hacc[c];
      // the table holds lookup_log of every cell (normalize_pool_kernel); eight
      // independent gathers in flight, added in point order
      auto slow_value = [&](int p) -> double {
        const float2 q = rotate_point(a[p], rot.c, rot.s);
        const int cx = cell_fast(cm, __fadd_rn(q.x, thx[kx]));
        const int cy = cell_fast(cm, __fadd_rn(q.y, thy[ky]));
        if (cx < tm.ax0 || cx > tm.ax1 || cy < tm.ay0 || cy > tm.ay1) return CSM_FLOOR_LOG;
        return __ldg(hi + ((size_t)(cy - tm.ay0) * tm.stride + (cx - tm.ax0)));
      };
      // groups of 8 points; the loads of the next RF_DEPTH - 1 groups are issued
      // before the adds of the current one (the sum itself is a serial chain)
      const int ng = pc >> 3;
      double buf[RF_DEPTH][8];
      bool fast[RF_DEPTH];
      auto load8 = [&](int g, double* t) {
        const int4 w0 = *reinterpret_cast<const int4*>(s_word + 8 * g);
        const int4 w1 = *reinterpret_cast<const int4*>(s_word + 8 * g + 4);
        t[0] = __ldg(hp + w0.x); t[1] = __ldg(hp + w0.y); t[2] = __ldg(hp + w0.z); t[3] = __ldg(hp + w0.w);
        t[4] = __ldg(hp + w1.x); t[5] = __ldg(hp + w1.y); t[6] = __ldg(hp + w1.z); t[7] = __ldg(hp + w1.w);
      };
#pragma unroll
      for (int d = 0; d < RF_DEPTH - 1; ++d) {
        fast[d] = d < ng && s_flag[d] == 0;
        if (fast[d]) load8(d, buf[d]);
      }
      bool dead = false;  // every candidate of this warp has fallen below the bound
#pragma unroll 1
      for (int g0 = 0; g0 < ng && !dead; g0 += RF_DEPTH) {
#pragma unroll
        for (int d = 0; d < RF_DEPTH; ++d) {
          const int g = g0 + d;
          if (g < ng) {
            constexpr int ahead = RF_DEPTH - 1;
            const int dn = (d + ahead) % RF_DEPTH;
            fast[dn] = g + ahead < ng && s_flag[g + ahead] == 0;
            if (fast[dn]) load8(g + ahead, buf[dn]);
            if (fast[d]) {
#pragma unroll
              for (int u = 0; u < 8; ++u) s = __dadd_rn(s, buf[d][u]);
            } else {
#pragma unroll 1
              for (int u = 0; u < 8; ++u) {
                const int wd = s_word[8 * g + u];
                s = __dadd_rn(s, wd != RF_SLOW ? __ldg(hp + wd) : slow_value(p0 + 8 * g + u));
              }
            }
          }
        }
        dead = !__any_sync(__activemask(), s >= bound);
      }
      if (dead) {
        s = -INFINITY;  // (below the bound: any value that cannot win will do)
      } else {
#pragma unroll 1
        for (int pl = 8 * ng; pl < pc; ++pl) {
          const int wd = s_word[pl];
          s = __dadd_rn(s, wd != RF_SLOW ? __ldg(hp + wd) : slow_value(p0 + pl));
        }
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
  return block_argmax(hb, s_red);
}

// State of one pair's best-first loop (meaningful in thread 0).
struct BestState {
  double prob;
  float x, y, theta;
  long long n_high, n_pops;
  int status;
};

__device__ __forceinline__ void take_if_better(BestState& st, const ArgBest& htop, const SearchGrid& sg,
                                               int i, int j, float theta) {
  if (htop.idx >= 0 && htop.score > st.prob) {
    const int nhy = sg.n_h[j];
    st.prob = htop.score;
    st.x = sg.t_high[i * sg.n_h_max + htop.idx / nhy];
    st.y = sg.t_high[j * sg.n_h_max + htop.idx % nhy];
    st.theta = theta;
  }
}

// The serial loop of cc:170-243 from the state `st` and the (partly invalidated) memo.
__device__ void best_first_loop(const SearchGrid& sg, const CellMap& cm, const float2* a, int n_pts,
                                const TableMeta& tm, const double* hi, const Rotation* rots,
                                const PairWork& pw, const RefineShared& sh, ArgBest* s_red,
                                int* s_ctl, double* s_bound, BestState& st) {
  const int n_t = sg.n_t, n_tr = n_t * n_t;
  for (;;) {
    const ArgBest top = memo_top(sg, sh.memo, n_tr, s_red);
    if (threadIdx.x == 0) {
      int stop = 0;
      if (top.idx < 0 || top.score < st.prob) stop = 1;
      else if (top.score == -INFINITY && st.prob == -INFINITY) { stop = 1; st.status = 1; }
      if (!stop) { sh.memo[top.idx] = -INFINITY; ++st.n_pops; }
      s_ctl[0] = stop;
      *s_bound = dead_bound(st.prob, n_pts);
    }
    __syncthreads();
    if (s_ctl[0]) break;
    const int i = top.idx / n_t, j = top.idx - i * n_t;  // tr = i*n_t + j, x outer
    const Rotation rot = rots[pw.rot_begin + sh.memo_rot[top.idx]];
    const ArgBest htop = score_window(sg, cm, a, n_pts, tm, hi, rot, i, j, *s_bound, sh, s_red);
    if (threadIdx.x == 0) {
      st.n_high += sg.n_h[i] * sg.n_h[j];
      take_if_better(st, htop, sg, i, j, rot.theta);
    }
    __syncthreads();
  }
}

__device__ __forceinline__ void write_result(PairResult* results, const PairWork& pw, int n_tr,
                                             const BestState& st) {
  PairResult r;
  r.score = st.prob; r.x = st.x; r.y = st.y; r.theta = st.theta;
  r.pad = st.status;
  r.n_low = (long long)pw.n_rot * n_tr;
  r.n_high = st.n_high; r.n_pops = st.n_pops;
  results[pw.out_index] = r;
}

// One kernel, one block per pair: the whole loop in sequence.
__global__ void __launch_bounds__(RF_THREADS) refine_kernel(
    Geometry g, SearchGrid sg, CellMap cm, const float2* __restrict__ points,
    const long long* __restrict__ offsets, const TableSlot* __restrict__ slots,
    const TableMeta* __restrict__ metas, const PairWork* __restrict__ pairs,
    const Rotation* __restrict__ rots, SearchBuffers sb, PairResult* __restrict__ results) {
  const PairWork pw = pairs[blockIdx.x];
  const int n_tr = sg.n_t * sg.n_t;
  extern __shared__ __align__(16) unsigned char s_raw[];
  const RefineShared sh = carve_refine(s_raw, sg.n_h_max, n_tr);
  __shared__ ArgBest s_red[RF_THREADS / 32];
  __shared__ int s_ctl[4];
  __shared__ double s_bound;

  const float2* a = points + offsets[pw.a_scan];
  const int n_pts = (int)(offsets[pw.a_scan + 1] - offsets[pw.a_scan]);
  const TableMeta tm = metas[pw.table];
  const double* __restrict__ hi = slots[pw.table].values;
  reduce_memo(sb, blockIdx.x, pw.n_rot, n_tr, sh.memo, sh.memo_rot);
  __syncthreads();
  BestState st = {-INFINITY, 0.f, 0.f, 0.f, 0, 0, 0};
  best_first_loop(sg, cm, a, n_pts, tm, hi, rots, pw, sh, s_red, s_ctl, &s_bound, st);
  if (threadIdx.x == 0) write_result(results, pw, n_tr, st);
}

// --- speculative form --------------------------------------------------------

__global__ void RF_BOUNDS_FIRST refine_first_kernel(
    Geometry g, SearchGrid sg, CellMap cm, const float2* __restrict__ points,
    const long long* __restrict__ offsets, const TableSlot* __restrict__ slots,
    const TableMeta* __restrict__ metas, const PairWork* __restrict__ pairs,
    const Rotation* __restrict__ rots, SearchBuffers sb) {
  const PairWork pw = pairs[blockIdx.x];
  const int n_t = sg.n_t, n_tr = n_t * n_t;
  extern __shared__ __align__(16) unsigned char s_raw[];
  const RefineShared sh = carve_refine(s_raw, sg.n_h_max, n_tr);
  __shared__ ArgBest s_red[RF_THREADS / 32];
  __shared__ int s_ctl[4];
  __shared__ double s_best;

  const float2* a = points + offsets[pw.a_scan];
  const int n_pts = (int)(offsets[pw.a_scan + 1] - offsets[pw.a_scan]);
  const TableMeta tm = metas[pw.table];
  const double* __restrict__ hi = slots[pw.table].values;
  reduce_memo(sb, blockIdx.x, pw.n_rot, n_tr, sh.memo, sh.memo_rot);
  __syncthreads();

  SpecPair sp;
  sp.best_prob = -INFINITY; sp.x = sp.y = sp.theta = 0.f;
  sp.status = 0; sp.n_list = 0; sp.n_more = 0; sp.item_base = 0; sp.first_tr = -1; sp.n_high = 0;
  const ArgBest top = memo_top(sg, sh.memo, n_tr, s_red);
  if (top.idx < 0 || top.score == -INFINITY) {  // nothing to score (cc:209-212 would abort)
    sp.status = top.idx < 0 ? 0 : 1;
    if (threadIdx.x == 0) sb.spec_pairs[blockIdx.x] = sp;
    return;
  }
  const int i = top.idx / n_t, j = top.idx - i * n_t;
  const Rotation rot = rots[pw.rot_begin + sh.memo_rot[top.idx]];
  const ArgBest htop = score_window(sg, cm, a, n_pts, tm, hi, rot, i, j, -INFINITY, sh, s_red);
  if (threadIdx.x == 0) {
    BestState st = {-INFINITY, 0.f, 0.f, 0.f, 0, 0, 0};
    take_if_better(st, htop, sg, i, j, rot.theta);
    s_best = st.prob;
    sp.best_prob = st.prob; sp.x = st.x; sp.y = st.y; sp.theta = st.theta;
    sp.n_high = sg.n_h[i] * sg.n_h[j];
    sp.first_tr = top.idx;
    s_ctl[1] = 0;
  }
  __syncthreads();
  // slots the reference may still pop: bound >= best1 (and not the one just popped)
  const double b1 = s_best;
  int q = 0;
  for (int tr = threadIdx.x; tr < n_tr; tr += RF_THREADS) q += (tr != top.idx && sh.memo[tr] >= b1);
  if (q) atomicAdd(&s_ctl[1], q);
  __syncthreads();
  const int n_q = s_ctl[1];
  const int n_list = min(n_q, sb.spec_cap);
  if (threadIdx.x == 0) {
    sp.n_list = n_list;
    sp.n_more = n_q - n_list;
    sp.item_base = n_list ? atomicAdd(&sb.spec_counters[0], n_list) : 0;
    s_ctl[2] = sp.item_base;
    sb.spec_pairs[blockIdx.x] = sp;
  }
  __syncthreads();
  if (n_list == 0) return;
  const int base = s_ctl[2];
  // position in pop order: bound descending, slot rank ascending on ties
  for (int tr = threadIdx.x; tr < n_tr; tr += RF_THREADS) {
    const double m = sh.memo[tr];
    if (tr == top.idx || !(m >= b1)) continue;
    const int rk = sg.slot_rank[tr];
    int pos = 0;
    for (int o = 0; o < n_tr; ++o) {
      const double mo = sh.memo[o];
      pos += (o != top.idx) && (mo > m || (mo == m && sg.slot_rank[o] < rk));
    }
    if (pos < n_list) {
      SpecItem it;
      it.pair = blockIdx.x; it.tr = tr; it.rot = sh.memo_rot[tr]; it.pad = 0; it.memo = m;
      sb.spec_items[base + pos] = it;
    }
  }
}

// SMALLWIN: windows of at most RF_THREADS candidates (the 10 x 10 ones of the 0.03 m grid)
// are a single pass of the block, which wants warps more than registers: 8 blocks per SM
// (64 registers).  Larger windows (30 x 30 at 0.01 m) keep the compiler's 80.
// (two entry points over one body: any explicit bound on the large-window kernel measured
// worse than none - 11.9 - 12.5 ms against 11.65 on the 0.01 m configuration)
__device__ __forceinline__ void refine_spec_body(
    const Geometry& g, const SearchGrid& sg, const CellMap& cm, const float2* __restrict__ points,
    const long long* __restrict__ offsets, const TableSlot* __restrict__ slots,
    const TableMeta* __restrict__ metas, const PairWork* __restrict__ pairs,
    const Rotation* __restrict__ rots, const SearchBuffers& sb) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  const RefineShared sh = carve_refine(s_raw, sg.n_h_max, 0);
  __shared__ ArgBest s_red[RF_THREADS / 32];
  __shared__ int s_item;
  const int n_items = sb.spec_counters[0];
  const int n_t = sg.n_t;
  for (;;) {
    if (threadIdx.x == 0) s_item = atomicAdd(&sb.spec_counters[1], 1);
    __syncthreads();
    const int item = s_item;
    if (item >= n_items) break;
    const SpecItem it = sb.spec_items[item];
    const PairWork pw = pairs[it.pair];
    const float2* a = points + offsets[pw.a_scan];
    const int n_pts = (int)(offsets[pw.a_scan + 1] - offsets[pw.a_scan]);
    const TableMeta tm = metas[pw.table];
    const double* __restrict__ hi = slots[pw.table].values;
    const int i = it.tr / n_t, j = it.tr - i * n_t;
    const Rotation rot = rots[pw.rot_begin + it.rot];
    const double bound = dead_bound(sb.spec_pairs[it.pair].best_prob, n_pts);
    const ArgBest htop = score_window(sg, cm, a, n_pts, tm, hi, rot, i, j, bound, sh, s_red);
    if (threadIdx.x == 0) {
      SpecOut o;
      o.score = htop.score; o.idx = htop.idx; o.pad = 0;
      sb.spec_out[item] = o;
    }
    __syncthreads();  // s_item is rewritten next turn
  }
}

template <bool SMALLWIN>
__global__ void refine_spec_kernel(Geometry g, SearchGrid sg, CellMap cm, const float2* __restrict__ points,
                                   const long long* __restrict__ offsets,
                                   const TableSlot* __restrict__ slots,
                                   const TableMeta* __restrict__ metas,
                                   const PairWork* __restrict__ pairs,
                                   const Rotation* __restrict__ rots, SearchBuffers sb);
template <>
__global__ void __launch_bounds__(RF_THREADS, 8) refine_spec_kernel<true>(
    Geometry g, SearchGrid sg, CellMap cm, const float2* __restrict__ points,
    const long long* __restrict__ offsets, const TableSlot* __restrict__ slots,
    const TableMeta* __restrict__ metas, const PairWork* __restrict__ pairs,
    const Rotation* __restrict__ rots, SearchBuffers sb) {
  refine_spec_body(g, sg, cm, points, offsets, slots, metas, pairs, rots, sb);
}
template <>
__global__ void __launch_bounds__(RF_THREADS) refine_spec_kernel<false>(
    Geometry g, SearchGrid sg, CellMap cm, const float2* __restrict__ points,
    const long long* __restrict__ offsets, const TableSlot* __restrict__ slots,
    const TableMeta* __restrict__ metas, const PairWork* __restrict__ pairs,
    const Rotation* __restrict__ rots, SearchBuffers sb) {
  refine_spec_body(g, sg, cm, points, offsets, slots, metas, pairs, rots, sb);
}

__global__ void __launch_bounds__(RF_THREADS) refine_finish_kernel(
    Geometry g, SearchGrid sg, CellMap cm, const float2* __restrict__ points,
    const long long* __restrict__ offsets, const TableSlot* __restrict__ slots,
    const TableMeta* __restrict__ metas, const PairWork* __restrict__ pairs,
    const Rotation* __restrict__ rots, SearchBuffers sb, PairResult* __restrict__ results) {
  const PairWork pw = pairs[blockIdx.x];
  const int n_t = sg.n_t, n_tr = n_t * n_t;
  extern __shared__ __align__(16) unsigned char s_raw[];
  const RefineShared sh = carve_refine(s_raw, sg.n_h_max, n_tr);
  __shared__ ArgBest s_red[RF_THREADS / 32];
  __shared__ int s_ctl[4];
  __shared__ double s_bound;

  const SpecPair sp = sb.spec_pairs[blockIdx.x];
  BestState st = {sp.best_prob, sp.x, sp.y, sp.theta, sp.n_high, sp.first_tr >= 0 ? 1 : 0, sp.status};
  if (threadIdx.x == 0) {
    // the reference's decisions over the listed slots, in pop order
    bool stopped = sp.first_tr < 0;
    for (int k = 0; k < sp.n_list && !stopped; ++k) {
      const SpecItem it = sb.spec_items[sp.item_base + k];
      if (it.memo < st.prob) { stopped = true; break; }
      if (it.memo == -INFINITY && st.prob == -INFINITY) { stopped = true; st.status = 1; break; }
      const int i = it.tr / n_t, j = it.tr - i * n_t;
      ++st.n_pops;
      st.n_high += sg.n_h[i] * sg.n_h[j];
      const SpecOut o = sb.spec_out[sp.item_base + k];
      ArgBest htop = {o.score, o.idx, o.idx};
      take_if_better(st, htop, sg, i, j, rots[pw.rot_begin + it.rot].theta);
    }
    // the list holds every slot with bound >= best1 unless it was cut
    s_ctl[1] = (!stopped && sp.n_more > 0) ? 1 : 0;
  }
  __syncthreads();
  if (s_ctl[1]) {
    const float2* a = points + offsets[pw.a_scan];
    const int n_pts = (int)(offsets[pw.a_scan + 1] - offsets[pw.a_scan]);
    const TableMeta tm = metas[pw.table];
    const double* __restrict__ hi = slots[pw.table].values;
    reduce_memo(sb, blockIdx.x, pw.n_rot, n_tr, sh.memo, sh.memo_rot);
    __syncthreads();
    if (threadIdx.x == 0) sh.memo[sp.first_tr] = -INFINITY;
    for (int k = threadIdx.x; k < sp.n_list; k += RF_THREADS)
      sh.memo[sb.spec_items[sp.item_base + k].tr] = -INFINITY;
    __syncthreads();
    best_first_loop(sg, cm, a, n_pts, tm, hi, rots, pw, sh, s_red, s_ctl, &s_bound, st);
  }
  if (threadIdx.x == 0) write_result(results, pw, n_tr, st);
}

static size_t refine_smem_bytes(int n_t, int nh_max, bool with_memo) {
  size_t b = (size_t)RF_POINT_CHUNK * sizeof(int) + RF_POINT_CHUNK / 8 +
             (size_t)nh_max * nh_max * sizeof(double);
  if (with_memo) b += (size_t)n_t * n_t * (sizeof(double) + sizeof(int));
  return b;
}

int launch_refine(const Geometry& g, const SearchGrid& sg, const CellMap& cm_high,
                  const float2* points, const long long* offsets, const TableSlot* slots,
                  const TableMeta* metas, const PairWork* pairs, const Rotation* rots,
                  int n_pairs, const SearchBuffers& sb, PairResult* results, bool serial,
                  int sm_count, cudaStream_t stream) {
  if (n_pairs <= 0) return 0;
  const size_t smem = refine_smem_bytes(sg.n_t, sg.n_h_max, true);
  static SmemOptIn opt_serial, opt_first, opt_spec, opt_spec_wide, opt_finish;
  if (serial) {
    opt_serial.allow(refine_kernel, smem);
    refine_kernel<<<n_pairs, RF_THREADS, smem, stream>>>(g, sg, cm_high, points, offsets, slots, metas,
                                                         pairs, rots, sb, results);
    return 1;
  }
  const size_t smem_spec = refine_smem_bytes(sg.n_t, sg.n_h_max, false);
  opt_first.allow(refine_first_kernel, smem);
  const bool small_win = sg.n_h_max * sg.n_h_max <= RF_THREADS;
  if (small_win) opt_spec.allow(refine_spec_kernel<true>, smem_spec);
  else opt_spec_wide.allow(refine_spec_kernel<false>, smem_spec);
  opt_finish.allow(refine_finish_kernel, smem);
  cudaMemsetAsync(sb.spec_counters, 0, 2 * sizeof(int), stream);
  refine_first_kernel<<<n_pairs, RF_THREADS, smem, stream>>>(g, sg, cm_high, points, offsets, slots,
                                                             metas, pairs, rots, sb);
  // persistent: as many blocks as stay resident, each fetching windows from the queue
  const int per_sm = small_win ? opt_spec.blocks_per_sm(refine_spec_kernel<true>, RF_THREADS, smem_spec)
                               : opt_spec_wide.blocks_per_sm(refine_spec_kernel<false>, RF_THREADS, smem_spec);
  const long long max_items = (long long)n_pairs * sb.spec_cap;
  const int grid = (int)std::min<long long>((long long)sm_count * per_sm, std::max<long long>(max_items, 1));
  if (small_win)
    refine_spec_kernel<true><<<grid, RF_THREADS, smem_spec, stream>>>(g, sg, cm_high, points, offsets,
                                                                      slots, metas, pairs, rots, sb);
  else
    refine_spec_kernel<false><<<grid, RF_THREADS, smem_spec, stream>>>(g, sg, cm_high, points, offsets,
                                                                       slots, metas, pairs, rots, sb);
  refine_finish_kernel<<<n_pairs, RF_THREADS, smem, stream>>>(g, sg, cm_high, points, offsets, slots,
                                                              metas, pairs, rots, sb, results);
  return 3;
}

}  // namespace csm
