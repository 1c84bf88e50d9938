// Lookup-table construction on sm_100a: rasterise -> Van Vliet IIR blur (x then
// y) -> normalise -> variable-window max-pool.  Replaces
//   CorrelativeScanMatcher::GetLookupTable       (CorrelativeScanMatcher.cc:21-30)
//   LookupTable::GaussianBlur / normalize        (CorrelativeScanMatcher.h:79-87)
//   CImg::vanvliet / _cimg_recursive_apply       (CImg.h:49248-49314, 49062-49115)
//   CImg::normalize                              (CImg.h:39499-39517)
//   CorrelativeScanMatcher::GetLookupTableLowRes (CorrelativeScanMatcher.cc:32-48)
//
// Layout.  A scan lights at most ~1e3 of the 4e6 cells and the filter's impulse
// response falls by 0.716 per cell, so the table is only computed on the
// "active rectangle" = occupied box of the scan grown by `margin` cells and
// clipped to the grid (TableMeta).  Inside it the recursion is the reference's,
// operation for operation; where the rectangle edge coincides with the grid
// edge the Neumann seed / Triggs boundary are applied exactly, and with
// margin < 0 the rectangle is the whole grid and the table is bit-identical.
// The rectangle is stored row-major (x fastest) in the slot's value buffer.
//
// All kernels take a wave of tables: blockIdx.y = table slot in the wave.
#include "csm_common.cuh"
#include "csm_kernels.h"
#include <cstdlib>

namespace csm {

// ---------------------------------------------------------------------------
// K1  rasterise: cell coordinates, occupied box, active rectangle, bitmap.
// One block per table.  The "image" the x-pass filters is never materialised:
// a bit per cell of the occupied box is enough (SetPointValue(p, 1) is
// idempotent and order-free, CorrelativeScanMatcher.cc:23-25).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) raster_kernel(
    Geometry g, const float2* __restrict__ points, const long long* __restrict__ offsets,
    const int* __restrict__ scan_ids, TableSlot* __restrict__ slots,
    TableMeta* __restrict__ metas) {
  const int t = blockIdx.x;
  const int scan = scan_ids[t];
  const long long p0 = offsets[scan], p1 = offsets[scan + 1];
  const TableSlot slot = slots[t];
  __shared__ int s_red[4][8];
  __shared__ int s_box[5];
  __shared__ int s_oldrect[4];  // stored rectangle of the table this slot held before
  if (threadIdx.x == 0) {
    const TableMeta old = metas[t];  // (slot_init_kernel made it "empty" at allocation)
    const bool had = old.cx1 >= old.cx0;
    s_oldrect[0] = had ? old.ax0 : 0; s_oldrect[1] = had ? old.ax1 : -1;
    s_oldrect[2] = had ? old.ay0 : 0; s_oldrect[3] = had ? old.ay1 : -1;
  }

  int lo_x = INT_MAX, hi_x = INT_MIN, lo_y = INT_MAX, hi_y = INT_MIN, cnt = 0;
  for (long long i = p0 + threadIdx.x; i < p1; i += blockDim.x) {
    const float2 p = points[i];
    const long long cx = cell_of(p.x, g.res_h, g.w), cy = cell_of(p.y, g.res_h, g.w);
    if (cx >= 0 && cx < g.w && cy >= 0 && cy < g.w) {
      lo_x = min(lo_x, (int)cx); hi_x = max(hi_x, (int)cx);
      lo_y = min(lo_y, (int)cy); hi_y = max(hi_y, (int)cy);
      ++cnt;
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    lo_x = min(lo_x, __shfl_xor_sync(0xffffffffu, lo_x, o));
    hi_x = max(hi_x, __shfl_xor_sync(0xffffffffu, hi_x, o));
    lo_y = min(lo_y, __shfl_xor_sync(0xffffffffu, lo_y, o));
    hi_y = max(hi_y, __shfl_xor_sync(0xffffffffu, hi_y, o));
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __shared__ int s_cnt[8];
  if (lane == 0) {
    s_red[0][warp] = lo_x; s_red[1][warp] = hi_x;
    s_red[2][warp] = lo_y; s_red[3][warp] = hi_y; s_cnt[warp] = cnt;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int n = 0;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) {
      lo_x = min(lo_x, s_red[0][k]); hi_x = max(hi_x, s_red[1][k]);
      lo_y = min(lo_y, s_red[2][k]); hi_y = max(hi_y, s_red[3][k]);
      n += s_cnt[k];
    }
    TableMeta m;
    m.n_inside = n;
    m.pad = 0;
    if (n == 0) {  // no point inside the grid: the table is all zeros
      m.cx0 = 0; m.cx1 = -1; m.cy0 = 0; m.cy1 = -1;
      m.ax0 = 0; m.ax1 = -1; m.ay0 = 0; m.ay1 = -1;
      m.bx1 = -1; m.by1 = -1;
      m.stride = 0; m.cropped = 1;
    } else {
      m.cx0 = lo_x; m.cx1 = hi_x; m.cy0 = lo_y; m.cy1 = hi_y;
      if (g.margin < 0) {
        m.ax0 = 0; m.ax1 = g.w - 1; m.ay0 = 0; m.ay1 = g.w - 1;
        m.bx1 = m.ax1; m.by1 = m.ay1;
      } else {
        // The recursions run margin cells past the occupied box so that the anti-causal
        // passes start from a negligible state.  What has to be STORED for matching is
        // less: the blur's impulse response g is positive and decreasing, so a cell more
        // than L cells outside the box holds at most n g(L) g(0) while the table maximum
        // is at least g(0)^2 (a point's own cell): with n tail[L] <= 0.25e-10 (factor 4
        // for the mirrored mass of a Neumann edge) it reads as the lookup floor, exactly
        // as in the reference.  The top / left margins feed nothing (the causal passes
        // start at the box), so they shrink to L outright; below / right only the stores do.
        int keep = g.margin;
        if (g.tight && g.tail) {
          for (int d = 0; d <= g.margin; ++d)
            if ((double)n * g.tail[d] <= 0.25e-10) { keep = d; break; }
        }
        m.ax0 = max(0, lo_x - keep); m.ax1 = min(g.w - 1, hi_x + keep);
        m.ay0 = max(0, lo_y - keep); m.ay1 = min(g.w - 1, hi_y + keep);
        m.bx1 = min(g.w - 1, hi_x + g.margin); m.by1 = min(g.w - 1, hi_y + g.margin);
      }
      m.cropped = !(m.ax0 == 0 && m.ay0 == 0 && m.ax1 == g.w - 1 && m.ay1 == g.w - 1);
      m.stride = (m.ax1 - m.ax0 + 1 + 3) & ~3;
    }
    m.lx0 = INT_MAX; m.lx1 = INT_MIN; m.ly0 = INT_MAX; m.ly1 = INT_MIN;
    m.max_key = 0ULL;            // below every key
    m.min_key = ~0ULL;           // above every key
    metas[t] = m;
    s_box[0] = m.cx0; s_box[1] = m.cx1; s_box[2] = m.cy0; s_box[3] = m.cy1; s_box[4] = n;
  }
  __syncthreads();
  const int cx0 = s_box[0], cx1 = s_box[1], cy0 = s_box[2], cy1 = s_box[3];

  // Low-res tables start as "all zero" / "all floor"; the pool kernels fill the cells
  // whose windows meet the stored rectangle.  The slot was all zero / all floor when it
  // was allocated (slot_init_kernel), so only what the PREVIOUS table of this slot wrote
  // has to be cleared: the low-res cells of the windows that met its rectangle.  A window
  // that holds hi-res cell c lands within two low-res cells of c's own (its first
  // coordinate lies less than low_res + high_res before c's); three are cleared.
  if (s_oldrect[1] >= s_oldrect[0] && s_oldrect[3] >= s_oldrect[2]) {
    const double ratio = g.res_h / g.res_l;
    auto low_of = [&](int c, int pad) {
      const int v = (int)floor(((double)c - 0.5 * g.w) * ratio) + g.wl / 2 + pad;
      return max(0, min(g.wl - 1, v));
    };
    const int x0 = low_of(s_oldrect[0] - 1, -3), x1 = low_of(s_oldrect[1] + 1, 3);
    const int y0 = low_of(s_oldrect[2] - 1, -3), y1 = low_of(s_oldrect[3] + 1, 3);
    const int nx = x1 - x0 + 1, ny = y1 - y0 + 1;
    const double floor_log = CSM_FLOOR_LOG;
    for (int i = threadIdx.x; i < nx * ny; i += blockDim.x) {
      const int x = x0 + i % nx, y = y0 + i / nx;
      slot.low[(size_t)y * g.wl + x] = 0.0;
      slot.lowlog[lowlog_index(g.wl, x, y)] = floor_log;
    }
  }
  if (s_box[4] == 0) return;

  const int rows = cy1 - cy0 + 1, words = ((cx1 - cx0) >> 5) + 1;
  for (int i = threadIdx.x; i < rows * words; i += blockDim.x)
    slot.bitmap[(size_t)(i / words) * g.bm_stride + (i % words)] = 0u;
  for (int i = threadIdx.x; i < rows; i += blockDim.x) slot.rowflag[i] = 0;
  __syncthreads();
  for (long long i = p0 + threadIdx.x; i < p1; i += blockDim.x) {
    const float2 p = points[i];
    const long long cx = cell_of(p.x, g.res_h, g.w), cy = cell_of(p.y, g.res_h, g.w);
    if (cx >= 0 && cx < g.w && cy >= 0 && cy < g.w) {
      const int bx = (int)cx - cx0, by = (int)cy - cy0;
      atomicOr(&slot.bitmap[(size_t)by * g.bm_stride + (bx >> 5)], 1u << (bx & 31));
      slot.rowflag[by] = 1;
    }
  }
}

// ---------------------------------------------------------------------------
// K2a  x-pass of CImg::blur -> vanvliet(sigma, 0, 'x') (CImg.h:49286-49291).
// One warp per 32 rows of the occupied box, lane = row.  Rows without a point
// stay identically zero in the reference and are skipped (the y-pass reads
// them as zero).  The causal pass starts at the box's first column: every
// sample left of it is an exact zero in the reference too.  Lanes run their
// serial recursions 32 columns at a time into a shared-memory tile which the
// warp then writes out row-contiguously (256 B per row segment).
// ---------------------------------------------------------------------------
#ifndef XP_WARPS_N
#define XP_WARPS_N 4
#endif
constexpr int XP_WARPS = XP_WARPS_N;

__global__ void __launch_bounds__(XP_WARPS * 32) xpass_stored_kernel(
    Geometry g, VanVliet f, const TableSlot* __restrict__ slots,
    const TableMeta* __restrict__ metas) {
  const int t = blockIdx.y;
  const TableMeta m = metas[t];
  if (m.cx1 < m.cx0) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int group = blockIdx.x * XP_WARPS + warp;
  const int rows = m.cy1 - m.cy0 + 1;
  if (group * 32 >= rows) return;
  const TableSlot slot = slots[t];
  const int row_local = group * 32 + lane;
  const bool has_row = row_local < rows && slot.rowflag[row_local] != 0;
  const unsigned mask = __ballot_sync(0xffffffffu, has_row);
  if (mask == 0u) return;

  __shared__ double s_tile[XP_WARPS][32][33];
  double (*tile)[33] = s_tile[warp];
  double* base = slot.values + (size_t)(m.cy0 + group * 32 - m.ay0) * m.stride;  // row rr adds rr*stride
  const unsigned* bm = slot.bitmap + (size_t)row_local * g.bm_stride;
  const int words = ((m.cx1 - m.cx0) >> 5) + 1;
  const double a1 = f.a1, a2 = f.a2, a3 = f.a3;
  const double sum = __dmul_rn(f.B, f.B);

  // Neumann seed val[1..3] = data[0] / B (CImg.h:49088-49090): data[0] is
  // column 0 of the grid, which only a box starting at column 0 can light.
  double v1 = 0.0, v2 = 0.0, v3 = 0.0;
  if (m.cx0 == 0 && has_row && (bm[0] & 1u)) v1 = v2 = v3 = __ddiv_rn(1.0, f.B);
  // iplus = data[N-1] before the causal pass (CImg.h:49085)
  double iplus = 0.0;
  if (m.cx1 == g.w - 1 && has_row) {
    const int b = g.w - 1 - m.cx0;
    if ((bm[b >> 5] >> (b & 31)) & 1u) iplus = 1.0;
  }

  // ---- causal pass, columns cx0 .. ax1
  const int n_fwd = m.ax1 - m.cx0 + 1;
  for (int c0 = 0; c0 < n_fwd; c0 += 32) {
    const int nk = min(32, n_fwd - c0);
    const unsigned word = (has_row && (c0 >> 5) < words) ? bm[c0 >> 5] : 0u;
#pragma unroll 4
    for (int k = 0; k < nk; ++k) {
      const double x = ((word >> k) & 1u) ? 1.0 : 0.0;
      const double v0 = vv_step(x, v1, v2, v3, a1, a2, a3);
      tile[lane][k] = v0;
      v3 = v2; v2 = v1; v1 = v0;
    }
    __syncwarp();
    if (lane < nk) {
      double* dst = base + (m.cx0 + c0 - m.ax0) + lane;
      for (int rr = 0; rr < 32; ++rr)
        if ((mask >> rr) & 1u) dst[(size_t)rr * m.stride] = tile[rr][lane];
    }
    __syncwarp();
  }

  // ---- boundary: Triggs at the grid edge (exact), zero state at a crop edge
  const bool at_edge = (m.ax1 == g.w - 1);
  double w0 = 0.0;
  if (at_edge) {
    double w1, w2;
    vv_triggs(f, iplus, v1, v2, v3, w0, w1, w2);
    v1 = w0; v2 = w1; v3 = w2;  // after the shift of CImg.h:49100
  } else {
    v1 = v2 = v3 = 0.0;
  }

  // ---- anti-causal pass, columns ax1 .. ax0, chunks aligned to cx0
  // floor division for possibly negative chunk indices
  const int rel_hi = m.ax1 - m.cx0;            // >= 0
  const int rel_lo = m.ax0 - m.cx0;            // <= 0
  const int ch_hi = rel_hi >> 5;
  const int ch_lo = (rel_lo >= 0) ? (rel_lo >> 5) : -((-rel_lo + 31) >> 5);
  for (int ch = ch_hi; ch >= ch_lo; --ch) {
    const int r0 = ch * 32;                     // relative column of tile column 0
    const int k_lo = max(0, rel_lo - r0), k_hi = min(31, rel_hi - r0);
    // load the causal outputs (zero left of the box)
    {
      const int rel = r0 + lane;
      const bool in = (lane >= k_lo && lane <= k_hi);
      const double* src = base + (m.cx0 + rel - m.ax0);
      for (int rr = 0; rr < 32; ++rr) {
        double v = 0.0;
        if (in && rel >= 0 && ((mask >> rr) & 1u)) v = src[(size_t)rr * m.stride];
        tile[rr][lane] = v;
      }
    }
    __syncwarp();
    for (int k = k_hi; k >= k_lo; --k) {
      if (at_edge && r0 + k == rel_hi) {  // data[N-1] = val[0] of the Triggs step
        tile[lane][k] = w0;
        continue;
      }
      const double v0 = vv_step(__dmul_rn(tile[lane][k], sum), v1, v2, v3, a1, a2, a3);
      tile[lane][k] = v0;
      v3 = v2; v2 = v1; v1 = v0;
    }
    __syncwarp();
    if (lane >= k_lo && lane <= k_hi) {
      double* dst = base + (m.cx0 + r0 + lane - m.ax0);
      for (int rr = 0; rr < 32; ++rr)
        if ((mask >> rr) & 1u) dst[(size_t)rr * m.stride] = tile[rr][lane];
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// K2b  y-pass of CImg::blur -> vanvliet(sigma, 0, 'y') (CImg.h:49292-49297),
// fused with the max/min reduction CImg::normalize needs (CImg.h:39505).
// One thread per column of the active rectangle; consecutive threads touch
// consecutive addresses on every row.  The causal pass starts at the occupied
// box's first row (exact zeros above it), runs to the rectangle's last row;
// the anti-causal pass walks back to the rectangle's first row.
// ---------------------------------------------------------------------------
#ifndef YP_THREADS_N
#define YP_THREADS_N 128
#endif
constexpr int YP_THREADS = YP_THREADS_N;
constexpr int YP_B = 8;  // rows loaded ahead of the recursion

__global__ void __launch_bounds__(YP_THREADS) ypass_stored_kernel(
    Geometry g, VanVliet f, const TableSlot* __restrict__ slots,
    TableMeta* __restrict__ metas) {
  const int t = blockIdx.y;
  const TableMeta m = metas[t];
  if (m.cx1 < m.cx0) return;
  const int aw = m.ax1 - m.ax0 + 1;
  if ((int)(blockIdx.x * YP_THREADS) >= aw) return;
  const int col = blockIdx.x * YP_THREADS + threadIdx.x;
  const bool live = col < aw;
  const TableSlot slot = slots[t];
  const double a1 = f.a1, a2 = f.a2, a3 = f.a3;
  const double sum = __dmul_rn(f.B, f.B);
  double* colp = slot.values + (live ? col : 0);
  const unsigned char* flag = slot.rowflag;
  const size_t stride = (size_t)m.stride;

  double vmax = -INFINITY, vmin = INFINITY;
  if (live) {
    // Neumann seed: data[0] = x-pass output on grid row 0 (zero unless the box
    // starts there and that row has a point).
    double v1 = 0.0, v2 = 0.0, v3 = 0.0;
    if (m.cy0 == 0 && flag[0]) {
      const double x0 = colp[(size_t)(0 - m.ay0) * stride];
      v1 = v2 = v3 = __ddiv_rn(x0, f.B);
    }
    double iplus = 0.0;  // x-pass output on the grid's last row
    if (m.cy1 == g.w - 1 && flag[m.cy1 - m.cy0])
      iplus = colp[(size_t)(g.w - 1 - m.ay0) * stride];

    // causal pass: rows cy0 .. ay1.  The recursion is serial but its inputs are
    // not: YP_B rows are loaded first (independent loads in flight), then stepped.
    for (int r0 = m.cy0; r0 <= m.ay1; r0 += YP_B) {
      double xin[YP_B];
#pragma unroll
      for (int u = 0; u < YP_B; ++u) {
        const int r = r0 + u;
        xin[u] = (r <= m.cy1 && flag[r - m.cy0]) ? colp[(size_t)(r - m.ay0) * stride] : 0.0;
      }
#pragma unroll
      for (int u = 0; u < YP_B; ++u) {
        const int r = r0 + u;
        if (r <= m.ay1) {
          const double v0 = vv_step(xin[u], v1, v2, v3, a1, a2, a3);
          colp[(size_t)(r - m.ay0) * stride] = v0;
          v3 = v2; v2 = v1; v1 = v0;
        }
      }
    }
    int r = m.ay1;
    if (m.ay1 == g.w - 1) {  // Triggs boundary at the grid edge
      double w0, w1, w2;
      vv_triggs(f, iplus, v1, v2, v3, w0, w1, w2);
      colp[(size_t)(r - m.ay0) * stride] = w0;
      vmax = fmax(vmax, w0); vmin = fmin(vmin, w0);
      v1 = w0; v2 = w1; v3 = w2;
      --r;
    } else {
      v1 = v2 = v3 = 0.0;
    }
    // anti-causal pass down to the rectangle's first row, same batching
    for (int r0 = r; r0 >= m.ay0; r0 -= YP_B) {
      double yin[YP_B];
#pragma unroll
      for (int u = 0; u < YP_B; ++u) {
        const int rr = r0 - u;
        yin[u] = (rr >= m.cy0) ? colp[(size_t)(rr - m.ay0) * stride] : 0.0;
      }
#pragma unroll
      for (int u = 0; u < YP_B; ++u) {
        const int rr = r0 - u;
        if (rr >= m.ay0) {
          const double v0 = vv_step(__dmul_rn(yin[u], sum), v1, v2, v3, a1, a2, a3);
          colp[(size_t)(rr - m.ay0) * stride] = v0;
          vmax = fmax(vmax, v0); vmin = fmin(vmin, v0);
          v3 = v2; v2 = v1; v1 = v0;
        }
      }
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    vmax = fmax(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    vmin = fmin(vmin, __shfl_xor_sync(0xffffffffu, vmin, o));
  }
  if ((threadIdx.x & 31) == 0 && vmax >= vmin) {
    atomicMax(&metas[t].max_key, order_key(vmax));
    atomicMin(&metas[t].min_key, order_key(vmin));
  }
}

// ---------------------------------------------------------------------------
// K2a' / K2b'  the same two passes WITHOUT the round trip of the causal outputs
// through HBM.  A line's causal pass is a serial recursion whose outputs the
// anti-causal pass consumes in reverse order; the stored variants above write
// them out and read them back (8 + 8 B per cell and pass).  Here the causal pass
// keeps only its state (the last three outputs) at every 32nd sample - a
// checkpoint list in thread-local memory, 24 B per 32 cells - and the
// anti-causal pass rebuilds each 32-sample chunk from its checkpoint just
// before it walks it backwards.  The rebuilt values are the same operations on
// the same operands, hence the same bits.  x-pass: write-only (its input is the
// bitmap); y-pass: reads the x-pass output twice, writes once.  MAXCH = checkpoint
// capacity in 32-sample chunks: 64 for grids up to 2048 cells per line (0.03 m
// tables), 192 for grids up to 6144 (the solver's 0.01 m tables, W = 6000,
// solver.cc:56); longer grids take the stored variants.
// ---------------------------------------------------------------------------
// Global-space accesses of the value buffer (the slot pointers come from memory,
// so the compiler would otherwise emit generic LD / ST).
__device__ __forceinline__ double ld_gbl(const double* p) {
  double v;
  asm volatile("ld.global.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_gbl(double* p, double v) {
  asm volatile("st.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
// Row u of a chunk: p + u * stride doubles with the product taken in 32 bits (a chunk
// spans at most 32 rows of at most 30 000 cells), so that the address is one
// IMAD.WIDE.U32 instead of a 64-bit multiply per access.
__device__ __forceinline__ const double* row_at(const double* p, unsigned stride8, int u) {
  return reinterpret_cast<const double*>(reinterpret_cast<const char*>(p) + stride8 * (unsigned)u);
}
__device__ __forceinline__ double* row_at(double* p, unsigned stride8, int u) {
  return reinterpret_cast<double*>(reinterpret_cast<char*>(p) + stride8 * (unsigned)u);
}

constexpr int RC_MAXCH_SMALL = 64, RC_MAXCH_LARGE = 192;

template <int MAXCH>
__global__ void __launch_bounds__(XP_WARPS * 32) xpass_kernel(
    Geometry g, VanVliet f, const TableSlot* __restrict__ slots,
    const TableMeta* __restrict__ metas) {
  const int t = blockIdx.y;
  const TableMeta m = metas[t];
  if (m.cx1 < m.cx0) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rows = m.cy1 - m.cy0 + 1;
  if (blockIdx.x * XP_WARPS * 32 >= rows) return;  // never more lit rows than rows
  const TableSlot slot = slots[t];

  // Lanes take the rows that hold a point (the others stay exact zeros and are
  // never touched): row bits of the occupied box, their prefix counts, and the
  // k-th lit row for lane k of the block's 128-row share.
  constexpr int WPL = MAXCH / 32;  // row words per lane of the prefix scan
  static_assert(MAXCH % 32 == 0, "the prefix scan gives every lane WPL words");
  __shared__ unsigned s_rows[MAXCH + 1];
  __shared__ int s_pref[MAXCH + 1];
  __shared__ int s_rowidx[XP_WARPS][32];
  for (int w = warp; w <= MAXCH; w += XP_WARPS) {
    const int r = w * 32 + lane;
    const unsigned b = __ballot_sync(0xffffffffu, r < rows && slot.rowflag[r] != 0);
    if (lane == 0) s_rows[w] = b;
  }
  __syncthreads();
  if (warp == 0) {
    int c[WPL], tot = 0;
#pragma unroll
    for (int u = 0; u < WPL; ++u) { c[u] = __popc(s_rows[WPL * lane + u]); tot += c[u]; }
    int incl = tot;
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    int run = incl - tot;
#pragma unroll
    for (int u = 0; u < WPL; ++u) { s_pref[WPL * lane + u] = run; run += c[u]; }
    if (lane == 31) s_pref[MAXCH] = incl;
  }
  __syncthreads();
  const int n_lit = s_pref[MAXCH];
  const int k_row = (blockIdx.x * XP_WARPS + warp) * 32 + lane;
  if (k_row - lane >= n_lit) return;
  const bool has_row = k_row < n_lit;
  int row_local = 0;
  if (has_row) {
    int w = 0, hi = MAXCH;  // s_pref[w] <= k_row < s_pref[hi]
    while (hi - w > 1) {
      const int mid = (w + hi) >> 1;
      if (s_pref[mid] <= k_row) w = mid; else hi = mid;
    }
    row_local = w * 32 + (int)__fns(s_rows[w], 0, k_row - s_pref[w] + 1);
  }
  s_rowidx[warp][lane] = row_local;
  const unsigned mask = __ballot_sync(0xffffffffu, has_row);
  __syncwarp();

  __shared__ double s_tile[XP_WARPS][32][33];
  double (*tile)[33] = s_tile[warp];
  double ck[3 * MAXCH];
  double* base = slot.values + (size_t)(m.cy0 - m.ay0) * m.stride;  // box row 0
  const int* rowidx = s_rowidx[warp];
  const unsigned* bm = slot.bitmap + (size_t)row_local * g.bm_stride;
  const int words = ((m.cx1 - m.cx0) >> 5) + 1;
  const double a1 = f.a1, a2 = f.a2, a3 = f.a3;
  const double sum = __dmul_rn(f.B, f.B);

  double v1 = 0.0, v2 = 0.0, v3 = 0.0;  // Neumann seed, CImg.h:49088-49090
  if (m.cx0 == 0 && has_row && (bm[0] & 1u)) v1 = v2 = v3 = __ddiv_rn(1.0, f.B);
  double iplus = 0.0;                   // data[N-1] before the causal pass, CImg.h:49085
  if (m.cx1 == g.w - 1 && has_row) {
    const int b = g.w - 1 - m.cx0;
    if ((bm[b >> 5] >> (b & 31)) & 1u) iplus = 1.0;
  }

  // ---- causal pass, columns cx0 .. bx1: states only
  const int n_fwd = m.bx1 - m.cx0 + 1;
  for (int c0 = 0, ci = 0; c0 < n_fwd; c0 += 32, ++ci) {
    ck[3 * ci] = v1; ck[3 * ci + 1] = v2; ck[3 * ci + 2] = v3;
    const int nk = min(32, n_fwd - c0);
    const unsigned word = (has_row && ci < words) ? bm[ci] : 0u;
#pragma unroll 8
    for (int k = 0; k < nk; ++k) {
      const double x = ((word >> k) & 1u) ? 1.0 : 0.0;
      const double v0 = vv_step(x, v1, v2, v3, a1, a2, a3);
      v3 = v2; v2 = v1; v1 = v0;
    }
  }

  const bool at_edge = (m.bx1 == g.w - 1);
  double w0 = 0.0;
  if (at_edge) {
    double w1, w2;
    vv_triggs(f, iplus, v1, v2, v3, w0, w1, w2);
    v1 = w0; v2 = w1; v3 = w2;
  } else {
    v1 = v2 = v3 = 0.0;
  }

  // ---- anti-causal pass, columns bx1 .. ax0 (stored from ax1 down), chunks aligned to cx0
  const int rel_hi = m.bx1 - m.cx0;
  const int rel_lo = m.ax0 - m.cx0;
  const int rel_st = m.ax1 - m.cx0;  // last stored column
  const int ch_hi = rel_hi >> 5;
  const int ch_lo = (rel_lo >= 0) ? (rel_lo >> 5) : -((-rel_lo + 31) >> 5);
  for (int ch = ch_hi; ch >= ch_lo; --ch) {
    const int r0 = ch * 32;
    const int k_lo = max(0, rel_lo - r0), k_hi = min(31, rel_hi - r0);
    if (ch >= 0) {  // rebuild the chunk's causal outputs (k_lo == 0 here)
      double f1 = ck[3 * ch], f2 = ck[3 * ch + 1], f3 = ck[3 * ch + 2];
      const unsigned word = (has_row && ch < words) ? bm[ch] : 0u;
#pragma unroll 8
      for (int k = 0; k <= k_hi; ++k) {
        const double x = ((word >> k) & 1u) ? 1.0 : 0.0;
        const double v0 = vv_step(x, f1, f2, f3, a1, a2, a3);
        tile[lane][k] = v0;
        f3 = f2; f2 = f1; f1 = v0;
      }
    }
    for (int k = k_hi; k >= k_lo; --k) {
      if (at_edge && r0 + k == rel_hi) {  // data[N-1] = val[0] of the Triggs step
        tile[lane][k] = w0;
        continue;
      }
      const double y = (ch >= 0) ? tile[lane][k] : 0.0;  // exact zeros left of the box
      const double v0 = vv_step(__dmul_rn(y, sum), v1, v2, v3, a1, a2, a3);
      tile[lane][k] = v0;
      v3 = v2; v2 = v1; v1 = v0;
    }
    __syncwarp();
    if (lane >= k_lo && lane <= k_hi && r0 + lane <= rel_st) {
      double* dst = base + (m.cx0 + r0 + lane - m.ax0);
      for (int rr = 0; rr < 32; ++rr)
        if ((mask >> rr) & 1u) st_gbl(dst + rowidx[rr] * m.stride, tile[rr][lane]);
    }
    __syncwarp();
  }
}

// The YC x-pass outputs of one column for rows r0 .. r0+YC-1 of the grid; `bits`
// has bit u set when row r0+u holds a point (rows without one are exact zeros and
// are not read).
template <int YC>
__device__ __forceinline__ void y_load(double (&a)[YC], const double* p, int stride,
                                       unsigned bits) {
#pragma unroll
  for (int u = 0; u < YC; ++u) a[u] = ((bits >> u) & 1u) ? ld_gbl(row_at(p, 8u * (unsigned)stride, u)) : 0.0;
}

// nrow causal steps over a[0 .. nrow-1]; KEEP: the outputs replace the inputs.
template <int YC, bool FULL, bool KEEP>
__device__ __forceinline__ void y_causal(double (&a)[YC], int nrow, double& v1, double& v2,
                                         double& v3, double a1, double a2, double a3) {
#pragma unroll
  for (int u = 0; u < YC; ++u) {
    if (FULL || u < nrow) {
      const double v0 = vv_step(a[u], v1, v2, v3, a1, a2, a3);
      if (KEEP) a[u] = v0;
      v3 = v2; v2 = v1; v1 = v0;
    }
  }
}

// Anti-causal steps over a[hi .. lo] (causal outputs in, final values stored to
// p + u * stride for u <= hs); edge_u = index whose value is the Triggs w0 (state
// already set), -1 when none.
template <int YC, bool FULL, bool NEEDMIN>
__device__ __forceinline__ void y_anticausal(const double (&a)[YC], int lo, int hi, int hs, int edge_u,
                                             double w0, double* p, int stride, double sum,
                                             double& v1, double& v2, double& v3, double a1,
                                             double a2, double a3, double& vmax, double& vmin) {
#pragma unroll
  for (int u = YC - 1; u >= 0; --u) {
    if (FULL || (u >= lo && u <= hi)) {
      double v0;
      if (!FULL && u == edge_u) {
        v0 = w0;
      } else {
        v0 = vv_step(__dmul_rn(a[u], sum), v1, v2, v3, a1, a2, a3);
        v3 = v2; v2 = v1; v1 = v0;
      }
      if (FULL || u <= hs) st_gbl(row_at(p, 8u * (unsigned)stride, u), v0);  // rows past ay1 are computed only
      vmax = v0 > vmax ? v0 : vmax;  // no NaNs here: a compare and a select instead of fmax
      if (NEEDMIN) vmin = v0 < vmin ? v0 : vmin;
    }
  }
}

// YC = rows per checkpointed chunk (32: 64 loads in flight per thread but ~150
// registers; 16: 3 KB of checkpoints per thread, ~90 registers).  The minimum is
// only tracked when the rectangle is the whole grid (a cropped table normalises
// against 0, see normalize_pool_kernel).
template <int YC, int MINB, int MAXCH>
__global__ void __launch_bounds__(YP_THREADS, MINB) ypass_kernel(
    Geometry g, VanVliet f, const TableSlot* __restrict__ slots,
    TableMeta* __restrict__ metas) {
  const int t = blockIdx.y;
  const TableMeta m = metas[t];
  if (m.cx1 < m.cx0) return;
  const int aw = m.ax1 - m.ax0 + 1;
  if ((int)(blockIdx.x * YP_THREADS) >= aw) return;
  const int col = blockIdx.x * YP_THREADS + threadIdx.x;
  const bool live = col < aw;
  const TableSlot slot = slots[t];
  const double a1 = f.a1, a2 = f.a2, a3 = f.a3;
  const double sum = __dmul_rn(f.B, f.B);
  double* colp = slot.values + (live ? col : 0);
  const unsigned char* flag = slot.rowflag;
  const int stride = m.stride;  // element offsets within a chunk stay 32-bit: one IMAD.WIDE per access

  // rows of the occupied box that hold a point, one bit each (zeros past the box)
  __shared__ unsigned s_rows[MAXCH + 1];
  {
    const int rows = m.cy1 - m.cy0 + 1;
    for (int w = threadIdx.x >> 5; w <= MAXCH; w += YP_THREADS / 32) {
      const int r = w * 32 + (threadIdx.x & 31);
      const unsigned b = __ballot_sync(0xffffffffu, r < rows && flag[r] != 0);
      if ((threadIdx.x & 31) == 0) s_rows[w] = b;
    }
  }
  __syncthreads();
  constexpr unsigned CMASK = YC == 32 ? 0xffffffffu : ((1u << (YC & 31)) - 1u);

  const bool needmin = !m.cropped;
  double vmax = -INFINITY, vmin = INFINITY;
  if (live) {
    double ck[3 * MAXCH * 32 / YC];
    double v1 = 0.0, v2 = 0.0, v3 = 0.0;  // Neumann seed from grid row 0
    if (m.cy0 == 0 && flag[0]) {
      const double x0 = ld_gbl(colp + (size_t)(0 - m.ay0) * stride);
      v1 = v2 = v3 = __ddiv_rn(x0, f.B);
    }
    double iplus = 0.0;  // x-pass output on the grid's last row
    if (m.cy1 == g.w - 1 && flag[m.cy1 - m.cy0])
      iplus = ld_gbl(colp + (size_t)(g.w - 1 - m.ay0) * stride);

    // causal pass, rows cy0 .. by1: states only
    const int n_fwd = m.by1 - m.cy0 + 1;
    const int rel_st = m.ay1 - m.cy0;  // last stored row
    const int nch = (n_fwd + YC - 1) / YC;
    {
      const double* p = colp + (size_t)(m.cy0 - m.ay0) * stride;
      for (int ci = 0; ci < nch; ++ci, p += YC * stride) {
        ck[3 * ci] = v1; ck[3 * ci + 1] = v2; ck[3 * ci + 2] = v3;
        const int b0 = ci * YC;
        const unsigned bits = (s_rows[b0 >> 5] >> (b0 & 31)) & CMASK;
        double a[YC];
        y_load<YC>(a, p, stride, bits);
        const int nrow = n_fwd - b0;
        if (nrow >= YC) y_causal<YC, true, false>(a, YC, v1, v2, v3, a1, a2, a3);
        else y_causal<YC, false, false>(a, nrow, v1, v2, v3, a1, a2, a3);
      }
    }
    const bool at_edge = (m.by1 == g.w - 1);
    double w0 = 0.0;
    if (at_edge) {  // Triggs boundary at the grid edge
      double w1, w2;
      vv_triggs(f, iplus, v1, v2, v3, w0, w1, w2);
      v1 = w0; v2 = w1; v3 = w2;
    } else {
      v1 = v2 = v3 = 0.0;
    }
    // anti-causal pass down to the rectangle's first row, chunk by chunk
    const int rel_lo = m.ay0 - m.cy0;  // <= 0
    const int ch_lo = -((-rel_lo + YC - 1) / YC);
    double* p = colp + (size_t)(m.cy0 + (nch - 1) * YC - m.ay0) * stride;
    for (int ch = nch - 1; ch >= ch_lo; --ch, p -= YC * stride) {
      const int b0 = ch * YC;  // row of a[0] relative to cy0
      double a[YC];
      if (ch >= 0) {  // rebuild the chunk's causal outputs from its checkpoint
        const unsigned bits = (s_rows[b0 >> 5] >> (b0 & 31)) & CMASK;
        y_load<YC>(a, p, stride, bits);
        double f1 = ck[3 * ch], f2 = ck[3 * ch + 1], f3 = ck[3 * ch + 2];
        const int nrow = n_fwd - b0;
        if (nrow >= YC) y_causal<YC, true, true>(a, YC, f1, f2, f3, a1, a2, a3);
        else y_causal<YC, false, true>(a, nrow, f1, f2, f3, a1, a2, a3);
      } else {
#pragma unroll
        for (int u = 0; u < YC; ++u) a[u] = 0.0;  // exact zeros above the box
      }
      const int lo = max(0, rel_lo - b0), hi = min(YC - 1, n_fwd - 1 - b0);
      const int hs = min(hi, rel_st - b0);  // rows of this chunk that are stored: lo .. hs
      const bool last = at_edge && ch == nch - 1;  // the chunk that holds the grid's last row
      if (lo == 0 && hs == YC - 1 && !last) {
        if (needmin)
          y_anticausal<YC, true, true>(a, 0, YC - 1, YC - 1, -1, w0, p, stride, sum, v1, v2, v3, a1,
                                       a2, a3, vmax, vmin);
        else
          y_anticausal<YC, true, false>(a, 0, YC - 1, YC - 1, -1, w0, p, stride, sum, v1, v2, v3, a1,
                                        a2, a3, vmax, vmin);
      } else {
        y_anticausal<YC, false, true>(a, lo, hi, hs, last ? hi : -1, w0, p, stride, sum, v1, v2, v3,
                                      a1, a2, a3, vmax, vmin);
      }
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    vmax = fmax(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    vmin = fmin(vmin, __shfl_xor_sync(0xffffffffu, vmin, o));
  }
  if ((threadIdx.x & 31) == 0 && vmax > -INFINITY) {
    atomicMax(&metas[t].max_key, order_key(vmax));
    if (needmin) atomicMin(&metas[t].min_key, order_key(vmin));
  }
}

// ---------------------------------------------------------------------------
// K2b'  two-run y-pass for matching tables (store_log, cropped or not, windows that
// tile the axis).  The one-run form above writes the blurred values, and the normalise
// pass then reads them back and writes their logs: 8 + 8 + 8 B per cell on top of the
// y recursion's own reads.  The divisor of the normalisation is the maximum of the y-pass's
// own output, so no single sweep can write final values; but a second RUN of the anti-causal
// sweep can, and it costs less than the round trip:
//   run 1 (YP2_MAX): causal sweep + anti-causal sweep as above, NOTHING stored but the
//          causal checkpoints (global scratch, TableSlot::ck); tracks the table maximum
//          and max-pools the raw values per low-res window (GetLookupTableLowRes,
//          cc:32-48) with one atomicMax per (warp, window) on order-preserving keys -
//          the maximum commutes with the monotone normalisation, as in
//          normalize_pool_kernel.
//   run 2 (YP2_LOG): rebuilds each chunk from run 1's checkpoints, repeats the
//          anti-causal sweep (same operations, same operands: same bits) and stores
//          lookup_log((v - m) / (M - m)) in place.
// low_finalize_kernel then turns the pooled keys into low / lowlog.  Every stored value
// is the one the one-run path + normalize_pool_kernel produce, bit for bit.
// Measured (B200, 2 000 tables of the bench / 400 of the 0.01 m configuration): 9.05 ms
// against 9.00 for the one-run path / 10.9 against 12.0, with 14.5 MB of DRAM traffic per
// table instead of 19.1.  Not kept: the same two runs with every column streamed through
// a cp.async ring in shared memory, 2 - 8 chunks deep (11.1 - 24.6 ms): the sweeps are
// bound by instruction issue and the dependent fp64 chain at 4 - 5 warps per scheduler,
// not by the latency of their loads, and the ring adds instructions while the shared
// memory it takes removes warps.
// ---------------------------------------------------------------------------

// ck layout: [chunk][3][column], chunk nch = the causal sweep's final state
__device__ __forceinline__ size_t ck_index(int chunk, int k, int col, int ckstride) {
  return ((size_t)chunk * 3 + k) * (size_t)ckstride + col;
}

constexpr int YP2_MAX = 1, YP2_LOG = 2;
// Resident blocks per SM the two runs are compiled for.  Both fit 80 registers without
// spilling (run 1 even 66); measured table stage of the bench, run 1 / run 2 at
// 5/5, 6/5, 6/6, 7/6, 8/6 blocks: 8.65 / 8.49 / 8.27 / 8.26 / 8.24 ms.
#ifndef YP2_MINB1_N
#define YP2_MINB1_N 6
#endif
#ifndef YP2_MINB2_N
#define YP2_MINB2_N 6
#endif
constexpr int YP2_MINB1 = YP2_MINB1_N, YP2_MINB2 = YP2_MINB2_N;
#ifndef YP1_MINB_N    // the one-run y-pass (6 and 7 blocks measured: no gain)
#define YP1_MINB_N 5
#endif
constexpr int YP1_MINB = YP1_MINB_N;
constexpr int YP2_MAXWIN = 1024;  // pooling windows per axis the shared copies hold

// Anti-causal steps over a[hi .. lo] (causal outputs in).  Run 1 folds the final values
// into the maxima - the table's, and the pooling window's (rows <= hs; a window is
// complete after its first row, bit u of wb, and moves to cap0, the one before to
// cap1); run 2 leaves them in a[] for the normalising stores.
template <int YC, int MODE, bool FULL, bool NEEDMIN>
__device__ __forceinline__ void y2_anticausal(double (&a)[YC], int lo, int hi, int hs, int edge_u,
                                              double w0, double sum, double& v1, double& v2,
                                              double& v3, double a1, double a2, double a3,
                                              double& vmax, double& vmin, double& dm, unsigned wb,
                                              double& cap0, double& cap1) {
#pragma unroll
  for (int u = YC - 1; u >= 0; --u) {
    if (FULL || (u >= lo && u <= hi)) {
      double v0;
      if (!FULL && u == edge_u) {
        v0 = w0;
      } else {
        v0 = vv_step(__dmul_rn(a[u], sum), v1, v2, v3, a1, a2, a3);
        v3 = v2; v2 = v1; v1 = v0;
      }
      if (MODE == YP2_MAX) {
        // (the table maximum is folded from the window maxima at their flush: the rows
        // past hs that take no part in them are below the lookup floor by construction)
        if (NEEDMIN) vmin = v0 < vmin ? v0 : vmin;
        if (FULL || u <= hs) {
          dm = v0 > dm ? v0 : dm;
          if ((wb >> u) & 1u) { cap1 = cap0; cap0 = dm; dm = -INFINITY; }
        }
      } else {
        a[u] = v0;
      }
    }
  }
}

template <int YC, int MINB, int MAXCH, int MODE>
__global__ void __launch_bounds__(YP_THREADS, MINB) ypass2_kernel(
    Geometry g, VanVliet f, const TableSlot* __restrict__ slots, TableMeta* __restrict__ metas,
    const int* __restrict__ pool_s, const int* __restrict__ pool_dst,
    const unsigned char* __restrict__ pool_win) {
  const int t = blockIdx.y;
  const TableMeta m = metas[t];
  if (m.cx1 < m.cx0) return;
  const int aw = m.ax1 - m.ax0 + 1;
  if ((int)(blockIdx.x * YP_THREADS) >= aw) return;
  const int col_raw = blockIdx.x * YP_THREADS + threadIdx.x;
  const bool live = col_raw < aw;
  // run 1 shuffles across the warp: a lane past the rectangle recomputes the last column
  // (reads only) instead of leaving the loop
  const int col = live ? col_raw : aw - 1;
  const TableSlot slot = slots[t];
  const double a1 = f.a1, a2 = f.a2, a3 = f.a3;
  const double sum = __dmul_rn(f.B, f.B);
  double* colp = slot.values + col;
  double* ckp = slot.ck;
  const int ckstride = m.stride;
  const unsigned char* flag = slot.rowflag;
  const int stride = m.stride;  // element offsets within a chunk stay 32-bit: one IMAD.WIDE per access
  const int lane = threadIdx.x & 31;

  __shared__ unsigned s_rows[MAXCH + 1];
  __shared__ int s_ps[MODE == YP2_MAX ? YP2_MAXWIN : 1];
  __shared__ int s_pd[MODE == YP2_MAX ? YP2_MAXWIN : 1];  // destination row, -1: window not pooled
  __shared__ double s_out[MODE == YP2_LOG ? YC : 1][YP_THREADS];
  {
    const int rows = m.cy1 - m.cy0 + 1;
    for (int w = threadIdx.x >> 5; w <= MAXCH; w += YP_THREADS / 32) {
      const int r = w * 32 + lane;
      const unsigned b = __ballot_sync(0xffffffffu, r < rows && flag[r] != 0);
      if (lane == 0) s_rows[w] = b;
    }
    if (MODE == YP2_MAX)
      for (int k = threadIdx.x; k < g.n_pool; k += YP_THREADS) {
        s_ps[k] = pool_s[k];
        s_pd[k] = pool_win[k] ? pool_dst[k] : -1;
      }
  }
  __syncthreads();
  if ((int)(blockIdx.x * YP_THREADS + (threadIdx.x & ~31)) >= aw) return;  // whole warp past the rectangle
  constexpr unsigned CMASK = YC == 32 ? 0xffffffffu : ((1u << (YC & 31)) - 1u);

  const int n_fwd = m.by1 - m.cy0 + 1;
  const int rel_st = m.ay1 - m.cy0;  // last stored row
  const int nch = (n_fwd + YC - 1) / YC;
  const bool at_edge = (m.by1 == g.w - 1);
  const int rel_lo = m.ay0 - m.cy0;  // <= 0
  const int ch_lo = -((-rel_lo + YC - 1) / YC);

  double v1 = 0.0, v2 = 0.0, v3 = 0.0;
  if (MODE == YP2_MAX) {
    if (m.cy0 == 0 && flag[0]) {  // Neumann seed from grid row 0
      const double x0 = ld_gbl(colp + (size_t)(0 - m.ay0) * stride);
      v1 = v2 = v3 = __ddiv_rn(x0, f.B);
    }
    const double* p = colp + (size_t)(m.cy0 - m.ay0) * stride;
    for (int ci = 0; ci < nch; ++ci, p += YC * stride) {
      if (live) {
        ckp[ck_index(ci, 0, col, ckstride)] = v1;
        ckp[ck_index(ci, 1, col, ckstride)] = v2;
        ckp[ck_index(ci, 2, col, ckstride)] = v3;
      }
      const int b0 = ci * YC;
      const unsigned bits = (s_rows[b0 >> 5] >> (b0 & 31)) & CMASK;
      double a[YC];
      y_load<YC>(a, p, stride, bits);
      const int nrow = n_fwd - b0;
      if (nrow >= YC) y_causal<YC, true, false>(a, YC, v1, v2, v3, a1, a2, a3);
      else y_causal<YC, false, false>(a, nrow, v1, v2, v3, a1, a2, a3);
    }
    if (live) {
      ckp[ck_index(nch, 0, col, ckstride)] = v1;
      ckp[ck_index(nch, 1, col, ckstride)] = v2;
      ckp[ck_index(nch, 2, col, ckstride)] = v3;
    }
  } else {
    v1 = ckp[ck_index(nch, 0, col, ckstride)];
    v2 = ckp[ck_index(nch, 1, col, ckstride)];
    v3 = ckp[ck_index(nch, 2, col, ckstride)];
  }
  double w0 = 0.0;
  if (at_edge) {  // Triggs boundary at the grid edge
    double iplus = 0.0;  // x-pass output on the grid's last row
    if (m.cy1 == g.w - 1 && flag[m.cy1 - m.cy0])
      iplus = ld_gbl(colp + (size_t)(g.w - 1 - m.ay0) * stride);
    double w1, w2;
    vv_triggs(f, iplus, v1, v2, v3, w0, w1, w2);
    v1 = w0; v2 = w1; v3 = w2;
  } else {
    v1 = v2 = v3 = 0.0;
  }

  // ---- run 1: pooling state.  Columns: the window of this thread's column and which
  // neighbours share it (segmented warp maximum); rows: the window of the row being
  // walked, a window is flushed after its first row (rows descend).
  int kx_dst = -1;
  unsigned same = 0;  // bit j: lane + (1 << j) holds a column of the same window
  bool head = false;
  int ky = 0;
  if (MODE == YP2_MAX) {
    int lo = 0, hi = g.n_pool - 1;  // last window with pool_s <= column
    const int cabs = m.ax0 + col;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (s_ps[mid] <= cabs) lo = mid; else hi = mid - 1;
    }
    const int kx = live ? lo : -1 - lane;  // lanes past the rectangle pool nothing
    kx_dst = live ? s_pd[lo] : -1;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      const int o = __shfl_down_sync(0xffffffffu, kx, 1 << j);
      if (lane + (1 << j) < 32 && o == kx) same |= 1u << j;
    }
    const int prev = __shfl_up_sync(0xffffffffu, kx, 1);
    head = lane == 0 || prev != kx;
    lo = 0; hi = g.n_pool - 1;  // window of the last stored row
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (s_ps[mid] <= m.ay1) lo = mid; else hi = mid - 1;
    }
    ky = lo;
  }
  unsigned long long* lowkeys = reinterpret_cast<unsigned long long*>(slot.low);
  double vmax = -INFINITY;
  auto flush = [&](double val, int kyw) {
    vmax = val > vmax ? val : vmax;
    const int dy = s_pd[kyw];
    if (dy < 0) return;  // uniform
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      const double o = __shfl_down_sync(0xffffffffu, val, 1 << j);
      if ((same >> j) & 1u) val = o > val ? o : val;
    }
    if (head && kx_dst >= 0 && val > -INFINITY)
      atomicMax(&lowkeys[(size_t)dy * g.wl + kx_dst], order_key(val));
  };

  // ---- run 2: normalisation constants
  double mn = 0.0, range = 1.0, sub_floor = 0.0, rinv = 1.0;
  bool constant = false;
  if (MODE == YP2_LOG) {
    const double M = key_to_double(m.max_key);
    mn = m.cropped ? 0.0 : key_to_double(m.min_key);
    constant = (M == mn);
    range = __dsub_rn(M, mn);
    sub_floor = __dmul_rn(range, CSM_MIN_VALUE_FOR_LOOKUP * (1.0 - 1e-9));
    rinv = __ddiv_rn(1.0, range);
  }

  const bool needmin = !m.cropped;
  double vmin = INFINITY, dm = -INFINITY;
  double* p = colp + (size_t)(m.cy0 + (nch - 1) * YC - m.ay0) * stride;
  for (int ch = nch - 1; ch >= ch_lo; --ch, p -= YC * stride) {
    const int b0 = ch * YC;  // row of a[0] relative to cy0
    double a[YC];
    if (ch >= 0) {  // rebuild the chunk's causal outputs from its checkpoint
      const unsigned bits = (s_rows[b0 >> 5] >> (b0 & 31)) & CMASK;
      y_load<YC>(a, p, stride, bits);
      double f1 = ckp[ck_index(ch, 0, col, ckstride)], f2 = ckp[ck_index(ch, 1, col, ckstride)],
             f3 = ckp[ck_index(ch, 2, col, ckstride)];
      const int nrow = n_fwd - b0;
      if (nrow >= YC) y_causal<YC, true, true>(a, YC, f1, f2, f3, a1, a2, a3);
      else y_causal<YC, false, true>(a, nrow, f1, f2, f3, a1, a2, a3);
    } else {
#pragma unroll
      for (int u = 0; u < YC; ++u) a[u] = 0.0;  // exact zeros above the box
    }
    const int lo = max(0, rel_lo - b0), hi = min(YC - 1, n_fwd - 1 - b0);
    const int hs = min(hi, rel_st - b0);  // rows of this chunk that are stored: lo .. hs
    const int edge_u = (at_edge && ch == nch - 1) ? hi : -1;
    // window starts among this chunk's rows (run 1): bit u = grid row cy0 + b0 + u starts a window
    unsigned wb = 0;
    if (MODE == YP2_MAX) {
      const int r0 = m.cy0 + b0;
      for (int k = ky; k >= 0 && s_ps[k] >= r0 + lo; --k)
        if (s_ps[k] <= r0 + hs) wb |= 1u << (s_ps[k] - r0);
    }
    double cap0 = -INFINITY, cap1 = -INFINITY;
    if (lo == 0 && hs == YC - 1 && edge_u < 0) {
      if (needmin)
        y2_anticausal<YC, MODE, true, true>(a, 0, YC - 1, YC - 1, -1, w0, sum, v1, v2, v3, a1, a2, a3,
                                            vmax, vmin, dm, wb, cap0, cap1);
      else
        y2_anticausal<YC, MODE, true, false>(a, 0, YC - 1, YC - 1, -1, w0, sum, v1, v2, v3, a1, a2, a3,
                                             vmax, vmin, dm, wb, cap0, cap1);
    } else {
      y2_anticausal<YC, MODE, false, true>(a, lo, hi, hs, edge_u, w0, sum, v1, v2, v3, a1, a2, a3, vmax,
                                           vmin, dm, wb, cap0, cap1);
    }
    if (MODE == YP2_MAX) {
      const int nb = __popc(wb);  // at most 2: windows are at least YC / 2 rows (host check)
      if (nb == 2) { flush(cap1, ky); flush(cap0, ky - 1); }
      else if (nb == 1) flush(cap0, ky);
      ky -= nb;
    } else {
      // the divisions and logs run in a ROLLED loop over a shared copy of the chunk (a
      // thread reads back its own words only): unrolled, the 16 inlined logs of the
      // three sweep variants outgrow the instruction cache (ncu: no_instruction stalls)
#pragma unroll
      for (int u = 0; u < YC; ++u) s_out[u][threadIdx.x] = a[u];
      if (live) {
#pragma unroll 1
        for (int u = lo; u <= hs; ++u) {
          const double d = __dsub_rn(s_out[u][threadIdx.x], mn);
          double out = CSM_FLOOR_LOG;
          if (!(constant || d < sub_floor)) out = lookup_log(div_by(d, range, rinv));
          st_gbl(row_at(p, 8u * (unsigned)stride, u), out);
        }
      }
    }
  }
  if (MODE == YP2_MAX) {
    // the window cut by the rectangle's first row (nothing pending right after a flush)
    if (ky >= 0 && __any_sync(0xffffffffu, dm > -INFINITY)) flush(dm, ky);
    for (int o = 16; o > 0; o >>= 1) {
      vmax = fmax(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
      vmin = fmin(vmin, __shfl_xor_sync(0xffffffffu, vmin, o));
    }
    if (lane == 0 && vmax > -INFINITY) {
      atomicMax(&metas[t].max_key, order_key(vmax));
      if (needmin) atomicMin(&metas[t].min_key, order_key(vmin));
    }
  }
}

// Pooled keys -> low / lowlog (the tail of normalize_pool_kernel): one block per table,
// over the windows that meet the stored rectangle.  A cell no column reached keeps its
// 0.0 (bit pattern 0, below every key).
__global__ void __launch_bounds__(256) low_finalize_kernel(
    Geometry g, const TableSlot* __restrict__ slots, TableMeta* __restrict__ metas,
    const int* __restrict__ pool_s, const int* __restrict__ pool_e,
    const int* __restrict__ pool_dst, const unsigned char* __restrict__ pool_win) {
  const int t = blockIdx.x;
  const TableMeta m = metas[t];
  if (m.cx1 < m.cx0) return;
  const TableSlot slot = slots[t];
  const double M = key_to_double(m.max_key);
  const double mn = m.cropped ? 0.0 : key_to_double(m.min_key);
  const bool constant = (M == mn);
  const double range = __dsub_rn(M, mn);
  // windows that meet the rectangle: [kx0, kx1] x [ky0, ky1] (the windows tile the axis)
  __shared__ int s_k[4];
  if (threadIdx.x < 4) {
    const int target = threadIdx.x == 0 ? m.ax0 : threadIdx.x == 1 ? m.ax1 : threadIdx.x == 2 ? m.ay0 : m.ay1;
    int lo = 0, hi = g.n_pool - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (pool_s[mid] <= target) lo = mid; else hi = mid - 1;
    }
    s_k[threadIdx.x] = lo;
  }
  __syncthreads();
  const int kx0 = s_k[0], nx = s_k[1] - kx0 + 1, ky0 = s_k[2], ny = s_k[3] - ky0 + 1;
  const unsigned long long* lowkeys = reinterpret_cast<const unsigned long long*>(slot.low);
  int bx0 = INT_MAX, bx1 = INT_MIN, by0 = INT_MAX, by1 = INT_MIN;
  for (int i = threadIdx.x; i < nx * ny; i += blockDim.x) {
    const int kx = kx0 + i % nx, ky = ky0 + i / nx;
    const int dx = pool_dst[kx], dy = pool_dst[ky];
    if (!pool_win[kx] || !pool_win[ky] || dx < 0 || dy < 0) continue;
    if (max(pool_s[kx], m.ax0) > min(pool_e[kx], m.ax1)) continue;
    if (max(pool_s[ky], m.ay0) > min(pool_e[ky], m.ay1)) continue;
    const size_t at = (size_t)dy * g.wl + dx;
    const unsigned long long key = lowkeys[at];
    double mx = 0.0;  // cells outside the grid (Dirichlet) / outside the rectangle
    if (key != 0ULL && !constant) mx = fmax(mx, __ddiv_rn(__dsub_rn(key_to_double(key), mn), range));
    slot.low[at] = mx;
    slot.lowlog[lowlog_index(g.wl, dx, dy)] = lookup_log(mx);
    if (mx > CSM_MIN_VALUE_FOR_LOOKUP) {
      bx0 = min(bx0, dx); bx1 = max(bx1, dx);
      by0 = min(by0, dy); by1 = max(by1, dy);
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    bx0 = min(bx0, __shfl_xor_sync(0xffffffffu, bx0, o));
    bx1 = max(bx1, __shfl_xor_sync(0xffffffffu, bx1, o));
    by0 = min(by0, __shfl_xor_sync(0xffffffffu, by0, o));
    by1 = max(by1, __shfl_xor_sync(0xffffffffu, by1, o));
  }
  if ((threadIdx.x & 31) == 0 && bx0 <= bx1) {
    atomicMin(&metas[t].lx0, bx0); atomicMax(&metas[t].lx1, bx1);
    atomicMin(&metas[t].ly0, by0); atomicMax(&metas[t].ly1, by1);
  }
}

// ---------------------------------------------------------------------------
// K3+K4  normalise (CImg::normalize(0,1), CImg.h:39499-39517) fused with
// GetLookupTableLowRes (CorrelativeScanMatcher.cc:32-48).
//
// The reference walks x,y over the low grid with an accumulating double and
// pools hi-res cells [cX(x), cX(x+low_res)-1]: windows of 9-11 cells that tile
// the axis (the end of window k is the start of window k+1 because the loop's
// next x IS x+low_res).  The 1-D window list is built on the host with the
// reference's own expressions (PoolList); blockIdx.x = window index along y.
// Each block normalises the rows of its window in place, keeps the per-column
// maxima in shared memory, then reduces them per x-window.  Normalisation is
// monotone, so max(normalise(v)) == normalise(max(v)) bit for bit.
// ---------------------------------------------------------------------------
constexpr int NP_B = 6;  // rows of a pooling window loaded ahead (windows are 9-11 rows)

// normalised value = (v - m) / (M - m) * (1 - 0) + 0 ; the last two operations
// are exact no-ops

__global__ void __launch_bounds__(256) normalize_pool_kernel(
    Geometry g, const TableSlot* __restrict__ slots, TableMeta* __restrict__ metas,
    const int* __restrict__ pool_s, const int* __restrict__ pool_e,
    const int* __restrict__ pool_dst, const unsigned char* __restrict__ pool_win,
    int store_log) {
  extern __shared__ double s_colmax[];
  const int t = blockIdx.y;
  const TableMeta m = metas[t];
  if (m.cx1 < m.cx0) return;
  const int ky = blockIdx.x;
  const int rs = max(pool_s[ky], m.ay0), re = min(pool_e[ky], m.ay1);
  if (rs > re) return;
  const TableSlot slot = slots[t];
  const double M = key_to_double(m.max_key);
  const double mn = m.cropped ? 0.0 : key_to_double(m.min_key);
  const bool constant = (M == mn);  // CImg fills with min_value (0) in that case
  const double range = __dsub_rn(M, mn);
  const int aw = m.ax1 - m.ax0 + 1;
  // v -> (v - mn) / range is monotone (subtraction and division by a positive
  // number round monotonically), so the window maximum is normalised once per
  // column instead of once per cell, and a cell whose difference lies clearly
  // below range * 1e-10 reads as the lookup floor without the division and the
  // log: d < range * 1e-10 * (1 - 1e-9) => d / range rounds to <= 1e-10.
  const double sub_floor = __dmul_rn(range, CSM_MIN_VALUE_FOR_LOOKUP * (1.0 - 1e-9));
  const double rinv = constant ? 1.0 : __ddiv_rn(1.0, range);
  for (int c = threadIdx.x; c < aw; c += blockDim.x) {
    double dm = -INFINITY;
    double* p = slot.values + (size_t)(rs - m.ay0) * m.stride + c;
    for (int r0 = rs; r0 <= re; r0 += NP_B, p += NP_B * m.stride) {
      double v[NP_B];  // the rows of a window are independent: all loads first
#pragma unroll
      for (int u = 0; u < NP_B; ++u) v[u] = (r0 + u <= re) ? ld_gbl(p + u * m.stride) : 0.0;
#pragma unroll
      for (int u = 0; u < NP_B; ++u) {
        if (r0 + u <= re) {
          const double d = __dsub_rn(v[u], mn);
          dm = d > dm ? d : dm;  // no NaNs: compare + select
          // matching only ever needs log(GetPointValue): take it once per cell here
          // instead of once per (candidate, point) in the search
          double out;
          if (store_log) out = (constant || d < sub_floor) ? CSM_FLOOR_LOG : lookup_log(div_by(d, range, rinv));
          else out = constant ? 0.0 : __ddiv_rn(d, range);
          st_gbl(p + u * m.stride, out);
        }
      }
    }
    s_colmax[c] = constant ? 0.0 : __ddiv_rn(dm, range);
  }
  __syncthreads();
  const int dy = pool_dst[ky];
  if (!pool_win[ky] || dy < 0) return;
  int bx0 = INT_MAX, bx1 = INT_MIN;
  for (int kx = threadIdx.x; kx < g.n_pool; kx += blockDim.x) {
    const int dx = pool_dst[kx];
    if (!pool_win[kx] || dx < 0) continue;
    const int cs = max(pool_s[kx], m.ax0), ce = min(pool_e[kx], m.ax1);
    if (cs > ce) continue;
    double mx = 0.0;  // cells outside the grid (Dirichlet) / outside the rectangle
    for (int c = cs; c <= ce; ++c) mx = fmax(mx, s_colmax[c - m.ax0]);
    slot.low[(size_t)dy * g.wl + dx] = mx;
    const bool above = mx > CSM_MIN_VALUE_FOR_LOOKUP;
    slot.lowlog[lowlog_index(g.wl, dx, dy)] = lookup_log(mx);
    if (above) { bx0 = min(bx0, dx); bx1 = max(bx1, dx); }
  }
  // box of low-res cells above the lookup floor (the search stages it in smem)
  for (int o = 16; o > 0; o >>= 1) {
    bx0 = min(bx0, __shfl_xor_sync(0xffffffffu, bx0, o));
    bx1 = max(bx1, __shfl_xor_sync(0xffffffffu, bx1, o));
  }
  if ((threadIdx.x & 31) == 0 && bx0 <= bx1) {
    atomicMin(&metas[t].lx0, bx0); atomicMax(&metas[t].lx1, bx1);
    atomicMin(&metas[t].ly0, dy); atomicMax(&metas[t].ly1, dy);
  }
}

// Generic GetLookupTableLowRes for caller-made tables (any width/height/values,
// no normalisation): one thread per (kx, ky) window pair, CImg::get_crop
// semantics (zero fill outside the image takes part in the max).  Launched in
// loop order position so that "later window overwrites earlier" is resolved by
// the host-computed winner flags.
__global__ void pool_generic_kernel(const double* __restrict__ hi, int w, int h,
                                    const int* __restrict__ xs, const int* __restrict__ xe,
                                    const int* __restrict__ xdst,
                                    const unsigned char* __restrict__ xwin, int nx,
                                    const int* __restrict__ ys, const int* __restrict__ ye,
                                    const int* __restrict__ ydst,
                                    const unsigned char* __restrict__ ywin, int ny,
                                    double* __restrict__ low, int wl) {
  const int kx = blockIdx.x * blockDim.x + threadIdx.x;
  const int ky = blockIdx.y;
  if (kx >= nx || ky >= ny) return;
  if (!xwin[kx] || !ywin[ky] || xdst[kx] < 0 || ydst[ky] < 0) return;
  const int x0 = xs[kx], x1 = xe[kx], y0 = ys[ky], y1 = ye[ky];
  const bool outside = x0 < 0 || y0 < 0 || x1 >= w || y1 >= h;
  double mx = outside ? 0.0 : -INFINITY;
  for (int y = max(y0, 0); y <= min(y1, h - 1); ++y)
    for (int x = max(x0, 0); x <= min(x1, w - 1); ++x)
      mx = fmax(mx, hi[(size_t)y * w + x]);
  low[(size_t)ydst[ky] * wl + xdst[kx]] = mx;
}

// Expands a slot's active rectangle into a dense w*w table (zeros outside).
__global__ void expand_kernel(Geometry g, const TableSlot* __restrict__ slots,
                              const TableMeta* __restrict__ metas, int t,
                              double* __restrict__ dense) {
  const TableMeta m = metas[t];
  const TableSlot slot = slots[t];
  const size_t n = (size_t)g.w * g.w;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % g.w), y = (int)(i / g.w);
    double v = 0.0;
    if (x >= m.ax0 && x <= m.ax1 && y >= m.ay0 && y <= m.ay1)
      v = slot.values[(size_t)(y - m.ay0) * m.stride + (x - m.ax0)];
    dense[i] = v;
  }
}

// Sum of the low-res tables of a wave (checksum for the table-build benchmark).
__global__ void low_checksum_kernel(Geometry g, const TableSlot* __restrict__ slots,
                                    int n_tables, double* __restrict__ out) {
  const int t = blockIdx.x;
  if (t >= n_tables) return;
  double acc = 0.0;
  for (int i = threadIdx.x; i < g.wl * g.wl; i += blockDim.x) acc += slots[t].low[i];
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) atomicAdd(out, acc);
}

// Fresh slots: low-res tables all zero / all floor (padding included), meta "no table".
__global__ void __launch_bounds__(256) slot_init_kernel(Geometry g, const TableSlot* __restrict__ slots,
                                                        TableMeta* __restrict__ metas) {
  const TableSlot slot = slots[blockIdx.x];
  const double floor_log = CSM_FLOOR_LOG;
  for (int i = threadIdx.x; i < g.wl * g.wl; i += blockDim.x) slot.low[i] = 0.0;
  for (int i = threadIdx.x; i < (int)lowlog_cells(g.wl); i += blockDim.x) slot.lowlog[i] = floor_log;
  if (threadIdx.x == 0) {
    TableMeta m = {};
    m.cx0 = 0; m.cx1 = -1; m.cy0 = 0; m.cy1 = -1;
    m.ax0 = 0; m.ax1 = -1; m.ay0 = 0; m.ay1 = -1;
    m.bx1 = -1; m.by1 = -1; m.cropped = 1;
    m.lx0 = INT_MAX; m.lx1 = INT_MIN; m.ly0 = INT_MAX; m.ly1 = INT_MIN;
    m.min_key = ~0ULL;
    metas[blockIdx.x] = m;
  }
}

void launch_slot_init(const Geometry& g, const TableSlot* slots, TableMeta* metas, int n_slots,
                      cudaStream_t stream) {
  if (n_slots > 0) slot_init_kernel<<<n_slots, 256, 0, stream>>>(g, slots, metas);
}

// ---- launchers --------------------------------------------------------------
// NAUTILUS_CSM_TUNING builds (scripts/ab_build.sh) read A/B knobs from the
// environment; the shipped library has no getenv on its launch paths.
#ifdef NAUTILUS_CSM_TUNING
static int tuning_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
}
#endif

bool table_two_run_supported(int w) { return w <= 32 * RC_MAXCH_LARGE; }

void launch_table_wave(const Geometry& g, const VanVliet& f, const float2* points,
                       const long long* offsets, const int* scan_ids, TableSlot* slots,
                       TableMeta* metas, const PoolListDev& pool, int n_tables, bool store_log,
                       cudaStream_t stream, int* launches, bool has_ck) {
  if (n_tables <= 0) return;
  bool stored = false;
#ifdef NAUTILUS_CSM_TUNING
  stored = tuning_int("NAUTILUS_CSM_TABLE_STORED", 0) != 0;
#endif
  const bool checkpointed = g.w <= 32 * RC_MAXCH_LARGE && !stored;
  Geometry gr = g;  // matching tables store only what can exceed the lookup floor
  gr.tight = (store_log && checkpointed && g.margin >= 0 && g.tail) ? 1 : 0;
#ifdef NAUTILUS_CSM_TUNING
  if (tuning_int("NAUTILUS_CSM_TABLE_TIGHT", 1) == 0) gr.tight = 0;
#endif
  raster_kernel<<<n_tables, 256, 0, stream>>>(gr, points, offsets, scan_ids, slots, metas);
  dim3 gx((g.w + XP_WARPS * 32 - 1) / (XP_WARPS * 32), n_tables);
  dim3 gy((g.w + YP_THREADS - 1) / YP_THREADS, n_tables);
  int np_threads = 128;  // 256 / 192 / 128 / 96 / 64 threads -> 10.24 / 10.30 / 10.04 / 10.24 / 10.59 ms table stage (r1)
#ifdef NAUTILUS_CSM_TUNING
  np_threads = tuning_int("NAUTILUS_CSM_NP_THREADS", np_threads);
#endif
  // matching tables: two-run y-pass instead of y-pass + normalise pass (see ypass2_kernel)
  // It trades a longer dependent chain per column (five sweeps instead of three) for a
  // quarter less DRAM traffic, which only pays when the wave fills the GPU several times
  // over: measured equal at 1 000 tables of the 0.03 m grid, ahead beyond (2 000: 8.8
  // against 9.0 ms; 400 tables of the 0.01 m grid, 9x the cells each: 10.7 against 12.0),
  // behind below (250: 1.50 against 1.33 ms; a single table 0.67 against 0.50 ms per pair).
  const double grid_cells = (double)n_tables * g.w * g.w;
  bool two_run = store_log && checkpointed && pool.two_run_ok && has_ck && grid_cells >= 4.8e9;
#ifdef NAUTILUS_CSM_TUNING
  if (tuning_int("NAUTILUS_CSM_TWO_RUN", -1) == 0) two_run = false;
  if (tuning_int("NAUTILUS_CSM_TWO_RUN", -1) == 1)
    two_run = store_log && checkpointed && pool.two_run_ok && has_ck;
#endif
  if (two_run) {
    if (g.w <= 32 * RC_MAXCH_SMALL) {
      xpass_kernel<RC_MAXCH_SMALL><<<gx, XP_WARPS * 32, 0, stream>>>(g, f, slots, metas);
      ypass2_kernel<16, YP2_MINB1, RC_MAXCH_SMALL, YP2_MAX><<<gy, YP_THREADS, 0, stream>>>(
          g, f, slots, metas, pool.s, pool.dst, pool.win);
      ypass2_kernel<16, YP2_MINB2, RC_MAXCH_SMALL, YP2_LOG><<<gy, YP_THREADS, 0, stream>>>(
          g, f, slots, metas, pool.s, pool.dst, pool.win);
    } else {
      xpass_kernel<RC_MAXCH_LARGE><<<gx, XP_WARPS * 32, 0, stream>>>(g, f, slots, metas);
      ypass2_kernel<16, YP2_MINB1, RC_MAXCH_LARGE, YP2_MAX><<<gy, YP_THREADS, 0, stream>>>(
          g, f, slots, metas, pool.s, pool.dst, pool.win);
      ypass2_kernel<16, YP2_MINB2, RC_MAXCH_LARGE, YP2_LOG><<<gy, YP_THREADS, 0, stream>>>(
          g, f, slots, metas, pool.s, pool.dst, pool.win);
    }
    low_finalize_kernel<<<n_tables, 256, 0, stream>>>(g, slots, metas, pool.s, pool.e, pool.dst,
                                                      pool.win);
    if (launches) *launches += 5;
    return;
  }
  if (g.w <= 32 * RC_MAXCH_SMALL && !stored) {
    xpass_kernel<RC_MAXCH_SMALL><<<gx, XP_WARPS * 32, 0, stream>>>(g, f, slots, metas);
    ypass_kernel<16, YP1_MINB, RC_MAXCH_SMALL><<<gy, YP_THREADS, 0, stream>>>(g, f, slots, metas);
  } else if (g.w <= 32 * RC_MAXCH_LARGE && !stored) {
    xpass_kernel<RC_MAXCH_LARGE><<<gx, XP_WARPS * 32, 0, stream>>>(g, f, slots, metas);
    ypass_kernel<16, YP1_MINB, RC_MAXCH_LARGE><<<gy, YP_THREADS, 0, stream>>>(g, f, slots, metas);
  } else {
    xpass_stored_kernel<<<gx, XP_WARPS * 32, 0, stream>>>(g, f, slots, metas);
    ypass_stored_kernel<<<gy, YP_THREADS, 0, stream>>>(g, f, slots, metas);
  }
  dim3 gn(g.n_pool, n_tables);
  const size_t np_smem = sizeof(double) * g.w;
  static SmemOptIn np_opt;
  np_opt.allow(normalize_pool_kernel, np_smem);
  normalize_pool_kernel<<<gn, np_threads, np_smem, stream>>>(
      g, slots, metas, pool.s, pool.e, pool.dst, pool.win, store_log ? 1 : 0);
  if (launches) *launches += 4;
}

void launch_pool_generic(const double* hi, int w, int h, const PoolListDev& px,
                         const PoolListDev& py, double* low, int wl, cudaStream_t stream) {
  dim3 grid((px.n + 127) / 128, py.n);
  pool_generic_kernel<<<grid, 128, 0, stream>>>(hi, w, h, px.s, px.e, px.dst, px.win, px.n,
                                                py.s, py.e, py.dst, py.win, py.n, low, wl);
}

void launch_expand(const Geometry& g, const TableSlot* slots, const TableMeta* metas, int t,
                   double* dense, cudaStream_t stream) {
  expand_kernel<<<592, 256, 0, stream>>>(g, slots, metas, t, dense);
}

void launch_low_checksum(const Geometry& g, const TableSlot* slots, int n_tables, double* out,
                         cudaStream_t stream) {
  low_checksum_kernel<<<n_tables, 256, 0, stream>>>(g, slots, n_tables, out);
}

}  // namespace csm
