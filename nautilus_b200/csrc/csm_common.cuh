// Shared device-side definitions for the CSM kernels (sm_100a).
//
// Everything numeric in here follows the reference arithmetic restated in
// SURVEY.md Appendix A; the whole library is compiled with --fmad=false so no
// multiply-add is contracted, matching the reference build (no -march flag =>
// no FMA, /root/reference/CMakeLists.txt:11,16).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#define CSM_MIN_VALUE_FOR_LOOKUP 1E-10  // CorrelativeScanMatcher.h:20
#define CSM_EPSILON 1e-6                // CorrelativeScanMatcher.cc:19

namespace csm {

// log(MIN_VALUE_FOR_LOOKUP) correctly rounded (what glibc's log returns, cc:72);
// one constant everywhere so that every floor lookup adds the same bits.
#define CSM_FLOOR_LOG (-23.025850929940457)

// Natural logarithm for the table kernels, which take it once per cell above the lookup
// floor (a third of a table): ~30 instructions instead of the ~65 of CUDA's log().
// Table-driven like the log of current C libraries (128 intervals of the mantissa range
// [0.6875, 1.375), z = x 2^-k, r = z / c_i - 1 by one FMA, |r| < 2^-7):
//   log(x) = k ln2 + log(c_i) + (r - r^2/2 + ... + r^7/7)
// with k ln2 + log(c_i) carried as a double-double, so the result is within 1 ulp like
// log()'s (tests/test_gpu_log.py compares the two over the table's range).  Arguments near
// 1, where the leading terms cancel, and anything not a positive normal number go to log().
#include "log_table.inc"
struct CsmLogEntry { double invc, logc_hi, logc_lo; };
static __device__ const CsmLogEntry csm_log_table[128] = {CSM_LOG_TABLE};

__device__ __forceinline__ double csm_log(double x) {
  if (!(x >= 1e-300 && x < 0.9375)) return log(x);
  const long long ix = __double_as_longlong(x);
  const long long tmp = ix - 0x3fe6000000000000LL;
  const int i = (int)(tmp >> 45) & 127;
  const int k = (int)(tmp >> 52);  // arithmetic: floor
  const double z = __longlong_as_double(ix - (tmp & (long long)0xfff0000000000000ULL));
  const CsmLogEntry* e = csm_log_table + i;
  const double invc = __ldg(&e->invc), lc_hi = __ldg(&e->logc_hi), lc_lo = __ldg(&e->logc_lo);
  const double r = __fma_rn(z, invc, -1.0);
  const double kd = (double)k;
  const double t = __dmul_rn(kd, CSM_LN2_HI);           // exact: LN2_HI has 42 bits
  const double w = __dadd_rn(t, lc_hi);
  const double w_err = __dadd_rn(__dsub_rn(t, w), lc_hi);  // |t| >= |lc_hi| or t == 0
  const double hi = __dadd_rn(w, r);
  double lo = __dadd_rn(__dsub_rn(w, hi), r);             // |w| >= |r| below 0.9375
  lo = __dadd_rn(lo, __dadd_rn(w_err, __fma_rn(kd, CSM_LN2_LO, lc_lo)));
  const double r2 = __dmul_rn(r, r);
  double p = __fma_rn(r, 1.0 / 7.0, -1.0 / 6.0);
  p = __fma_rn(r, p, 1.0 / 5.0);
  p = __fma_rn(r, p, -1.0 / 4.0);
  p = __fma_rn(r, p, 1.0 / 3.0);
  p = __fma_rn(r, p, -0.5);
  return __dadd_rn(hi, __fma_rn(r2, p, lo));
}

// d / b for many d and one b (the normalisation divides a table by its range): with
// binv = RN(1 / b), two residual corrections by FMA give the correctly rounded quotient
// (Markstein: q1 is faithful, hence q2 = RN(q1 + rem binv) = RN(d / b)) in 5 operations
// instead of the ~20 of a full division.  For normal d, b, d / b.
__device__ __forceinline__ double div_by(double d, double b, double binv) {
  const double q0 = __dmul_rn(d, binv);
  const double q1 = __fma_rn(__fma_rn(-q0, b, d), binv, q0);
  return __fma_rn(__fma_rn(-q1, b, d), binv, q1);
}

// log(GetPointValue) of a table cell value (h:51-59 + cc:72).
__device__ __forceinline__ double lookup_log(double v) {
  return v > CSM_MIN_VALUE_FOR_LOOKUP ? csm_log(v) : CSM_FLOOR_LOG;
}

// Floor padding of the pooled log table: a full tile of the low-res scorer (14 x
// translations by 16 y translations, csm_search.cu LS_TI x LS_TJ) on every side, so that
// the 14 x 16 patch of cells any (point, rotation) addresses lies inside the padded
// array wherever it falls relative to the table - entirely outside patches are clamped
// onto the padding - and needs no bounds test.
#define CSM_LOW_PADX 14
#define CSM_LOW_PADY 16
__host__ __device__ __forceinline__ int lowlog_stride(int wl) { return (wl + 2 * CSM_LOW_PADX) | 1; }
__host__ __device__ __forceinline__ int lowlog_rows(int wl) { return wl + 2 * CSM_LOW_PADY; }
__host__ __device__ __forceinline__ size_t lowlog_cells(int wl) {
  return (size_t)lowlog_rows(wl) * (size_t)lowlog_stride(wl);
}
__host__ __device__ __forceinline__ size_t lowlog_index(int wl, int x, int y) {
  return (size_t)(y + CSM_LOW_PADY) * (size_t)lowlog_stride(wl) + (size_t)(x + CSM_LOW_PADX);
}

// Van Vliet-Young order-0 coefficients + Triggs-Sdika matrix
// (CImg.h:49267-49284, 49066-49081); filled on the host.
struct VanVliet {
  double B, a1, a2, a3;
  double M[9];
};

// Per table-slot metadata, written by the raster kernel.
struct TableMeta {
  // occupied box of the scan, in hi-res cells, inclusive; cx1 < cx0 => no point
  int cx0, cx1, cy0, cy1;
  // active rectangle the blurred table is STORED on (inclusive): what the normalise
  // pass, the search kernels and the table download see; cells outside it read as 0 /
  // the lookup floor
  int ax0, ax1, ay0, ay1;
  // last column / row the recursions are COMPUTED on (>= ax1 / ay1): the anti-causal
  // passes start here.  Tables built for matching store less than they compute, see
  // raster_kernel.
  int bx1, by1;
  int stride;    // row stride of the slot's value buffer, in doubles
  int cropped;   // 1 when the active rectangle is not the whole grid
  int n_inside;  // points that fell inside the grid
  int pad;
  unsigned long long max_key;  // order-preserving keys of the raw max / min
  unsigned long long min_key;
  // box of low-res cells whose value exceeds the lookup floor (1e-10); every
  // other low-res cell reads as the floor.  lx1 < lx0 => none.
  int lx0, lx1, ly0, ly1;
};

// One resident table slot.
struct TableSlot {
  double* values;      // active rectangle, row-major, stride meta.stride; holds the
                       // normalised values, or their lookup_log when the wave was
                       // built for matching (launch_table_wave store_log)
  unsigned* bitmap;    // occupied-box bitmap, bm_stride words per row
  unsigned char* rowflag;  // 1 per occupied-box row that has a point
  double* low;         // wl*wl low-res values (max-pool of the normalised table)
  double* lowlog;      // lookup_log of the low-res table in the padded layout the
                       // scorer reads: lowlog_rows(wl) rows of lowlog_stride(wl)
                       // doubles, cell (x, y) at lowlog_index(wl, x, y); the padding
                       // (CSM_LOW_PADX columns / CSM_LOW_PADY rows each side) holds the floor
  double* ck;          // scratch of the two-run y-pass: causal states per 16-row chunk,
                       // [chunk][3][column], ck_doubles(w) doubles; null when the grid takes
                       // the stored kernels
};

// chunks of 16 rows + the final state, three doubles per column
__host__ __device__ inline size_t ck_doubles(size_t w) {
  return ((w + 15) / 16 + 1) * 3 * ((w + 3) & ~(size_t)3);
}

// Grid geometry shared by all kernels.
struct Geometry {
  int w;          // hi-res table width = height  (LookupTable ctor, h:32-35)
  int wl;         // low-res width
  double res_h;   // high_res_
  double res_l;   // low_res_
  int bm_stride;  // bitmap words per row = ceil(w/32)
  int margin;     // crop margin in cells; <0 => never crop
  int n_pool;     // number of pooling windows per axis (loop trips of cc:37)
  // tail[d] = sup over d' >= d of g(d') / g(0), g = 1-D impulse response of the blur
  // (device array of margin + 1 doubles, null when margin < 0): lets the raster kernel
  // bound how far from a scan its table can exceed the lookup floor
  const double* tail;
  int tight;      // 1: tables of this wave store only what can exceed the floor
};

// Exact cell lookup without a double division.  cell(v) = dim/2 + floor((double)v / res)
// is monotone in v, so it is determined by the float thresholds
//   lowf[q - qlo] = smallest float v with floor((double)v / res) >= q,
// which the host finds with the reference's own expression.  A float estimate of
// the quotient is corrected by at most one step against them.
struct CellMap {
  const float* lowf;  // thresholds for q = qlo .. qhi + 1
  int qlo, qhi;       // quotients of the first / last in-grid cell
  int half;           // dim / 2
  float inv_res;
};

// Returns the cell in [0, dim) or -1 when v falls outside the grid.  The float
// estimate e = v * (float)(1/res) is within |q| * 1.2e-7 <= 2e-3 of the true
// quotient (|q| <= 15000, enforced by create()), so it is only ambiguous when
// its fractional part is within CELL_EPS of an integer; only then (and at the
// grid edge) are the thresholds consulted, and one correction step suffices.
// The host twin of this function is checked against the division form in
// tests/test_abi.py.
#define CSM_CELL_EPS 4e-3f
__host__ __device__ __forceinline__ int cell_fast(const CellMap& m, float v) {
#ifdef __CUDA_ARCH__
  const float e = __fmul_rn(v, m.inv_res);
  int q = __float2int_rd(e);  // saturating
#else
  const float e = v * m.inv_res;
  const float fl0 = floorf(e);
  int q = !(fl0 > -2.0e9f) ? -2147483647 - 1 : (fl0 > 2.0e9f ? 2147483647 : (int)fl0);
#endif
  if (!(v == v)) return -1;      // NaN: outside, like the reference's double -> uint64 conversion
  const float fr = e - (float)q;  // exact while |q| < 2^24; irrelevant once saturated
  // strictly inside the grid and clear of both cell boundaries: the estimate is the cell
  const bool inside = (unsigned)(q - m.qlo - 1) < (unsigned)(m.qhi - m.qlo - 1);
  if (inside && fabsf(fr - 0.5f) < 0.5f - CSM_CELL_EPS) return q + m.half;
  q = q < m.qlo ? m.qlo : (q > m.qhi ? m.qhi : q);
  const float* t = m.lowf + (q - m.qlo);
  if (v < t[0]) --q;
  else if (v >= t[1]) ++q;
  return (q < m.qlo || q > m.qhi) ? -1 : q + m.half;
}

// convertX/convertY (CorrelativeScanMatcher.h:43-49): dim/2 + floor(x/res) in
// double.  IEEE division (div.rn.f64) => identical to the host result.
__device__ __forceinline__ long long cell_of(float v, double res, int dim) {
  const double q = floor(__ddiv_rn((double)v, res));
  const double c = (double)(dim / 2) + q;
  if (!(c >= -4.0e18)) return -4000000000000000000LL;  // also NaN: outside the grid
  if (c > 4.0e18) return 4000000000000000000LL;
  return (long long)c;
}

// Order-preserving map double -> uint64 (for atomicMax / atomicMin).
__device__ __forceinline__ unsigned long long order_key(double v) {
  unsigned long long b = (unsigned long long)__double_as_longlong(v);
  return (b & 0x8000000000000000ULL) ? ~b : (b | 0x8000000000000000ULL);
}
__device__ __host__ __forceinline__ double key_to_double(unsigned long long k) {
  unsigned long long b = (k & 0x8000000000000000ULL) ? (k & 0x7FFFFFFFFFFFFFFFULL) : ~k;
#ifdef __CUDA_ARCH__
  return __longlong_as_double((long long)b);
#else
  double d;
  memcpy(&d, &b, sizeof(d));
  return d;
#endif
}

// One step of the order-0 recursion, accumulation order of CImg.h:49103-49105:
// ((x + v1*a1) + v2*a2) + v3*a3, every product and sum rounded separately.
__device__ __forceinline__ double vv_step(double x, double v1, double v2, double v3,
                                          double a1, double a2, double a3) {
  double v0 = __dadd_rn(x, __dmul_rn(v1, a1));
  v0 = __dadd_rn(v0, __dmul_rn(v2, a2));
  v0 = __dadd_rn(v0, __dmul_rn(v3, a3));
  return v0;
}

// Triggs boundary of CImg.h:49092-49099.  in: v1..v3 = last three causal
// outputs, iplus = last input sample.  out: w0..w2.
__device__ __forceinline__ void vv_triggs(const VanVliet& f, double iplus, double v1,
                                          double v2, double v3, double& w0, double& w1,
                                          double& w2) {
  const double sum = __dmul_rn(f.B, f.B);
  const double den = __dsub_rn(__dsub_rn(__dsub_rn(1.0, f.a1), f.a2), f.a3);
  const double uplus = __ddiv_rn(iplus, den);
  const double vplus = __ddiv_rn(uplus, den);
  const double unp = __dsub_rn(v1, uplus), unp1 = __dsub_rn(v2, uplus),
               unp2 = __dsub_rn(v3, uplus);
  const double* M = f.M;
  w0 = __dmul_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(M[0], unp), __dmul_rn(M[1], unp1)),
                                     __dmul_rn(M[2], unp2)), vplus), sum);
  w1 = __dmul_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(M[3], unp), __dmul_rn(M[4], unp1)),
                                     __dmul_rn(M[5], unp2)), vplus), sum);
  w2 = __dmul_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(M[6], unp), __dmul_rn(M[7], unp1)),
                                     __dmul_rn(M[8], unp2)), vplus), sum);
}

// Opt-in to more than 48 KB of dynamic shared memory.  The attribute belongs to
// the (kernel, device) pair, so the largest size already granted is remembered per
// device ordinal of the calling thread's current device; handles on different
// devices and on different host threads may launch concurrently.
struct SmemOptIn {
  static constexpr int kMaxDev = 64;
  std::mutex mu;
  size_t granted[kMaxDev] = {};
  int per_sm[kMaxDev] = {};      // resident blocks per SM at per_sm_smem (persistent launches)
  size_t per_sm_smem[kMaxDev] = {};
  template <typename K>
  cudaError_t allow(K kernel, size_t smem) {
    if (smem <= 48 * 1024) return cudaSuccess;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> g(mu);
    if (dev < 0 || dev >= kMaxDev || smem > granted[dev]) {
      e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e == cudaSuccess && dev >= 0 && dev < kMaxDev) granted[dev] = smem;
    }
    return e;
  }
  template <typename K>
  int blocks_per_sm(K kernel, int threads, size_t smem) {
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> g(mu);
    if (dev >= 0 && dev < kMaxDev && per_sm[dev] > 0 && per_sm_smem[dev] == smem) return per_sm[dev];
    int n = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, threads, smem);
    if (n < 1) n = 1;
    if (dev >= 0 && dev < kMaxDev) { per_sm[dev] = n; per_sm_smem[dev] = smem; }
    return n;
  }
};

}  // namespace csm
