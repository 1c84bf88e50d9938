// C ABI of libnautilus_csm_b200.so (include/nautilus_csm.h): handle, device
// buffer ownership, host-built index lists, wave scheduling.  No CPU fallback:
// every compute entry point launches the sm_100a kernels or fails.
//
// Host-built lists.  Every coordinate the reference derives from an
// accumulating double, a float narrowing or a floor/round is one-dimensional
// and tiny, so it is produced here with the reference's own expressions
//  and uploaded: pooling windows, low-res translations
// and their memo slots, hi-res window steps, rotation lists with libm
// sinf/cosf.  The device only evaluates floor(x / res) on point + translation.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <numeric>
#include <string>
#include <utility>
#include <vector>

#include "csm_handle.h"

using namespace csm;
using namespace csmh;

namespace csmh {
std::string& error_text() {
  thread_local std::string g_error;
  return g_error;
}
}  // namespace csmh

namespace {

// convertX / convertY on the host (CorrelativeScanMatcher.h:43-49)
long long host_cell(float v, double res, uint64_t dim) {
  const double c = (double)(dim / 2) + floor((double)v / res);
  if (!(c >= -4.0e18)) return -4000000000000000000LL;  // also NaN
  if (c > 4.0e18) return 4000000000000000000LL;
  return (long long)c;
}

// tail[d] = sup_{d' >= d} g(d') / g(0) for the 1-D impulse response g of the blur
// (order-0 Van Vliet recursion, forward then backward, CImg.h:49084-49114, on a line
// long enough that both ends hold zeros).  The raster kernel uses it to bound how far
// from a scan a table can exceed the lookup floor (csm_tables.cu).
std::vector<double> blur_tail(const VanVliet& f, int margin) {
  const int n = 4 * margin + 65, c = n / 2;
  std::vector<double> x((size_t)n, 0.0);
  x[c] = 1.0;
  double v1 = 0, v2 = 0, v3 = 0;
  for (int i = 0; i < n; ++i) {
    const double v0 = ((x[i] + v1 * f.a1) + v2 * f.a2) + v3 * f.a3;
    x[i] = v0; v3 = v2; v2 = v1; v1 = v0;
  }
  const double sum = f.B * f.B;
  v1 = v2 = v3 = 0;  // zero state at the far end (the Triggs term of a zero tail)
  for (int i = n - 1; i >= 0; --i) {
    const double v0 = ((x[i] * sum + v1 * f.a1) + v2 * f.a2) + v3 * f.a3;
    x[i] = v0; v3 = v2; v2 = v1; v1 = v0;
  }
  std::vector<double> tail((size_t)margin + 1, 0.0);
  double run = 0.0;
  for (int d = n - 1 - c; d >= 0; --d) {
    run = std::max(run, std::max(fabs(x[c + d]), fabs(x[c - d])));
    if (d <= margin) tail[d] = x[c] > 0 ? run / x[c] : 1.0;
  }
  return tail;
}

// Float thresholds of the cell map (see csm::CellMap): entry k is the smallest
// float v with floor((double)v / res) >= qlo + k, found with the reference's
// own expression (CorrelativeScanMatcher.h:43-49) by stepping nextafterf.
std::vector<float> make_thresholds(double res, int dim, int* qlo_out, int* qhi_out) {
  const int half = dim / 2, qlo = -half, qhi = dim - 1 - half;
  std::vector<float> thr;
  for (int q = qlo; q <= qhi + 1; ++q) {
    auto ok = [&](float x) { return floor((double)x / res) >= (double)q; };
    float v = (float)((double)q * res);
    if (ok(v)) {
      for (;;) {
        const float d = nextafterf(v, -INFINITY);
        if (!ok(d)) break;
        v = d;
      }
    } else {
      while (!ok(v)) v = nextafterf(v, INFINITY);
    }
    thr.push_back(v);
  }
  *qlo_out = qlo;
  *qhi_out = qhi;
  return thr;
}

PoolListHost make_pool_list(double range, double low_res, double hi_res, uint64_t hi_dim,
                            uint64_t low_dim) {
  PoolListHost pl;
  for (double x = -range + CSM_EPSILON; x < range; x += low_res) {
    // MaxArea(x, ., x + low_res, .): doubles narrowed to float by convertX
    const uint64_t usx = (uint64_t)host_cell((float)x, hi_res, hi_dim);
    const uint64_t uex = (uint64_t)host_cell((float)(x + low_res), hi_res, hi_dim);
    const int x0 = (int)usx, x1 = (int)(uex - 1);  // get_crop(int, ...) narrowing
    pl.s.push_back(std::min(x0, x1));
    pl.e.push_back(std::max(x0, x1));
    const long long d = host_cell((float)x, low_res, low_dim);  // SetPointValue(Vector2f(x, y))
    pl.dst.push_back((d >= 0 && d < (long long)low_dim) ? (int)d : -1);
    if (pl.s.size() > 10000000) break;  // degenerate parameters
  }
  const int n = (int)pl.s.size();
  pl.win.assign(n, 0);
  std::map<int, int> last;
  for (int k = 0; k < n; ++k)
    if (pl.dst[k] >= 0) last[pl.dst[k]] = k;
  for (auto& kv : last) pl.win[kv.second] = 1;
  bool tiles = n > 0 && pl.s[0] <= 0 && pl.e[n - 1] >= (int)hi_dim - 1;
  for (int k = 0; tiles && k < n; ++k) {
    if (pl.s[k] > pl.e[k]) tiles = false;
    if (k + 1 < n && pl.e[k] + 1 != pl.s[k + 1]) tiles = false;
  }
  pl.tiles = tiles;
  return pl;
}

}  // namespace


namespace {

void free_slots(NautilusCsm* h) {
  cudaFree(h->slab_values); cudaFree(h->slab_bitmap); cudaFree(h->slab_rowflag);
  cudaFree(h->slab_low); cudaFree(h->slab_lowlog); cudaFree(h->slab_ck);
  h->slab_values = nullptr; h->slab_bitmap = nullptr; h->slab_rowflag = nullptr;
  h->slab_low = nullptr; h->slab_lowlog = nullptr; h->slab_ck = nullptr;
  h->n_slots = 0;
}

// Whether the handle's matching tables can take the two-run y-pass (csm_tables.cu): only
// then do the slots carry its checkpoint scratch.
bool wants_ck(const NautilusCsm* h) {
  return h->geo.margin >= 0 && h->geo.tail && table_two_run_supported(h->geo.w) &&
         h->pool_dev.d.two_run_ok;
}

size_t slot_bytes(const NautilusCsm* h) {
  const Geometry& g = h->geo;
  const size_t w = g.w, wl = g.wl;
  return sizeof(double) * ((w + 3) & ~(size_t)3) * w + sizeof(unsigned) * g.bm_stride * w + w +
         sizeof(double) * (wl * wl + lowlog_cells((int)wl)) +
         (wants_ck(h) ? sizeof(double) * ck_doubles(w) : 0);
}

// Makes at least `want` table slots resident (bounded by params.max_tables and
// by free device memory); returns the slot count through h->n_slots.
int ensure_slots(NautilusCsm* h, int want) {
  const Geometry& g = h->geo;
  int cap = h->params.max_tables > 0 ? h->params.max_tables : 4096;
  if (std::max(1, std::min(want, cap)) <= h->n_slots) return NAUTILUS_OK;  // steady state: no driver query
  size_t free_b = 0, total_b = 0;
  CU_TRY(cudaMemGetInfo(&free_b, &total_b));
  const size_t per = slot_bytes(h);
  const size_t have = (size_t)h->n_slots * per;
  const size_t budget = (free_b + have) / 2;  // leave half of what is free to the caller
  cap = (int)std::max<size_t>(1, std::min<size_t>((size_t)cap, budget / per));
  want = std::max(1, std::min(want, cap));
  if (want <= h->n_slots) return NAUTILUS_OK;
  free_slots(h);
  const size_t w = g.w, wl = g.wl, vstride = ((w + 3) & ~(size_t)3) * w;
  CU_TRY(cudaMalloc(&h->slab_values, sizeof(double) * vstride * want));
  CU_TRY(cudaMalloc(&h->slab_bitmap, sizeof(unsigned) * (size_t)g.bm_stride * w * want));
  CU_TRY(cudaMalloc(&h->slab_rowflag, w * (size_t)want));
  CU_TRY(cudaMalloc(&h->slab_low, sizeof(double) * wl * wl * want));
  CU_TRY(cudaMalloc(&h->slab_lowlog, sizeof(double) * lowlog_cells((int)wl) * want));
  const bool with_ck = wants_ck(h);
  if (with_ck) CU_TRY(cudaMalloc(&h->slab_ck, sizeof(double) * ck_doubles(w) * want));
  h->slots_host.resize(want);
  for (int i = 0; i < want; ++i) {
    TableSlot& s = h->slots_host[i];
    s.values = h->slab_values + vstride * i;
    s.bitmap = h->slab_bitmap + (size_t)g.bm_stride * w * i;
    s.rowflag = h->slab_rowflag + w * (size_t)i;
    s.low = h->slab_low + wl * wl * (size_t)i;
    s.lowlog = h->slab_lowlog + lowlog_cells((int)wl) * (size_t)i;
    s.ck = with_ck ? h->slab_ck + ck_doubles(w) * (size_t)i : nullptr;
  }
  int rc = h->slots_dev.reserve(want);
  if (rc) return rc;
  rc = h->metas_dev.reserve(want);
  if (rc) return rc;
  CU_TRY(cudaMemcpy(h->slots_dev.p, h->slots_host.data(), sizeof(TableSlot) * want,
                    cudaMemcpyHostToDevice));
  launch_slot_init(g, h->slots_dev.p, h->metas_dev.p, want, nullptr);
  CU_TRY(cudaDeviceSynchronize());  // (allocation path only) before any stream of the handle uses the slots
  h->n_slots = want;
  return NAUTILUS_OK;
}

// Low-res translation grid, memo slots and hi-res window steps.
int build_search_grid(NautilusCsm* h) {
  const double T = h->params.trans_range, rl = h->params.low_res, rh = h->params.high_res;
  const uint64_t wl = (uint64_t)h->geo.wl;
  std::vector<double> t_low_d;
  for (double t = -T + CSM_EPSILON; t < T; t += rl) {  // cc:96-99
    t_low_d.push_back(t);
    if (t_low_d.size() > 100000) return fail(NAUTILUS_EINVAL, "translation grid too large");
  }
  const int n_t = (int)t_low_d.size();
  if (n_t == 0) return fail(NAUTILUS_EINVAL, "empty translation grid (trans_range <= 0?)");
  h->t_low_host.resize(n_t);
  for (int i = 0; i < n_t; ++i) h->t_low_host[i] = (float)t_low_d[i];
  // AbsCoords (h:104-109) of every (x_i, y_j); max_element visits slots in this order
  std::vector<std::pair<unsigned long long, int>> order;
  order.reserve((size_t)n_t * n_t);
  for (int i = 0; i < n_t; ++i)
    for (int j = 0; j < n_t; ++j) {
      const size_t row = (size_t)(((double)(wl / 2) + round(t_low_d[j] / rl)) * (double)wl);
      const size_t col = (size_t)((double)(wl / 2) + round(t_low_d[i] / rl));
      order.emplace_back((unsigned long long)(row + col), i * n_t + j);
    }
  std::sort(order.begin(), order.end());
  for (size_t k = 1; k < order.size(); ++k)
    if (order[k].first == order[k - 1].first)
      return fail(NAUTILUS_EINVAL,
                  "two low-res translations share a memo slot (AbsCoords collision); "
                  "unsupported parameter combination");
  h->slot_rank_host.assign((size_t)n_t * n_t, 0);
  for (size_t k = 0; k < order.size(); ++k) h->slot_rank_host[order[k].second] = (int)k;
  // hi-res window opened at low index i (cc:195-202, 126-129)
  std::vector<std::vector<float>> steps(n_t);
  int nh_max = 1;
  for (int i = 0; i < n_t; ++i) {
    const float fx = h->t_low_host[i];
    const double lo = std::max((double)fx, -T), hi = std::min(fx + rl, T);
    for (double x = lo + CSM_EPSILON; x < hi; x += rh) {
      steps[i].push_back((float)x);
      if (steps[i].size() > 4096) return fail(NAUTILUS_EINVAL, "hi-res window too large");
    }
    nh_max = std::max(nh_max, (int)steps[i].size());
  }
  h->n_h_host.resize(n_t);
  h->t_high_host.assign((size_t)n_t * nh_max, 0.f);
  for (int i = 0; i < n_t; ++i) {
    h->n_h_host[i] = (int)steps[i].size();
    std::copy(steps[i].begin(), steps[i].end(), h->t_high_host.begin() + (size_t)i * nh_max);
  }
  int rc;
  if ((rc = h->t_low_dev.reserve(n_t))) return rc;
  if ((rc = h->t_high_dev.reserve((size_t)n_t * nh_max))) return rc;
  if ((rc = h->slot_rank_dev.reserve((size_t)n_t * n_t))) return rc;
  if ((rc = h->n_h_dev.reserve(n_t))) return rc;
  CU_TRY(cudaMemcpy(h->t_low_dev.p, h->t_low_host.data(), sizeof(float) * n_t, cudaMemcpyHostToDevice));
  CU_TRY(cudaMemcpy(h->t_high_dev.p, h->t_high_host.data(), sizeof(float) * n_t * nh_max,
                    cudaMemcpyHostToDevice));
  CU_TRY(cudaMemcpy(h->slot_rank_dev.p, h->slot_rank_host.data(), sizeof(int) * n_t * n_t,
                    cudaMemcpyHostToDevice));
  CU_TRY(cudaMemcpy(h->n_h_dev.p, h->n_h_host.data(), sizeof(int) * n_t, cudaMemcpyHostToDevice));
  h->grid.n_t = n_t;
  h->grid.t_low = h->t_low_dev.p;
  h->grid.slot_rank = h->slot_rank_dev.p;
  h->grid.n_h_max = nh_max;
  h->grid.n_h = h->n_h_dev.p;
  h->grid.t_high = h->t_high_dev.p;
  {
    // Regular-cell shortcut of the low-res scorer.  With x_k = (double)fl(p + t_k) / rl
    // the exact quotient of translation k, x_k - x_k0 - (k - k0) is bounded by the
    // two float additions (one ulp at the largest |p + t|) plus how far the stored
    // float translations are from an exact one-cell step; the device estimate of
    // x_k0 (float product with the rounded 1/rl) is within 2^-22 * |x|.  reg_eps
    // is four times the sum, so floor(x_k) = floor(estimate) + (k - k0).
    const double vmax = h->params.scanner_range + T + rl;
    const double qmax = vmax / rl + 2.0;
    double dev = 0.0;
    for (int k0 = 0; k0 < n_t; ++k0)
      for (int k = k0; k < std::min(n_t, k0 + 16); ++k)
        dev = std::max(dev, fabs((double)h->t_low_host[k] - (double)h->t_low_host[k0] - (k - k0) * rl));
    const double ulp = (double)nextafterf((float)vmax, INFINITY) - (double)(float)vmax;
    const double eps = 4.0 * (qmax * ldexp(1.0, -22) + (ulp + dev) / rl) + 1e-6;
    h->grid.reg_eps = (float)std::min(eps, 0.49);  // >= 0.5 never passes: always cell_fast
    if (eps >= 0.49) h->grid.reg_eps = 0.5f;
    h->grid.reg_qmax = (float)qmax;
    // hi-res windows: step k of a window against its first step
    const double qmax_h = vmax / rh + 2.0;
    double dev_h = 0.0;
    for (int i = 0; i < n_t; ++i)
      for (size_t k = 0; k < steps[i].size(); ++k)
        dev_h = std::max(dev_h, fabs((double)steps[i][k] - (double)steps[i][0] - (double)k * rh));
    const double eps_h = 4.0 * (qmax_h * ldexp(1.0, -22) + (ulp + dev_h) / rh) + 1e-6;
    h->grid.reg_eps_h = eps_h >= 0.49 ? 0.5f : (float)eps_h;
    h->grid.reg_qmax_h = (float)qmax_h;
  }
  // exact cell maps for the two grids
  int rc2;
  h->thr_low_host = make_thresholds(rl, h->geo.wl, &h->cm_low.qlo, &h->cm_low.qhi);
  h->thr_high_host = make_thresholds(rh, h->geo.w, &h->cm_high.qlo, &h->cm_high.qhi);
  if ((rc2 = h->thr_low_dev.reserve(h->thr_low_host.size()))) return rc2;
  if ((rc2 = h->thr_high_dev.reserve(h->thr_high_host.size()))) return rc2;
  CU_TRY(cudaMemcpy(h->thr_low_dev.p, h->thr_low_host.data(), sizeof(float) * h->thr_low_host.size(),
                    cudaMemcpyHostToDevice));
  CU_TRY(cudaMemcpy(h->thr_high_dev.p, h->thr_high_host.data(),
                    sizeof(float) * h->thr_high_host.size(), cudaMemcpyHostToDevice));
  h->cm_low.lowf = h->thr_low_dev.p;
  h->cm_low.half = h->geo.wl / 2;
  h->cm_low.inv_res = (float)(1.0 / rl);
  h->cm_high.lowf = h->thr_high_dev.p;
  h->cm_high.half = h->geo.w / 2;
  h->cm_high.inv_res = (float)(1.0 / rh);
  return NAUTILUS_OK;
}

// `pinned_source`: xy lives in the handle's own page-locked staging, so the copy is a
// true asynchronous DMA and the stream need not be drained before returning (the
// staging is not rewritten before the batch that uses it has been collected).
int upload_scan_set(NautilusCsm* h, ScanSet& ss, const float* xy, const int64_t* offsets,
                    int32_t n_scans, bool pinned_source = false) {
  if (n_scans < 0 || (n_scans > 0 && (!offsets))) return fail(NAUTILUS_EINVAL, "bad scan set");
  const long long total = n_scans > 0 ? offsets[n_scans] : 0;
  if (n_scans > 0 && offsets[0] != 0) return fail(NAUTILUS_EINVAL, "offsets[0] must be 0");
  for (int i = 0; i < n_scans; ++i)
    if (offsets[i + 1] < offsets[i]) return fail(NAUTILUS_EINVAL, "offsets must be non-decreasing");
  if (total > 0 && !xy) return fail(NAUTILUS_EINVAL, "xy is null");
  if ((size_t)total > ss.cap_points) {
    cudaFree(ss.points); ss.points = nullptr; ss.cap_points = 0;
    CU_TRY(cudaMalloc(&ss.points, sizeof(float2) * std::max<size_t>(total, 1024)));
    ss.cap_points = std::max<size_t>(total, 1024);
  }
  if ((size_t)n_scans + 1 > ss.cap_offsets) {
    cudaFree(ss.offsets); ss.offsets = nullptr; ss.cap_offsets = 0;
    CU_TRY(cudaMalloc(&ss.offsets, sizeof(long long) * std::max<size_t>(n_scans + 1, 64)));
    ss.cap_offsets = std::max<size_t>(n_scans + 1, 64);
  }
  ss.host_offsets.assign(n_scans + 1, 0);
  for (int i = 0; i <= n_scans && n_scans > 0; ++i) ss.host_offsets[i] = offsets[i];
  if (total > 0)
    CU_TRY(cudaMemcpyAsync(ss.points, xy, sizeof(float2) * total, cudaMemcpyHostToDevice, h->stream));
  if (pinned_source) {
    int rc = ss.offsets_pin.reserve((size_t)n_scans + 1);
    if (rc) return rc;
    for (int i = 0; i <= n_scans; ++i) ss.offsets_pin.p[i] = ss.host_offsets[i];
    CU_TRY(cudaMemcpyAsync(ss.offsets, ss.offsets_pin.p, sizeof(long long) * (n_scans + 1),
                           cudaMemcpyHostToDevice, h->stream));
  } else {
    CU_TRY(cudaMemcpyAsync(ss.offsets, ss.host_offsets.data(), sizeof(long long) * (n_scans + 1),
                           cudaMemcpyHostToDevice, h->stream));
    CU_TRY(cudaStreamSynchronize(h->stream));
  }
  ss.n = n_scans;
  return NAUTILUS_OK;
}

cudaEvent_t get_event(NautilusCsm* h, size_t k) {
  while (h->events.size() <= k) {
    cudaEvent_t e;
    cudaEventCreate(&e);
    h->events.push_back(e);
  }
  return h->events[k];
}

}  // namespace

extern "C" {

const char* nautilus_csm_last_error(void) { return error_text().c_str(); }
const char* nautilus_csm_version(void) { return "nautilus_csm_b200 sm_100a r1"; }

int64_t nautilus_csm_cell(double resolution, uint64_t dim, float v) {
  return host_cell(v, resolution, dim);
}

int64_t nautilus_csm_cell_fast(double resolution, uint64_t dim, float v) {
  // host twin of the device cell map (csm::cell_fast); -1 = outside the grid
  static thread_local double c_res = 0;
  static thread_local uint64_t c_dim = 0;
  static thread_local std::vector<float> c_thr;
  static thread_local CellMap c_map;
  if (c_res != resolution || c_dim != dim) {
    c_thr = make_thresholds(resolution, (int)dim, &c_map.qlo, &c_map.qhi);
    c_map.half = (int)dim / 2;
    c_map.inv_res = (float)(1.0 / resolution);
    c_res = resolution; c_dim = dim;
  }
  c_map.lowf = c_thr.data();
  return cell_fast(c_map, v);
}

static bool host_cell_clear(const CellMap& m, float v, float eps, float qmax, int* cell) {
  const volatile float e = v * m.inv_res;  // one rounded float product, like __fmul_rn
  const float ev = e, fl = floorf(ev), fr = ev - fl;
  *cell = (int)fl + m.half;
  return fabsf(ev) < qmax && fr > eps && fr < 1.0f - eps;
}

int nautilus_csm_regular_check(NautilusCsm* h, int64_t n_samples, uint64_t seed,
                               int64_t* n_checked, int64_t* n_bad_low, int64_t* n_bad_high) {
  if (!h || !n_checked || !n_bad_low || !n_bad_high) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  CellMap ml = h->cm_low, mh = h->cm_high;
  ml.lowf = h->thr_low_host.data();
  mh.lowf = h->thr_high_host.data();
  const SearchGrid& sg = h->grid;
  const int n_t = sg.n_t, nh_max = sg.n_h_max;
  const double vmax = h->params.scanner_range;
  uint64_t st = seed * 6364136223846793005ull + 1442695040888963407ull;
  auto rnd = [&]() {  // uniform in [0, 1)
    st = st * 6364136223846793005ull + 1442695040888963407ull;
    return (double)(st >> 11) * (1.0 / 9007199254740992.0);
  };
  int64_t checked = 0, bad_l = 0, bad_h = 0;
  for (int64_t s = 0; s < n_samples; ++s) {
    // a rotated point coordinate: uniform, or hugging a cell edge of either grid
    const int mode = (int)(s % 3);
    const int k0 = (int)(rnd() * n_t) % n_t;
    const int wi = (int)(rnd() * n_t) % n_t;
    float p;
    if (mode == 0) {
      p = (float)((2.0 * rnd() - 1.0) * vmax);
    } else {
      const double res = mode == 1 ? h->params.low_res : h->params.high_res;
      const float t0 = mode == 1 ? h->t_low_host[k0] : h->t_high_host[(size_t)wi * nh_max];
      const double edge = floor((2.0 * rnd() - 1.0) * vmax / res) * res;
      p = (float)(edge - (double)t0 + (2.0 * rnd() - 1.0) * 0.01 * res);
      if (rnd() < 0.25) p = nextafterf((float)(edge - (double)t0), rnd() < 0.5 ? -INFINITY : INFINITY);
    }
    int c0;
    if (host_cell_clear(ml, p + h->t_low_host[k0], sg.reg_eps, sg.reg_qmax, &c0)) {
      for (int k = k0; k < std::min(n_t, k0 + 16); ++k) {
        const int want = cell_fast(ml, p + h->t_low_host[k]);
        const int got = c0 + (k - k0);
        const int got_in = got >= 0 && got < h->geo.wl ? got : -1;
        ++checked;
        if (want != got_in) ++bad_l;
      }
    }
    const float* th = h->t_high_host.data() + (size_t)wi * nh_max;
    if (host_cell_clear(mh, p + th[0], sg.reg_eps_h, sg.reg_qmax_h, &c0)) {
      for (int k = 0; k < h->n_h_host[wi]; ++k) {
        const int want = cell_fast(mh, p + th[k]);
        const int got = c0 + k;
        const int got_in = got >= 0 && got < h->geo.w ? got : -1;
        ++checked;
        if (want != got_in) ++bad_h;
      }
    }
  }
  *n_checked = checked; *n_bad_low = bad_l; *n_bad_high = bad_h;
  return NAUTILUS_OK;
}

void nautilus_csm_rotation_window(double rot_a, double rot_b, double restriction,
                                  double* rot_min, double* rot_max) {
  // math_util::AngleDiff / AngleMod (math_util.h:79-88), cc:252-257
  const double two_pi = 2.0 * M_PI;
  double d = rot_a - rot_b;
  d -= two_pi * rint(d / two_pi);
  double lo = d - restriction / 2;
  lo -= two_pi * rint(lo / two_pi);
  double hi = d + restriction / 2;
  hi -= two_pi * rint(hi / two_pi);
  if (hi < lo) hi += 2 * M_PI;
  if (rot_min) *rot_min = lo;
  if (rot_max) *rot_max = hi;
}

int nautilus_csm_create(const NautilusCsmParams* p, NautilusCsm** out) {
  if (!p || !out) return fail(NAUTILUS_EINVAL, "null argument");
  *out = nullptr;
  if (!(p->scanner_range > 0) || !(p->low_res > 0) || !(p->high_res > 0) || !(p->trans_range > 0))
    return fail(NAUTILUS_EINVAL, "ranges and resolutions must be positive");
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0)
    return fail(NAUTILUS_ECUDA, std::string("no CUDA device: ") + cudaGetErrorString(e));
  if (p->device < 0 || p->device >= n_dev) return fail(NAUTILUS_EINVAL, "bad device ordinal");
  CU_TRY(cudaSetDevice(p->device));
  cudaDeviceProp prop;
  CU_TRY(cudaGetDeviceProperties(&prop, p->device));
  if (prop.major < 10)
    return fail(NAUTILUS_ECUDA, "device is not sm_100-class; this library is built for sm_100a only");
  const double wd = floor((p->scanner_range * 2.0) / p->high_res);  // h:32-35
  const double wld = floor((p->scanner_range * 2.0) / p->low_res);
  if (wd < 1 || wd > 30000 || wld < 1 || wld > 65000)
    return fail(NAUTILUS_EINVAL, "table dimensions out of the supported range");
  if (p->world_size > 1 && (p->world_rank < 0 || p->world_rank >= p->world_size))
    return fail(NAUTILUS_EINVAL, "world_rank outside [0, world_size)");
  NautilusCsm* h = new NautilusCsm();
  h->params = *p;
  h->device = p->device;
  h->sm_count = prop.multiProcessorCount;
  h->rot_step = p->rot_step > 0 ? p->rot_step : M_PI / 180;
  h->world_size = p->world_size > 1 ? p->world_size : 1;
  h->world_rank = p->world_size > 1 ? p->world_rank : 0;
  h->geo.w = (int)wd;
  h->geo.wl = (int)wld;
  h->geo.res_h = p->high_res;
  h->geo.res_l = p->low_res;
  h->geo.bm_stride = (h->geo.w + 31) / 32;
  h->geo.margin = p->margin_cells < 0 ? -1 : (p->margin_cells == 0 ? 128 : p->margin_cells);
  // blur(sigma, sigma, 0, true, true): sigma is a float parameter there
  const float sigma_f = (float)(p->blur_sigma > 0 ? p->blur_sigma : 4.0);
  if (sigma_f < 0.5f) { delete h; return fail(NAUTILUS_EINVAL, "blur_sigma < 0.5 is a no-op in CImg; unsupported"); }
  vanvliet_coeffs((double)sigma_f, &h->vv);
  if (cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete h;
    return fail(NAUTILUS_ECUDA, "cudaStreamCreate failed");
  }
  h->stream = h->own_stream;
  {
    int lo = 0, hi = 0;  // table builds get the highest priority: they fill the slots
    cudaDeviceGetStreamPriorityRange(&lo, &hi);  // the searching kernels leave free
    if (cudaStreamCreateWithPriority(&h->table_stream, cudaStreamNonBlocking, hi) != cudaSuccess)
      h->table_stream = nullptr;
  }
  h->geo.tail = nullptr;
  h->geo.tight = 0;
  if (h->geo.margin >= 0) {
    const std::vector<double> tail = blur_tail(h->vv, h->geo.margin);
    if (h->tail_dev.reserve(tail.size()) ||
        cudaMemcpy(h->tail_dev.p, tail.data(), sizeof(double) * tail.size(), cudaMemcpyHostToDevice) != cudaSuccess) {
      nautilus_csm_destroy(h);
      return fail(NAUTILUS_ECUDA, "upload of the blur tail failed");
    }
    h->geo.tail = h->tail_dev.p;
  }
  h->pool_host = make_pool_list(p->scanner_range, p->low_res, p->high_res, (uint64_t)h->geo.w,
                                (uint64_t)h->geo.wl);
  h->geo.n_pool = (int)h->pool_host.s.size();
  int rc = h->pool_dev.upload(h->pool_host);
  if (rc) { nautilus_csm_destroy(h); return rc; }
  // A translation grid the memo cannot index (e.g. trans_range > scanner_range)
  // only disables matching; the table entry points stay usable.
  h->grid_rc = build_search_grid(h);
  if (h->grid_rc == NAUTILUS_ENOMEM || h->grid_rc == NAUTILUS_ECUDA) {
    rc = h->grid_rc;
    nautilus_csm_destroy(h);
    return rc;
  }
  if (h->grid_rc) h->grid_error = error_text();
  *out = h;
  return NAUTILUS_OK;
}

void nautilus_csm_destroy(NautilusCsm* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  nautilus_internal_nn_cancel(h);
  nautilus_internal_comm_release(h);
  h->gather_dev.release(); h->gather_pin.release();
  free_slots(h);
  h->nn.release();
  h->pool_dev.release();
  h->slots_dev.release(); h->metas_dev.release(); h->dense_scratch.release();
  h->scalar_scratch.release();
  h->thr_low_dev.release(); h->thr_high_dev.release(); h->tail_dev.release();
  h->t_low_dev.release(); h->t_high_dev.release(); h->slot_rank_dev.release(); h->n_h_dev.release();
  h->user_scans.release(); h->tmp_scans.release();
  h->table_scan_dev.release(); h->pairs_dev.release(); h->rots_dev.release();
  h->results_dev.release(); h->chunk_best_dev.release(); h->chunk_rot_dev.release();
  h->residual_arena.release();
  h->spec_pairs_dev.release(); h->spec_items_dev.release(); h->spec_out_dev.release(); h->spec_counters_dev.release();
  h->single_pin.release(); h->user_scans.offsets_pin.release(); h->tmp_scans.offsets_pin.release();
  h->targets_pin.release(); h->sorted_pin.release(); h->results_pin.release(); h->spec_counts_pin.release();
  for (cudaEvent_t ev : h->events) cudaEventDestroy(ev);
  if (h->table_stream) cudaStreamDestroy(h->table_stream);
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  delete h;
}

int nautilus_csm_set_stream(NautilusCsm* h, void* cuda_stream) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  h->stream = cuda_stream ? (cudaStream_t)cuda_stream : h->own_stream;
  return NAUTILUS_OK;
}

int nautilus_csm_table_dims(const NautilusCsm* h, uint64_t* w_high, uint64_t* w_low) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  if (w_high) *w_high = (uint64_t)h->geo.w;
  if (w_low) *w_low = (uint64_t)h->geo.wl;
  return NAUTILUS_OK;
}

int nautilus_csm_upload_scans(NautilusCsm* h, const float* xy, const int64_t* offsets,
                              int32_t n_scans) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  CU_TRY(cudaSetDevice(h->device));
  h->nn.built = false;  // the NN index belongs to the previous scan set
  nautilus_internal_nn_cancel(h);  // a begun host build reads the set that is being replaced
  int rc = upload_scan_set(h, h->user_scans, xy, offsets, n_scans);
  if (rc) return rc;
  // the host copy the NN path needs for its KD-tree layout is fetched from the
  // device when nautilus_nn_build asks for it (nautilus_internal_host_xy)
  h->user_scans.host_xy.clear();
  return NAUTILUS_OK;
}

int nautilus_host_alloc(uint64_t bytes, void** out) {
  if (!out) return fail(NAUTILUS_EINVAL, "null argument");
  *out = nullptr;
  if (bytes == 0) return NAUTILUS_OK;
  CU_TRY(cudaHostAlloc(out, (size_t)bytes, cudaHostAllocPortable));
  return NAUTILUS_OK;
}

int nautilus_host_free(void* p) {
  if (p) CU_TRY(cudaFreeHost(p));
  return NAUTILUS_OK;
}

int nautilus_csm_last_timings(const NautilusCsm* h, NautilusCsmTimings* out) {
  if (!h || !out) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  *out = h->timings;
  return NAUTILUS_OK;
}

}  // extern "C"

namespace {

int require_tiling(NautilusCsm* h) {
  if (!h->pool_host.tiles)
    return fail(NAUTILUS_EINVAL,
                "low/high resolution pair whose pooling windows do not tile the grid is not "
                "supported by the fused normalise+pool kernel");
  return NAUTILUS_OK;
}

}  // namespace

// Core of GetTransformation for a batch over one scan set, first half: validates,
// plans and launches everything; the records land in h->results_dev (room for
// result_capacity >= n_pairs of them) in caller order.  Nothing is copied back and
// the stream is not synchronised: nautilus_internal_match_collect does that, and a
// sharded call puts its all-gather in between.
int nautilus_internal_match_launch(NautilusCsm* h, ScanSet& ss, const int32_t* a_ids,
                                   const int32_t* b_ids, const double* rot_min,
                                   const double* rot_max, int32_t n_pairs,
                                   size_t result_capacity) {
  if (n_pairs < 0) return fail(NAUTILUS_EINVAL, "n_pairs < 0");
  const HostTimer plan_timer;
  h->timings = NautilusCsmTimings{};
  h->pending = PendingBatch{};
  if (n_pairs == 0) return NAUTILUS_OK;
  if (!a_ids || !b_ids || !rot_min || !rot_max) return fail(NAUTILUS_EINVAL, "null argument");
  if (ss.n == 0) return fail(NAUTILUS_ESTATE, "no scans uploaded");
  if (h->grid_rc) return fail(h->grid_rc, h->grid_error);
  int rc = require_tiling(h);
  if (rc) return rc;
  const int n_t = h->grid.n_t, n_tr = n_t * n_t;

  // ---- validate, and find the distinct targets (ascending scan id) without
  // sorting: the table kernels only need that list, so they are launched first
  // and the rest of the host-side planning below overlaps them.
  std::vector<int>& tindex = h->plan_tindex;
  tindex.assign(ss.n, -1);
  for (int p = 0; p < n_pairs; ++p) {
    if (a_ids[p] < 0 || a_ids[p] >= ss.n || b_ids[p] < 0 || b_ids[p] >= ss.n)
      return fail(NAUTILUS_EINVAL, "scan id out of range");
    if (ss.host_offsets[a_ids[p] + 1] == ss.host_offsets[a_ids[p]])
      return fail(NAUTILUS_EEMPTY, "empty query cloud: the reference does not terminate on it");
    if (!(rot_min[p] < rot_max[p]))
      return fail(NAUTILUS_EEMPTY, "empty rotation window: the reference does not terminate on it");
    tindex[b_ids[p]] = 0;
  }
  // distinct target scans, ascending: every target's tables are built once per call
  if ((rc = h->targets_pin.reserve(std::min<size_t>(ss.n, n_pairs)))) return rc;
  int* targets = h->targets_pin.p;
  int n_targets = 0;
  for (int s = 0; s < ss.n; ++s)
    if (tindex[s] == 0) { tindex[s] = n_targets; targets[n_targets++] = s; }

  rc = ensure_slots(h, n_targets);
  if (rc) return rc;
  // ---- waves.  A batch whose targets fit the resident slots is one wave.  With
  // more targets than slots the slots are reused wave after wave.  Splitting a
  // batch that fits into several waves (NAUTILUS_CSM_WAVES) builds the tables of
  // wave w+1 on a second stream while wave w is searched.
  const int n_slots = h->n_slots;
  const bool ring = n_targets > n_slots;
  int wave_tables;
  if (ring) {
    wave_tables = n_slots;
  } else {
    int n_sub = 1;
#ifdef NAUTILUS_CSM_TUNING
    if (const char* env = getenv("NAUTILUS_CSM_WAVES")) n_sub = atoi(env);
#endif
    n_sub = std::max(1, std::min(n_sub, n_targets));
    wave_tables = (n_targets + n_sub - 1) / n_sub;
  }
  const int n_waves = (n_targets + wave_tables - 1) / wave_tables;
  auto slot_base = [&](int w) { return ring ? 0 : w * wave_tables; };

  if ((rc = h->table_scan_dev.reserve(n_targets))) return rc;
  cudaStream_t st = h->stream;
  CU_TRY(cudaMemcpyAsync(h->table_scan_dev.p, targets, sizeof(int) * n_targets,
                         cudaMemcpyHostToDevice, st));
  int launches = 0;
  // events: 0 = start; per wave w: 1+5w tables begin, +1 tables end (table stream),
  // +2 search begin, +3 low-res end, +4 refine end (caller stream)
  auto evt = [&](int w, int k) { return get_event(h, 1 + 5 * (size_t)w + k); };
  cudaStream_t tst = h->table_stream ? h->table_stream : st;
  CU_TRY(cudaEventRecord(get_event(h, 0), st));
  CU_TRY(cudaStreamWaitEvent(tst, get_event(h, 0), 0));  // the target list is uploaded on st
  auto issue_tables = [&](int w) -> int {
    const int t0 = w * wave_tables, t1 = std::min(n_targets, t0 + wave_tables);
    const int base = slot_base(w);
    if (ring && w >= 1) CU_TRY(cudaStreamWaitEvent(tst, evt(w - 1, 4), 0));  // slots still being searched
    CU_TRY(cudaEventRecord(evt(w, 0), tst));
    launch_table_wave(h->geo, h->vv, ss.points, ss.offsets, h->table_scan_dev.p + t0,
                      h->slots_dev.p + base, h->metas_dev.p + base, h->pool_dev.d, t1 - t0, true,
                      tst, &launches, h->slab_ck != nullptr);
    CU_TRY(cudaEventRecord(evt(w, 1), tst));
    return NAUTILUS_OK;
  };
  if ((rc = issue_tables(0))) return rc;
  if (!ring && n_waves > 1 && (rc = issue_tables(1))) return rc;

  // ---- rotation lists (cc:92-93: double accumulation, used as float), cached
  // per distinct window; sinf/cosf from the host libm like Rotation2Df does.
  std::map<std::pair<double, double>, std::pair<int, int>> rot_cache;
  std::vector<Rotation> rots;
  std::vector<PairWork>& work = h->plan_work;
  work.resize(n_pairs);
  int max_rot = 0;
  for (int p = 0; p < n_pairs; ++p) {
    auto key = std::make_pair(rot_min[p], rot_max[p]);
    auto it = rot_cache.find(key);
    if (it == rot_cache.end()) {
      const int begin = (int)rots.size();
      for (double r = rot_min[p]; r < rot_max[p]; r += h->rot_step) {
        Rotation ro;
        ro.theta = (float)r;
        ro.c = cosf(ro.theta);
        ro.s = sinf(ro.theta);
        ro.pad = 0.f;
        rots.push_back(ro);
        if (rots.size() > 50000000) return fail(NAUTILUS_EINVAL, "rotation list too large");
      }
      it = rot_cache.emplace(key, std::make_pair(begin, (int)rots.size() - begin)).first;
    }
    work[p].a_scan = a_ids[p];
    work[p].rot_begin = it->second.first;
    work[p].n_rot = it->second.second;
    work[p].out_index = p;
    work[p].pad = 0;
    max_rot = std::max(max_rot, work[p].n_rot);
  }

  // ---- group the pairs by target (counting sort over the target index, stable)
  std::vector<int>& first_pair = h->plan_first_pair;  // index into `sorted` of each target's first pair
  first_pair.assign(n_targets + 1, 0);
  for (int p = 0; p < n_pairs; ++p) ++first_pair[tindex[b_ids[p]] + 1];
  for (int t = 0; t < n_targets; ++t) first_pair[t + 1] += first_pair[t];
  if ((rc = h->sorted_pin.reserve(n_pairs))) return rc;
  PairWork* sorted = h->sorted_pin.p;
  {
    std::vector<int>& fill = h->plan_fill;
    fill.assign(first_pair.begin(), first_pair.end() - 1);
    for (int p = 0; p < n_pairs; ++p) {
      const int t = tindex[b_ids[p]];
      PairWork& w = sorted[fill[t]++];
      w = work[p];
      w.table = t;  // made wave-relative below
    }
  }
  int max_wave_pairs = 0;
  for (int w = 0; w < n_waves; ++w) {
    const int t0 = w * wave_tables, t1 = std::min(n_targets, t0 + wave_tables);
    for (int k = first_pair[t0]; k < first_pair[t1]; ++k) sorted[k].table -= t0;
    max_wave_pairs = std::max(max_wave_pairs, first_pair[t1] - first_pair[t0]);
  }

  // ---- theta chunking.  A block's warps take rotations round-robin, so a chunk is a
  // whole number of rounds (low_score_warps() rotations each); a small batch is cut
  // into more chunks until the grid is ~16 waves of resident blocks, which keeps the
  // last partial wave short (1250 pairs on one of 8 GPUs: 3 chunks instead of 1).
  const int ls_warps = low_score_warps();
  const int rounds = (max_rot + ls_warps - 1) / ls_warps;
  const long long want_blocks = 16LL * h->sm_count * 2;
  int chunks = (int)std::max<long long>(1, std::min<long long>(rounds, (want_blocks + max_wave_pairs - 1) / max_wave_pairs));
  while (chunks > 1 && (size_t)max_wave_pairs * chunks * n_tr * 12 > ((size_t)1 << 30)) --chunks;  // memo buffers
  const int rot_per_chunk = ((rounds + chunks - 1) / chunks) * ls_warps;
  chunks = (max_rot + rot_per_chunk - 1) / rot_per_chunk;
  SearchBuffers sb;
  sb.chunks_per_pair = chunks;
  sb.rot_per_chunk = rot_per_chunk;
  if ((rc = h->chunk_best_dev.reserve((size_t)max_wave_pairs * chunks * n_tr))) return rc;
  if ((rc = h->chunk_rot_dev.reserve((size_t)max_wave_pairs * chunks * n_tr))) return rc;
  sb.chunk_best = h->chunk_best_dev.p;
  sb.chunk_rot = h->chunk_rot_dev.p;
  bool rf_serial = false;
#ifdef NAUTILUS_CSM_TUNING
  if (const char* rf_env = getenv("NAUTILUS_CSM_REFINE")) rf_serial = !strcmp(rf_env, "serial");
#endif
  sb.spec_cap = spec_cap_for(n_tr);
  sb.spec_pairs = nullptr; sb.spec_items = nullptr; sb.spec_out = nullptr; sb.spec_counters = nullptr;
  if (!rf_serial) {
    if ((rc = h->spec_pairs_dev.reserve(max_wave_pairs))) return rc;
    if ((rc = h->spec_items_dev.reserve((size_t)max_wave_pairs * spec_cap_for(n_tr)))) return rc;
    if ((rc = h->spec_out_dev.reserve((size_t)max_wave_pairs * spec_cap_for(n_tr)))) return rc;
    if ((rc = h->spec_counters_dev.reserve(2 * (size_t)n_waves))) return rc;
    sb.spec_pairs = h->spec_pairs_dev.p; sb.spec_items = h->spec_items_dev.p;
    sb.spec_out = h->spec_out_dev.p; sb.spec_counters = h->spec_counters_dev.p;
  }

  if (h->geo.wl + 1 > 65535)
    return fail(NAUTILUS_EINVAL, "low-res grid too large for the 16-bit row lists of the search");
  // staged-table budget: three blocks per SM when the translation grid is the
  // reference's 14 x 14, whatever fits otherwise
  int table_cells = h->params.stage_cells > 0 ? h->params.stage_cells : 11776;
  while (table_cells > 1024 && low_score_smem_bytes(n_t, h->geo.wl, table_cells) > 220 * 1024)
    table_cells -= 512;
  if (low_score_smem_bytes(n_t, h->geo.wl, table_cells) > 220 * 1024)
    return fail(NAUTILUS_EINVAL, "translation grid too large for shared memory");

  if ((rc = h->pairs_dev.reserve(n_pairs))) return rc;
  if ((rc = h->rots_dev.reserve(rots.size()))) return rc;
  if ((rc = h->results_dev.reserve(std::max<size_t>(n_pairs, result_capacity)))) return rc;
  CU_TRY(cudaMemcpyAsync(h->pairs_dev.p, sorted, sizeof(PairWork) * n_pairs,
                         cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(h->rots_dev.p, rots.data(), sizeof(Rotation) * rots.size(),
                         cudaMemcpyHostToDevice, st));

  for (int w = 0; w < n_waves; ++w) {
    const int t0 = w * wave_tables, t1 = std::min(n_targets, t0 + wave_tables);
    const int k0 = first_pair[t0], k1 = first_pair[t1];
    const int base = slot_base(w);
    CU_TRY(cudaStreamWaitEvent(st, evt(w, 1), 0));
    CU_TRY(cudaEventRecord(evt(w, 2), st));
    launch_low_score(h->geo, h->grid, h->cm_low, ss.points, ss.offsets, h->slots_dev.p + base,
                     h->metas_dev.p + base, h->pairs_dev.p + k0, h->rots_dev.p, k1 - k0, sb,
                     table_cells, st);
    CU_TRY(cudaEventRecord(evt(w, 3), st));
    if (!rf_serial) sb.spec_counters = h->spec_counters_dev.p + 2 * w;  // the wave's queue counters
    launches += 1 + launch_refine(h->geo, h->grid, h->cm_high, ss.points, ss.offsets,
                                  h->slots_dev.p + base, h->metas_dev.p + base, h->pairs_dev.p + k0,
                                  h->rots_dev.p, k1 - k0, sb, h->results_dev.p, rf_serial,
                                  h->sm_count, st);
    CU_TRY(cudaEventRecord(evt(w, 4), st));
    const int next = ring ? w + 1 : w + 2;  // issued after the event it may wait on is recorded
    if (next < n_waves && (rc = issue_tables(next))) return rc;
  }
  CU_TRY(cudaGetLastError());
  h->pending.n_pairs = n_pairs;
  h->pending.n_waves = n_waves;
  h->pending.launches = launches;
  h->pending.rf_serial = rf_serial;
  h->pending.ms_host_plan = plan_timer.ms();
  h->last_n_pairs = n_pairs;
  return NAUTILUS_OK;
}

// Second half, after the caller synchronised the stream: timings of the launched
// batch from its events and the speculative-queue counters.
int nautilus_internal_match_timings(NautilusCsm* h, const int* spec_counts) {
  const int n_waves = h->pending.n_waves;
  auto evt = [&](int w, int k) { return get_event(h, 1 + 5 * (size_t)w + k); };
  h->timings.launches = h->pending.launches;
  h->timings.waves = n_waves;
  h->timings.ms_host_plan += h->pending.ms_host_plan;
  for (int w = 0; w < n_waves; ++w) h->timings.spec_windows += spec_counts ? spec_counts[2 * w] : 0;
  for (int w = 0; w < n_waves; ++w) {
    float a = 0, b = 0, c = 0;
    cudaEventElapsedTime(&a, evt(w, 0), evt(w, 1));
    cudaEventElapsedTime(&b, evt(w, 2), evt(w, 3));
    cudaEventElapsedTime(&c, evt(w, 3), evt(w, 4));
    h->timings.ms_tables += a; h->timings.ms_low += b; h->timings.ms_refine += c;
  }
  if (n_waves > 0) cudaEventElapsedTime(&h->timings.ms_total, get_event(h, 0), evt(n_waves - 1, 4));
  return NAUTILUS_OK;
}

// Queues the D2H copy of the speculative-queue counters of the launched batch.
int nautilus_internal_queue_spec_counts(NautilusCsm* h, int** out) {
  const int n_waves = h->pending.n_waves;
  int rc;
  if ((rc = h->spec_counts_pin.reserve(2 * (size_t)std::max(1, n_waves)))) return rc;
  int* spec_counts = h->spec_counts_pin.p;
  for (int k = 0; k < 2 * n_waves; ++k) spec_counts[k] = 0;
  if (!h->pending.rf_serial && n_waves > 0)
    CU_TRY(cudaMemcpyAsync(spec_counts, h->spec_counters_dev.p, sizeof(int) * 2 * n_waves,
                           cudaMemcpyDeviceToHost, h->stream));
  *out = spec_counts;
  return NAUTILUS_OK;
}

namespace {

int match_batch_impl(NautilusCsm* h, ScanSet& ss, const int32_t* a_ids, const int32_t* b_ids,
                     const double* rot_min, const double* rot_max, int32_t n_pairs,
                     double* score, float* x, float* y, float* theta, NautilusCsmStats* stats) {
  if (n_pairs > 0 && (!score || !x || !y || !theta)) return fail(NAUTILUS_EINVAL, "null argument");
  int rc = nautilus_internal_match_launch(h, ss, a_ids, b_ids, rot_min, rot_max, n_pairs, 0);
  if (rc || n_pairs == 0) return rc;
  cudaStream_t st = h->stream;
  if ((rc = h->results_pin.reserve(n_pairs))) return rc;
  PairResult* res = h->results_pin.p;
  int* spec_counts = nullptr;
  CU_TRY(cudaMemcpyAsync(res, h->results_dev.p, sizeof(PairResult) * n_pairs,
                         cudaMemcpyDeviceToHost, st));
  if ((rc = nautilus_internal_queue_spec_counts(h, &spec_counts))) return rc;
  const HostTimer wait_timer;
  CU_TRY(cudaStreamSynchronize(st));
  const float ms_wait = wait_timer.ms();
  const HostTimer collect_timer;
  nautilus_internal_match_timings(h, spec_counts);
  int bad = 0;
  for (int p = 0; p < n_pairs; ++p) {
    score[p] = res[p].score; x[p] = res[p].x; y[p] = res[p].y; theta[p] = res[p].theta;
    if (stats) { stats[p].n_low = res[p].n_low; stats[p].n_high = res[p].n_high; stats[p].n_pops = res[p].n_pops; }
    bad |= res[p].pad;
  }
  h->timings.ms_host_wait = ms_wait;
  h->timings.ms_host_collect = collect_timer.ms();
  if (bad) return fail(NAUTILUS_EEMPTY, "a pair had nothing to score");
  return NAUTILUS_OK;
}

}  // namespace

extern "C" {

int nautilus_csm_match_batch(NautilusCsm* h, const int32_t* a_ids, const int32_t* b_ids,
                             const double* rot_min, const double* rot_max, int32_t n_pairs,
                             double* score, float* x, float* y, float* theta,
                             NautilusCsmStats* stats) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  CU_TRY(cudaSetDevice(h->device));
  return match_batch_impl(h, h->user_scans, a_ids, b_ids, rot_min, rot_max, n_pairs, score, x, y,
                          theta, stats);
}

int nautilus_csm_copy_results_dev(NautilusCsm* h, void* dst_dev, int32_t n_pairs) {
  if (!h || !dst_dev) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  if (n_pairs < 0 || n_pairs > h->last_n_pairs)
    return fail(NAUTILUS_ESTATE, "no batch result of that size is resident");
  CU_TRY(cudaSetDevice(h->device));
  CU_TRY(cudaMemcpyAsync(dst_dev, h->results_dev.p, sizeof(PairResult) * (size_t)n_pairs,
                         cudaMemcpyDeviceToDevice, h->stream));
  return NAUTILUS_OK;
}

}  // extern "C"

namespace {

// Matches every queued single-pair request as ONE batch (the caller holds h->mu).
// Clouds that several requests pass by the same pointer and size (one target against
// many queries, GetLocalUncertaintyStats' pattern, cc:381-386) are uploaded once, so
// their lookup tables are built once.
void run_single_batch(NautilusCsm* h, std::vector<SingleRequest*>& reqs) {
  const int n = (int)reqs.size();
  std::vector<int32_t> a_ids, b_ids;
  std::vector<double> rmin, rmax;
  std::vector<int> live;
  std::vector<int64_t> offs(1, 0);
  std::vector<std::pair<const float*, int64_t>> clouds;  // distinct (pointer, size)
  auto scan_of = [&](const float* p, int64_t cnt) {
    for (size_t k = 0; k < clouds.size(); ++k)
      if (clouds[k].first == p && clouds[k].second == cnt) return (int32_t)k;
    clouds.emplace_back(p, cnt);
    offs.push_back(offs.back() + cnt);
    return (int32_t)(clouds.size() - 1);
  };
  for (int k = 0; k < n; ++k) {
    SingleRequest& r = *reqs[k];
    r.rc = NAUTILUS_OK;
    if (r.na < 0 || r.nb < 0 || (r.na > 0 && !r.a) || (r.nb > 0 && !r.b)) {
      r.rc = NAUTILUS_EINVAL; r.error = "bad cloud";
    } else if (r.na == 0) {
      r.rc = NAUTILUS_EEMPTY; r.error = "empty query cloud: the reference does not terminate on it";
    } else if (!(r.rot_min < r.rot_max)) {
      r.rc = NAUTILUS_EEMPTY; r.error = "empty rotation window: the reference does not terminate on it";
    }
    if (r.rc) continue;
    live.push_back(k);
    a_ids.push_back(scan_of(r.a, r.na));
    b_ids.push_back(scan_of(r.b, r.nb));
    rmin.push_back(r.rot_min);
    rmax.push_back(r.rot_max);
  }
  if (live.empty()) return;
  const int m = (int)live.size();
  int rc = cudaSetDevice(h->device) == cudaSuccess ? NAUTILUS_OK : fail(NAUTILUS_ECUDA, "cudaSetDevice");
  // clouds -> page-locked staging -> one H2D
  if (!rc) rc = h->single_pin.reserve((size_t)2 * std::max<int64_t>(offs.back(), 1));
  if (!rc) {
    for (size_t k = 0; k < clouds.size(); ++k)
      if (clouds[k].second)
        memcpy(h->single_pin.p + 2 * offs[k], clouds[k].first, sizeof(float) * 2 * clouds[k].second);
    rc = upload_scan_set(h, h->tmp_scans, h->single_pin.p, offs.data(), (int32_t)clouds.size(), true);
  }
  std::vector<double> score(m);
  std::vector<float> x(m), y(m), th(m);
  std::vector<NautilusCsmStats> st(m);
  if (!rc)
    rc = match_batch_impl(h, h->tmp_scans, a_ids.data(), b_ids.data(), rmin.data(), rmax.data(), m,
                          score.data(), x.data(), y.data(), th.data(), st.data());
  const std::string err = rc ? error_text() : std::string();
  for (int j = 0; j < m; ++j) {
    SingleRequest& r = *reqs[live[j]];
    r.rc = rc;
    r.error = err;
    if (rc) continue;
    if (r.score) *r.score = score[j];
    if (r.x) *r.x = x[j];
    if (r.y) *r.y = y[j];
    if (r.theta) *r.theta = th[j];
    if (r.stats) *r.stats = st[j];
  }
}

}  // namespace

extern "C" {

// GetTransformation for one pair of host clouds.  The reference object is shared by
// OpenMP threads (bag_avg_matcher.cc:104-141, CorrelativeScanMatcher.cc:381-386) and a
// single pair cannot fill a B200, so concurrent callers are COMBINED: a caller queues
// its request; whoever finds no batch in flight becomes the leader, takes everything
// queued so far and matches it as one batch, then hands the results back and repeats
// while requests keep arriving.  A lone caller is a batch of one.
int nautilus_csm_match(NautilusCsm* h, const float* a, int64_t na, const float* b, int64_t nb,
                       double rot_min, double rot_max, double* score, float* x, float* y,
                       float* theta, NautilusCsmStats* stats) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  SingleRequest req;
  req.a = a; req.na = na; req.b = b; req.nb = nb; req.rot_min = rot_min; req.rot_max = rot_max;
  req.score = score; req.x = x; req.y = y; req.theta = theta; req.stats = stats;
  std::unique_lock<std::mutex> q(h->single_mu);
  h->single_queue.push_back(&req);
  if (h->single_leader) {
    h->single_cv.wait(q, [&] { return req.done; });
  } else {
    h->single_leader = true;
    while (!h->single_queue.empty()) {
      std::vector<SingleRequest*> batch;
      batch.swap(h->single_queue);
      q.unlock();
      {
        NAUTILUS_GUARD(h);
        run_single_batch(h, batch);
        h->single_batches += 1;
        h->single_requests += (long long)batch.size();
      }
      q.lock();
      for (SingleRequest* r : batch) r->done = true;
      h->single_cv.notify_all();
    }
    h->single_leader = false;
  }
  q.unlock();
  if (req.rc) return fail(req.rc, req.error);
  return NAUTILUS_OK;
}

int nautilus_csm_single_queue_stats(const NautilusCsm* h, int64_t* n_requests, int64_t* n_batches) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  std::lock_guard<std::mutex> q(h->single_mu);
  if (n_requests) *n_requests = h->single_requests;
  if (n_batches) *n_batches = h->single_batches;
  return NAUTILUS_OK;
}

int nautilus_csm_uncertainty(NautilusCsm* h, const float* a, int64_t na, const float* b, int64_t nb,
                             int32_t dims, double rotation, float* out) {
  if (!h || !out) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  if (dims != 2 && dims != 3) return fail(NAUTILUS_EINVAL, "dims must be 2 or 3");
  if (na <= 0 || nb < 0) return fail(NAUTILUS_EEMPTY, "empty query cloud");
  if (h->grid_rc) return fail(h->grid_rc, h->grid_error);
  CU_TRY(cudaSetDevice(h->device));
  int rc = require_tiling(h);
  if (rc) return rc;
  std::vector<float> xy((size_t)2 * (na + nb));
  memcpy(xy.data(), a, sizeof(float) * 2 * na);
  if (nb) memcpy(xy.data() + 2 * na, b, sizeof(float) * 2 * nb);
  const int64_t offs[3] = {0, na, na + nb};
  if ((rc = upload_scan_set(h, h->tmp_scans, xy.data(), offs, 2))) return rc;
  ScanSet& ss = h->tmp_scans;
  // rotations: 0 .. 2 pi (cc:281,286) or the single given rotation
  // (MemoizeLowRes(rotation, rotation + EPSILON), cc:341-343)
  const double rmin = dims == 3 ? 0.0 : rotation;
  const double rmax = dims == 3 ? 2 * M_PI : rotation + CSM_EPSILON;
  std::vector<Rotation> rots;
  for (double r = rmin; r < rmax; r += h->rot_step) {
    Rotation ro;
    ro.theta = (float)r; ro.c = cosf(ro.theta); ro.s = sinf(ro.theta); ro.pad = 0.f;
    rots.push_back(ro);
    if (rots.size() > 1000000) return fail(NAUTILUS_EINVAL, "rotation list too large");
  }
  const int n_rot = (int)rots.size(), n_t = h->grid.n_t, n_tr = n_t * n_t;
  if (n_rot == 0) return fail(NAUTILUS_EEMPTY, "empty rotation window");
  PairWork pw;
  pw.a_scan = 0; pw.table = 0; pw.rot_begin = 0; pw.n_rot = n_rot; pw.out_index = 0; pw.pad = 0;
  if ((rc = ensure_slots(h, 1))) return rc;
  const int chunks_want = std::max(1, std::min(n_rot, h->sm_count * 2));
  SearchBuffers sb;
  sb.rot_per_chunk = (n_rot + chunks_want - 1) / chunks_want;
  sb.chunks_per_pair = (n_rot + sb.rot_per_chunk - 1) / sb.rot_per_chunk;
  if ((rc = h->chunk_best_dev.reserve((size_t)sb.chunks_per_pair * n_tr))) return rc;
  if ((rc = h->chunk_rot_dev.reserve((size_t)sb.chunks_per_pair * n_tr))) return rc;
  sb.chunk_best = h->chunk_best_dev.p;
  sb.chunk_rot = h->chunk_rot_dev.p;
  int table_cells = h->params.stage_cells > 0 ? h->params.stage_cells : 11776;
  while (table_cells > 1024 && low_score_smem_bytes(n_t, h->geo.wl, table_cells) > 220 * 1024)
    table_cells -= 512;
  if (low_score_smem_bytes(n_t, h->geo.wl, table_cells) > 220 * 1024)
    return fail(NAUTILUS_EINVAL, "translation grid too large for shared memory");
  if ((rc = h->table_scan_dev.reserve(1))) return rc;
  if ((rc = h->pairs_dev.reserve(1))) return rc;
  if ((rc = h->rots_dev.reserve(rots.size()))) return rc;
  if ((rc = h->dense_scratch.reserve((size_t)n_rot * n_tr + 16))) return rc;
  cudaStream_t st = h->stream;
  const int one = 1;  // the table of cloud b
  CU_TRY(cudaMemcpyAsync(h->table_scan_dev.p, &one, sizeof(int), cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(h->pairs_dev.p, &pw, sizeof(PairWork), cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(h->rots_dev.p, rots.data(), sizeof(Rotation) * rots.size(),
                         cudaMemcpyHostToDevice, st));
  launch_table_wave(h->geo, h->vv, ss.points, ss.offsets, h->table_scan_dev.p, h->slots_dev.p,
                    h->metas_dev.p, h->pool_dev.d, 1, true, st, nullptr, h->slab_ck != nullptr);
  launch_low_score(h->geo, h->grid, h->cm_low, ss.points, ss.offsets, h->slots_dev.p,
                   h->metas_dev.p, h->pairs_dev.p, h->rots_dev.p, 1, sb, table_cells, st);
  double* costs = h->dense_scratch.p;
  float* out_dev = reinterpret_cast<float*>(costs + (size_t)n_rot * n_tr);
  launch_uncertainty(h->grid, h->cm_high, ss.points, ss.offsets, h->slots_dev.p, h->metas_dev.p, pw,
                     h->rots_dev.p, sb, log(1e-4) /* UNCERTAINTY_USELESS_THRESHOLD, cc:12 */, costs,
                     dims, out_dev, st);
  CU_TRY(cudaGetLastError());
  CU_TRY(cudaMemcpyAsync(out, out_dev, sizeof(float) * dims * dims, cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  return NAUTILUS_OK;
}

}  // extern "C"

namespace {
int uncertainty_batch_impl(NautilusCsm* h, ScanSet& ss, const int32_t* a_ids, const int32_t* b_ids,
                           const double* rotation, int32_t n_pairs, float* out) {
  if (n_pairs < 0) return fail(NAUTILUS_EINVAL, "n_pairs < 0");
  if (n_pairs == 0) return NAUTILUS_OK;
  if (!a_ids || !b_ids || !rotation || !out) return fail(NAUTILUS_EINVAL, "null argument");
  if (ss.n == 0) return fail(NAUTILUS_ESTATE, "no scans uploaded");
  if (h->grid_rc) return fail(h->grid_rc, h->grid_error);
  CU_TRY(cudaSetDevice(h->device));
  int rc = require_tiling(h);
  if (rc) return rc;
  const int n_t = h->grid.n_t, n_tr = n_t * n_t;
  // distinct targets, ascending; pairs grouped by target (stable)
  std::vector<int> tindex(ss.n, -1), targets;
  for (int p = 0; p < n_pairs; ++p) {
    if (a_ids[p] < 0 || a_ids[p] >= ss.n || b_ids[p] < 0 || b_ids[p] >= ss.n)
      return fail(NAUTILUS_EINVAL, "scan id out of range");
    if (ss.host_offsets[a_ids[p] + 1] == ss.host_offsets[a_ids[p]])
      return fail(NAUTILUS_EEMPTY, "empty query cloud");
    tindex[b_ids[p]] = 0;
  }
  for (int s = 0; s < ss.n; ++s)
    if (tindex[s] == 0) { tindex[s] = (int)targets.size(); targets.push_back(s); }
  const int n_targets = (int)targets.size();
  if ((rc = ensure_slots(h, n_targets))) return rc;
  const int wave_tables = std::min(h->n_slots, n_targets);
  std::vector<int> first(n_targets + 1, 0);
  for (int p = 0; p < n_pairs; ++p) ++first[tindex[b_ids[p]] + 1];
  for (int t = 0; t < n_targets; ++t) first[t + 1] += first[t];
  std::vector<PairWork> sorted(n_pairs);
  std::vector<Rotation> rots(n_pairs);
  {
    std::vector<int> fill(first.begin(), first.end() - 1);
    for (int p = 0; p < n_pairs; ++p) {
      const int t = tindex[b_ids[p]];
      const int k = fill[t]++;
      PairWork& w = sorted[k];
      w.a_scan = a_ids[p]; w.table = t % wave_tables; w.rot_begin = k; w.n_rot = 1;
      w.out_index = p; w.pad = 0;
      // MemoizeLowRes(rotation, rotation + EPSILON) (cc:341-343): the one rotation, as float
      rots[k].theta = (float)rotation[p]; rots[k].c = cosf(rots[k].theta);
      rots[k].s = sinf(rots[k].theta); rots[k].pad = 0.f;
    }
  }
  int max_wave_pairs = 0;
  for (int t0 = 0; t0 < n_targets; t0 += wave_tables)
    max_wave_pairs = std::max(max_wave_pairs, first[std::min(n_targets, t0 + wave_tables)] - first[t0]);
  SearchBuffers sb{};
  sb.chunks_per_pair = 1;
  sb.rot_per_chunk = low_score_warps();
  if ((rc = h->chunk_best_dev.reserve((size_t)max_wave_pairs * n_tr))) return rc;
  if ((rc = h->chunk_rot_dev.reserve((size_t)max_wave_pairs * n_tr))) return rc;
  sb.chunk_best = h->chunk_best_dev.p;
  sb.chunk_rot = h->chunk_rot_dev.p;
  int table_cells = h->params.stage_cells > 0 ? h->params.stage_cells : 11776;
  while (table_cells > 1024 && low_score_smem_bytes(n_t, h->geo.wl, table_cells) > 220 * 1024)
    table_cells -= 512;
  if (low_score_smem_bytes(n_t, h->geo.wl, table_cells) > 220 * 1024)
    return fail(NAUTILUS_EINVAL, "translation grid too large for shared memory");
  if ((rc = h->table_scan_dev.reserve(n_targets))) return rc;
  if ((rc = h->pairs_dev.reserve(n_pairs))) return rc;
  if ((rc = h->rots_dev.reserve(n_pairs))) return rc;
  if ((rc = h->dense_scratch.reserve((size_t)max_wave_pairs * n_tr + 2 * (size_t)n_pairs + 16))) return rc;
  cudaStream_t st = h->stream;
  CU_TRY(cudaMemcpyAsync(h->table_scan_dev.p, targets.data(), sizeof(int) * n_targets, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(h->pairs_dev.p, sorted.data(), sizeof(PairWork) * n_pairs, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(h->rots_dev.p, rots.data(), sizeof(Rotation) * n_pairs, cudaMemcpyHostToDevice, st));
  double* costs = h->dense_scratch.p;
  float* out_dev = reinterpret_cast<float*>(costs + (size_t)max_wave_pairs * n_tr);
  for (int t0 = 0; t0 < n_targets; t0 += wave_tables) {
    const int t1 = std::min(n_targets, t0 + wave_tables);
    const int k0 = first[t0], k1 = first[t1];
    launch_table_wave(h->geo, h->vv, ss.points, ss.offsets, h->table_scan_dev.p + t0, h->slots_dev.p,
                      h->metas_dev.p, h->pool_dev.d, t1 - t0, true, st, nullptr, h->slab_ck != nullptr);
    launch_low_score(h->geo, h->grid, h->cm_low, ss.points, ss.offsets, h->slots_dev.p, h->metas_dev.p,
                     h->pairs_dev.p + k0, h->rots_dev.p, k1 - k0, sb, table_cells, st);
    launch_uncertainty_batch(h->grid, h->cm_high, ss.points, ss.offsets, h->slots_dev.p, h->metas_dev.p,
                             h->pairs_dev.p + k0, h->rots_dev.p, k1 - k0, sb,
                             log(1e-4) /* UNCERTAINTY_USELESS_THRESHOLD, cc:12 */, costs, out_dev, st);
  }
  CU_TRY(cudaGetLastError());
  CU_TRY(cudaMemcpyAsync(out, out_dev, sizeof(float) * 4 * (size_t)n_pairs, cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  return NAUTILUS_OK;
}
}  // namespace

extern "C" {

int nautilus_csm_uncertainty_batch(NautilusCsm* h, const int32_t* a_ids, const int32_t* b_ids,
                                   const double* rotation, int32_t n_pairs, float* out) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  return uncertainty_batch_impl(h, h->user_scans, a_ids, b_ids, rotation, n_pairs, out);
}

int nautilus_csm_local_uncertainty(NautilusCsm* h, const float* target_xy, int64_t n_target,
                                   const float* comp_xy, const int64_t* comp_offsets,
                                   int32_t n_comp, float* u_out, float* theta_out) {
  if (!h || !u_out) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  if (n_comp < 0 || n_target < 0 || (n_comp > 0 && !comp_offsets))
    return fail(NAUTILUS_EINVAL, "bad cloud list");
  if (n_comp == 0) return NAUTILUS_OK;
  if (n_target == 0) return fail(NAUTILUS_EEMPTY, "empty query cloud: the reference does not terminate on it");
  CU_TRY(cudaSetDevice(h->device));
  // scan 0 = the target cloud (the query of every match, cc:384-385), 1.. = the comparison clouds
  const int64_t n_cpts = comp_offsets[n_comp];
  std::vector<float> xy((size_t)2 * (n_target + n_cpts));
  memcpy(xy.data(), target_xy, sizeof(float) * 2 * n_target);
  if (n_cpts) memcpy(xy.data() + 2 * n_target, comp_xy, sizeof(float) * 2 * n_cpts);
  std::vector<int64_t> offs((size_t)n_comp + 2);
  offs[0] = 0;
  for (int k = 0; k <= n_comp; ++k) offs[k + 1] = n_target + comp_offsets[k];
  int rc = upload_scan_set(h, h->tmp_scans, xy.data(), offs.data(), n_comp + 1);
  if (rc) return rc;
  std::vector<int32_t> a_ids(n_comp, 0), b_ids(n_comp);
  std::vector<double> rmin(n_comp, 0.0), rmax(n_comp, 2 * M_PI), score(n_comp), rot(n_comp);
  std::vector<float> x(n_comp), y(n_comp), th(n_comp);
  for (int k = 0; k < n_comp; ++k) b_ids[k] = k + 1;
  // GetTransformation(targetCloud, comparisonCloud) over the full rotation (cc:384-385) ...
  rc = match_batch_impl(h, h->tmp_scans, a_ids.data(), b_ids.data(), rmin.data(), rmax.data(), n_comp,
                        score.data(), x.data(), y.data(), th.data(), nullptr);
  if (rc) return rc;
  // ... then GetUncertaintyMatrix(targetCloud, comparisonCloud, result.second.second) (cc:386-387)
  for (int k = 0; k < n_comp; ++k) rot[k] = (double)th[k];
  rc = uncertainty_batch_impl(h, h->tmp_scans, a_ids.data(), b_ids.data(), rot.data(), n_comp, u_out);
  if (rc) return rc;
  if (theta_out) memcpy(theta_out, th.data(), sizeof(float) * n_comp);
  return NAUTILUS_OK;
}

int nautilus_csm_highres_table(NautilusCsm* h, const float* xy, int64_t n_points, double* out_high,
                               double* out_low) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  if (n_points < 0 || (n_points > 0 && !xy)) return fail(NAUTILUS_EINVAL, "bad cloud");
  CU_TRY(cudaSetDevice(h->device));
  int rc = require_tiling(h);
  if (rc) return rc;
  const int64_t offs[2] = {0, n_points};
  if ((rc = upload_scan_set(h, h->tmp_scans, xy, offs, 1))) return rc;
  if ((rc = ensure_slots(h, 1))) return rc;
  if ((rc = h->table_scan_dev.reserve(1))) return rc;
  const int zero = 0;
  cudaStream_t st = h->stream;
  CU_TRY(cudaMemcpyAsync(h->table_scan_dev.p, &zero, sizeof(int), cudaMemcpyHostToDevice, st));
  launch_table_wave(h->geo, h->vv, h->tmp_scans.points, h->tmp_scans.offsets, h->table_scan_dev.p,
                    h->slots_dev.p, h->metas_dev.p, h->pool_dev.d, 1, false, st, nullptr);
  const size_t w = h->geo.w, wl = h->geo.wl;
  if (out_high) {
    if ((rc = h->dense_scratch.reserve(w * w))) return rc;
    launch_expand(h->geo, h->slots_dev.p, h->metas_dev.p, 0, h->dense_scratch.p, st);
    CU_TRY(cudaMemcpyAsync(out_high, h->dense_scratch.p, sizeof(double) * w * w,
                           cudaMemcpyDeviceToHost, st));
  }
  if (out_low)
    CU_TRY(cudaMemcpyAsync(out_low, h->slots_host[0].low, sizeof(double) * wl * wl,
                           cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaGetLastError());
  CU_TRY(cudaStreamSynchronize(st));
  return NAUTILUS_OK;
}

int nautilus_csm_lowres_from_table(NautilusCsm* h, const double* high, uint64_t width,
                                   uint64_t height, double resolution, double* out_low) {
  if (!h || !high || !out_low) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  if (width == 0 || height == 0)
    return fail(NAUTILUS_EINVAL, "empty table (CImg::get_crop throws on an empty image)");
  if (!(resolution > 0) || width > 30000 || height > 30000) return fail(NAUTILUS_EINVAL, "bad table");
  CU_TRY(cudaSetDevice(h->device));
  const NautilusCsmParams& p = h->params;
  const size_t wl = h->geo.wl;
  PoolListHost px = make_pool_list(p.scanner_range, p.low_res, resolution, width, wl);
  PoolListHost py = make_pool_list(p.scanner_range, p.low_res, resolution, height, wl);
  DevPool dx, dy;
  int rc = dx.upload(px);
  if (!rc) rc = dy.upload(py);
  if (!rc) rc = h->dense_scratch.reserve(width * height + wl * wl);
  if (rc) { dx.release(); dy.release(); return rc; }
  cudaStream_t st = h->stream;
  double* d_hi = h->dense_scratch.p;
  double* d_low = d_hi + width * height;
  cudaError_t e = cudaMemcpyAsync(d_hi, high, sizeof(double) * width * height, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_low, 0, sizeof(double) * wl * wl, st);
  if (e == cudaSuccess) {
    launch_pool_generic(d_hi, (int)width, (int)height, dx.d, dy.d, d_low, (int)wl, st);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(out_low, d_low, sizeof(double) * wl * wl, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  dx.release(); dy.release();
  if (e != cudaSuccess) return fail(NAUTILUS_ECUDA, cudaGetErrorString(e));
  return NAUTILUS_OK;
}

int nautilus_csm_build_tables(NautilusCsm* h, const int32_t* scan_ids, int32_t n, double* checksum) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  if (n < 0 || (n > 0 && !scan_ids)) return fail(NAUTILUS_EINVAL, "bad scan id list");
  CU_TRY(cudaSetDevice(h->device));
  h->timings = NautilusCsmTimings{};
  if (n == 0) { if (checksum) *checksum = 0; return NAUTILUS_OK; }
  ScanSet& ss = h->user_scans;
  for (int i = 0; i < n; ++i)
    if (scan_ids[i] < 0 || scan_ids[i] >= ss.n) return fail(NAUTILUS_EINVAL, "scan id out of range");
  int rc = require_tiling(h);
  if (rc) return rc;
  if ((rc = ensure_slots(h, n))) return rc;
  if ((rc = h->table_scan_dev.reserve(n))) return rc;
  if ((rc = h->scalar_scratch.reserve(1))) return rc;
  cudaStream_t st = h->stream;
  CU_TRY(cudaMemcpyAsync(h->table_scan_dev.p, scan_ids, sizeof(int) * n, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemsetAsync(h->scalar_scratch.p, 0, sizeof(double), st));
  int launches = 0, waves = 0;
  CU_TRY(cudaEventRecord(get_event(h, 0), st));
  for (int t0 = 0; t0 < n; t0 += h->n_slots, ++waves) {
    const int cnt = std::min(n - t0, h->n_slots);
    launch_table_wave(h->geo, h->vv, ss.points, ss.offsets, h->table_scan_dev.p + t0,
                      h->slots_dev.p, h->metas_dev.p, h->pool_dev.d, cnt, false, st, &launches);
    if (checksum) { launch_low_checksum(h->geo, h->slots_dev.p, cnt, h->scalar_scratch.p, st); ++launches; }
  }
  CU_TRY(cudaEventRecord(get_event(h, 1), st));
  CU_TRY(cudaGetLastError());
  double cs = 0;
  CU_TRY(cudaMemcpyAsync(&cs, h->scalar_scratch.p, sizeof(double), cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  if (checksum) *checksum = cs;
  cudaEventElapsedTime(&h->timings.ms_total, h->events[0], h->events[1]);
  h->timings.ms_tables = h->timings.ms_total;
  h->timings.launches = launches;
  h->timings.waves = waves;
  return NAUTILUS_OK;
}

}  // extern "C"
