/* TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the reference's
 * correlative-scan-matching path in plain C.  It is the checker for the CUDA
 * path and the instrumented (candidate-counting) CPU baseline.  It is never
 * linked into, loaded by or called from the product path (nautilus_b200/).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may use it.
 *
 * Parity status: PINNED.  tests/test_oracle_*.py check this file against
 *   (1) the 11 golden cases of
 *       /root/reference/third_party/csm/test/correlative_scan_matcher_tests.cc:12-106
 *   (2) the unmodified reference compiled into oracle/_ref/ (tables bit-for-bit,
 *       match results bit-for-bit) and committed fixtures generated from it
 *       (tests/golden/, generator tests/golden/make_golden.py).
 *
 * All paths below are relative to /root/reference/third_party/csm/src/.
 * Compile WITHOUT -march / -ffast-math / -ffp-contract=fast so that no FMA is
 * formed, like the reference build (CMakeLists.txt:11,16).
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MIN_VALUE_FOR_LOOKUP 1E-10 /* CorrelativeScanMatcher.h:20 */
#define CSM_EPSILON 1e-6           /* CorrelativeScanMatcher.cc:19 */

/* ---- struct LookupTable (CorrelativeScanMatcher.h:27-110) ------------------ */

/* LookupTable ctor, CorrelativeScanMatcher.h:32-35 */
uint64_t oracle_table_dim(double range, double resolution) {
  return (uint64_t)floor((range * 2.0) / resolution);
}

/* convertX / convertY, CorrelativeScanMatcher.h:43-49.  The reference returns
 * `width / 2 + floor(x / resolution)` (a double) as uint64_t; a negative value
 * turns into a huge unsigned number there (x86-64 cvttsd2si keeps the signed
 * bit pattern) and fails the `>= width` test; here the signed value itself is
 * returned and `inside()` rejects negatives. */
int64_t oracle_cell(float x, double resolution, uint64_t dim) {
  double v = (double)(dim / 2) + floor((double)x / resolution);
  if (v <= -9.0e18) return INT64_MIN;
  if (v >= 9.0e18) return INT64_MAX;
  return (int64_t)v;
}

static inline int inside(int64_t c, uint64_t dim) {
  return c >= 0 && (uint64_t)c < dim;
}

/* GetPointValue, CorrelativeScanMatcher.h:51-59 */
double oracle_get_point_value(const double* values, uint64_t w, uint64_t h,
                              double resolution, float px, float py) {
  int64_t x = oracle_cell(px, resolution, w);
  int64_t y = oracle_cell(py, resolution, h);
  if (!inside(x, w) || !inside(y, h) ||
      values[x + y * (int64_t)w] <= MIN_VALUE_FOR_LOOKUP) {
    return MIN_VALUE_FOR_LOOKUP;
  }
  return values[x + y * (int64_t)w];
}

/* SetPointValue, CorrelativeScanMatcher.h:70-77 */
void oracle_set_point_value(double* values, uint64_t w, uint64_t h,
                            double resolution, float px, float py, double v) {
  int64_t x = oracle_cell(px, resolution, w);
  int64_t y = oracle_cell(py, resolution, h);
  if (!inside(x, w) || !inside(y, h)) return;
  values[x + y * (int64_t)w] = v;
}

/* MaxArea, CorrelativeScanMatcher.h:91-101, with CImg::get_crop's behaviour
 * (CImg.h:45516-45564: reversed bounds are swapped, cells outside the image
 * read as 0 -- Dirichlet) and CImg::max (CImg.h:35236-35260).  The double
 * arguments are narrowed to float by convertX(float). */
double oracle_max_area(const double* values, uint64_t w, uint64_t h,
                       double resolution, double start_x, double start_y,
                       double end_x, double end_y) {
  /* uint64 arithmetic then narrowing to int, as the call
   * get_crop(sx, sy, ex - 1, ey - 1) with int parameters does. */
  uint64_t usx = (uint64_t)oracle_cell((float)start_x, resolution, w);
  uint64_t usy = (uint64_t)oracle_cell((float)start_y, resolution, h);
  uint64_t uex = (uint64_t)oracle_cell((float)end_x, resolution, w);
  uint64_t uey = (uint64_t)oracle_cell((float)end_y, resolution, h);
  int x0 = (int)usx, y0 = (int)usy, x1 = (int)(uex - 1), y1 = (int)(uey - 1);
  int nx0 = x0 < x1 ? x0 : x1, nx1 = x0 < x1 ? x1 : x0;
  int ny0 = y0 < y1 ? y0 : y1, ny1 = y0 < y1 ? y1 : y0;
  int touches_outside =
      nx0 < 0 || ny0 < 0 || nx1 >= (int)w || ny1 >= (int)h;
  int cx0 = nx0 < 0 ? 0 : nx0, cy0 = ny0 < 0 ? 0 : ny0;
  int cx1 = nx1 >= (int)w ? (int)w - 1 : nx1;
  int cy1 = ny1 >= (int)h ? (int)h - 1 : ny1;
  double m = 0.0;
  int have = 0;
  if (touches_outside) { m = 0.0; have = 1; } /* zero fill takes part in max */
  for (int y = cy0; y <= cy1; ++y)
    for (int x = cx0; x <= cx1; ++x) {
      double v = values[(int64_t)x + (int64_t)y * (int64_t)w];
      if (!have || v > m) { m = v; have = 1; }
    }
  return m;
}

/* AbsCoords, CorrelativeScanMatcher.h:104-109 */
uint64_t oracle_abs_coords(uint64_t w, uint64_t h, double resolution, double x,
                           double y) {
  uint64_t row = (uint64_t)(((double)(h / 2) + round(y / resolution)) * (double)w);
  uint64_t col = (uint64_t)((double)(w / 2) + round(x / resolution));
  return row + col;
}

/* ---- CImg numeric kernels used by LookupTable ------------------------------ */

typedef struct {
  double B, a1, a2, a3; /* filter[0..3] of CImg.h:49279-49284 */
  double M[9];          /* Triggs matrix, CImg.h:49071-49081 */
} VanVliet;

/* Coefficients of CImg<T>::vanvliet, CImg.h:49267-49284 (sigma >= 0.5) together
 * with the Triggs matrix of _cimg_recursive_apply, CImg.h:49066-49081. */
void oracle_vanvliet_coeffs(double sigma, VanVliet* f) {
  const double nnsigma = sigma < 0.5 ? 0.5 : sigma, m0 = 1.16680, m1 = 1.10783,
               m2 = 1.40586, m1sq = m1 * m1, m2sq = m2 * m2,
               q = (nnsigma < 3.556
                        ? -0.2568 + 0.5784 * nnsigma + 0.0561 * nnsigma * nnsigma
                        : 2.5091 + 0.9804 * (nnsigma - 3.556)),
               qsq = q * q, scale = (m0 + q) * (m1sq + m2sq + 2 * m1 * q + qsq),
               b1 = -q *
                    (2 * m0 * m1 + m1sq + m2sq + (2 * m0 + 4 * m1) * q + 3 * qsq) /
                    scale,
               b2 = qsq * (m0 + 2 * m1 + 3 * q) / scale, b3 = -qsq * q / scale,
               B = (m0 * (m1sq + m2sq)) / scale;
  f->B = B;
  f->a1 = -b1;
  f->a2 = -b2;
  f->a3 = -b3;
  const double a1 = f->a1, a2 = f->a2, a3 = f->a3,
               scaleM = 1. / ((1. + a1 - a2 + a3) * (1. - a1 - a2 - a3) *
                              (1. + a2 + (a1 - a3) * a3));
  f->M[0] = scaleM * (-a3 * a1 + 1. - a3 * a3 - a2);
  f->M[1] = scaleM * (a3 + a1) * (a2 + a3 * a1);
  f->M[2] = scaleM * a3 * (a1 + a3 * a2);
  f->M[3] = scaleM * (a1 + a3 * a2);
  f->M[4] = -scaleM * (a2 - 1.) * (a2 + a3 * a1);
  f->M[5] = -scaleM * a3 * (a3 * a1 + a3 * a3 + a2 - 1.);
  f->M[6] = scaleM * (a3 * a1 + a2 + a1 * a1 - a2 * a2);
  f->M[7] = scaleM * (a1 * a2 + a3 * a2 * a2 - a1 * a3 * a3 - a3 * a3 * a3 -
                      a3 * a2 + a3);
  f->M[8] = scaleM * a3 * (a1 + a3 * a2);
}

/* _cimg_recursive_apply, order 0, Neumann boundary: CImg.h:49084-49114.
 * One line of N samples, element stride `off`, filtered in place. */
void oracle_vanvliet_line(double* data, int64_t n_samples, int64_t off,
                          const VanVliet* f) {
  const double sumsq = f->B, sum = sumsq * sumsq, a1 = f->a1, a2 = f->a2,
               a3 = f->a3;
  const double* M = f->M;
  const int64_t N = n_samples;
  const double iplus = data[(N - 1) * off];
  double v1, v2, v3;
  v1 = v2 = v3 = data[0] / sumsq;
  for (int64_t n = 0; n < N; ++n) { /* causal pass */
    double v0 = data[n * off];
    v0 += v1 * a1;
    v0 += v2 * a2;
    v0 += v3 * a3;
    data[n * off] = v0;
    v3 = v2; v2 = v1; v1 = v0;
  }
  /* Triggs boundary, CImg.h:49092-49101 */
  const double uplus = iplus / (1. - a1 - a2 - a3),
               vplus = uplus / (1. - a1 - a2 - a3), unp = v1 - uplus,
               unp1 = v2 - uplus, unp2 = v3 - uplus;
  const double w0 = (M[0] * unp + M[1] * unp1 + M[2] * unp2 + vplus) * sum;
  const double w1 = (M[3] * unp + M[4] * unp1 + M[5] * unp2 + vplus) * sum;
  const double w2 = (M[6] * unp + M[7] * unp1 + M[8] * unp2 + vplus) * sum;
  data[(N - 1) * off] = w0;
  v1 = w0; v2 = w1; v3 = w2;
  for (int64_t n = N - 2; n >= 0; --n) { /* anti-causal pass */
    double v0 = data[n * off];
    v0 *= sum;
    v0 += v1 * a1;
    v0 += v2 * a2;
    v0 += v3 * a3;
    data[n * off] = v0;
    v3 = v2; v2 = v1; v1 = v0;
  }
}

/* LookupTable::GaussianBlur -> CImg::blur(sigma, sigma, 0, true, true):
 * CorrelativeScanMatcher.h:81-87, CImg.h:49337-49351 (x pass then y pass; an
 * axis of length 1 is skipped). */
void oracle_blur(double* values, uint64_t w, uint64_t h, double sigma) {
  if (w == 0 || h == 0) return;
  VanVliet f;
  /* blur() takes float sigmas; vanvliet() skips sigma < 0.5 at order 0. */
  const float fs = (float)sigma;
  if (fs < 0.5f) return;
  oracle_vanvliet_coeffs((double)fs, &f);
  if (w > 1)
    for (uint64_t y = 0; y < h; ++y) oracle_vanvliet_line(values + y * w, (int64_t)w, 1, &f);
  if (h > 1)
    for (uint64_t x = 0; x < w; ++x) oracle_vanvliet_line(values + x, (int64_t)h, (int64_t)w, &f);
}

/* LookupTable::normalize -> CImg::normalize(0, 1): CorrelativeScanMatcher.h:79,
 * CImg.h:39499-39517. */
void oracle_normalize(double* values, uint64_t n) {
  if (n == 0) return;
  double m = values[0], M = values[0];
  for (uint64_t i = 1; i < n; ++i) {
    if (values[i] < m) m = values[i];
    if (values[i] > M) M = values[i];
  }
  if (m == M) {
    for (uint64_t i = 0; i < n; ++i) values[i] = 0.0;
    return;
  }
  if (m != 0.0 || M != 1.0)
    for (uint64_t i = 0; i < n; ++i) values[i] = (values[i] - m) / (M - m) * 1.0 + 0.0;
}

/* ---- class CorrelativeScanMatcher ------------------------------------------ */

/* GetLookupTable, CorrelativeScanMatcher.cc:21-30.  `values` is w*w, zeroed here. */
void oracle_highres_table(double range, double resolution, const float* xy,
                          int64_t n, double* values) {
  const uint64_t w = oracle_table_dim(range, resolution);
  memset(values, 0, sizeof(double) * w * w);
  for (int64_t i = 0; i < n; ++i)
    oracle_set_point_value(values, w, w, resolution, xy[2 * i], xy[2 * i + 1], 1.0);
  oracle_blur(values, w, w, 4.0); /* DEFAULT_GAUSSIAN_SIGMA, .h:19 */
  oracle_normalize(values, w * w);
}

/* GetLookupTableLowRes, CorrelativeScanMatcher.cc:32-48.  `low` is wl*wl. */
void oracle_lowres_table(double range, double low_res, const double* hi,
                         uint64_t hw, uint64_t hh, double hi_res, double* low) {
  const uint64_t wl = oracle_table_dim(range, low_res);
  memset(low, 0, sizeof(double) * wl * wl);
  for (double x = -range + CSM_EPSILON; x < range; x += low_res) {
    for (double y = -range + CSM_EPSILON; y < range; y += low_res) {
      double max_area =
          oracle_max_area(hi, hw, hh, hi_res, x, y, x + low_res, y + low_res);
      oracle_set_point_value(low, wl, wl, low_res, (float)x, (float)y, max_area);
    }
  }
}

/* RotatePointcloud, CorrelativeScanMatcher.cc:55-64: Rotation2Df narrows the
 * angle to float, evaluates sinf/cosf, and the 2x2 * 2x1 float product is
 * c*x + (-s)*y ; s*x + c*y with separate roundings (no FMA). */
void oracle_rotate(const float* xy, int64_t n, double rotation, float* out) {
  const float ang = (float)rotation;
  const float s = sinf(ang), c = cosf(ang);
  const float ms = -s;
  for (int64_t i = 0; i < n; ++i) {
    const float x = xy[2 * i], y = xy[2 * i + 1];
    const float t0 = c * x, t1 = ms * y, t2 = s * x, t3 = c * y;
    out[2 * i] = t0 + t1;
    out[2 * i + 1] = t2 + t3;
  }
}

/* CalculatePointcloudCost, CorrelativeScanMatcher.cc:66-77 */
double oracle_cost(const float* xy, int64_t n, double x_trans, double y_trans,
                   const double* values, uint64_t w, double resolution) {
  double probability = 0.0;
  const float fx = (float)x_trans, fy = (float)y_trans; /* Vector2f(x_trans, y_trans) */
  for (int64_t i = 0; i < n; ++i) {
    const float px = xy[2 * i] + fx, py = xy[2 * i + 1] + fy;
    probability += log(oracle_get_point_value(values, w, w, resolution, px, py));
  }
  return probability / (double)n;
}

typedef struct {
  double prob;
  float x, y, theta;
} TransProb;

/* MemoizeLowRes, CorrelativeScanMatcher.cc:84-115.  Returns a malloc'd memo of
 * *memo_len entries; n_theta / n_trans receive the loop trip counts. */
static TransProb* memoize_low_res(const float* a, int64_t na, const double* low,
                                  uint64_t wl, double resolution, double range,
                                  double rot_min, double rot_max, double rot_step,
                                  uint64_t* memo_len, int64_t* n_scored) {
  const uint64_t len = oracle_abs_coords(wl, wl, resolution, range, range) + 1;
  TransProb* memo = (TransProb*)malloc(sizeof(TransProb) * len);
  for (uint64_t i = 0; i < len; ++i) {
    memo[i].prob = -INFINITY;
    memo[i].x = 0; memo[i].y = 0; memo[i].theta = 0;
  }
  float* rotated = (float*)malloc(sizeof(float) * 2 * (na > 0 ? na : 1));
  for (double rotation = rot_min; rotation < rot_max; rotation += rot_step) {
    oracle_rotate(a, na, rotation, rotated);
    for (double x_trans = -range + CSM_EPSILON; x_trans < range; x_trans += resolution) {
      for (double y_trans = -range + CSM_EPSILON; y_trans < range; y_trans += resolution) {
        const double cost = oracle_cost(rotated, na, x_trans, y_trans, low, wl, resolution);
        const uint64_t slot = oracle_abs_coords(wl, wl, resolution, x_trans, y_trans);
        ++*n_scored;
        /* std::max(old, new, LessThanTransProb): keeps old unless old < new */
        if (memo[slot].prob < cost) {
          memo[slot].prob = cost;
          memo[slot].x = (float)x_trans;
          memo[slot].y = (float)y_trans;
          memo[slot].theta = (float)rotation;
        }
      }
    }
  }
  free(rotated);
  *memo_len = len;
  return memo;
}

/* GetProbAndTransformation, CorrelativeScanMatcher.cc:117-143 */
static TransProb prob_and_transformation(const float* rotated, int64_t na,
                                         const double* hi, uint64_t w,
                                         double resolution, double x_min,
                                         double x_max, double y_min, double y_max,
                                         double rotation, int64_t* n_scored) {
  TransProb best;
  best.prob = -INFINITY;
  best.x = (float)x_min; best.y = (float)y_min; best.theta = (float)rotation;
  for (double x_trans = x_min + CSM_EPSILON; x_trans < x_max; x_trans += resolution) {
    for (double y_trans = y_min + CSM_EPSILON; y_trans < y_max; y_trans += resolution) {
      const double p = oracle_cost(rotated, na, x_trans, y_trans, hi, w, resolution);
      ++*n_scored;
      if (p > best.prob) {
        best.prob = p;
        best.x = (float)x_trans; best.y = (float)y_trans; best.theta = (float)rotation;
      }
    }
  }
  return best;
}

/* Search half of GetTransformation (CorrelativeScanMatcher.cc:157-245) on
 * prebuilt tables.  rot_step is M_PI/180 in the reference (cc:92); it is a
 * parameter here only so BASELINE config 4 (0.25 degree) has a CPU checker.
 * stats[0] = low-res candidates scored, stats[1] = hi-res candidates scored,
 * stats[2] = windows popped.  Returns 0, or -1 for an empty query cloud (the
 * reference loops forever on that input). */
int oracle_search(double trans_range, double low_res, double high_res,
                  const float* a, int64_t na, const double* hi, uint64_t w,
                  const double* low, uint64_t wl, double rot_min, double rot_max,
                  double rot_step, double* score, float* x, float* y, float* theta,
                  int64_t* stats) {
  if (na <= 0) return -1;
  int64_t n_low = 0, n_hi = 0, n_pop = 0;
  uint64_t len = 0;
  TransProb* memo = memoize_low_res(a, na, low, wl, low_res, trans_range, rot_min,
                                    rot_max, rot_step, &len, &n_low);
  float* rotated = (float*)malloc(sizeof(float) * 2 * na);
  double current_probability = 1.0, best_probability = -INFINITY;
  TransProb best;
  best.prob = -INFINITY; best.x = 0; best.y = 0; best.theta = 0;
  while (current_probability >= best_probability) {
    uint64_t arg = 0; /* std::max_element: first maximum */
    for (uint64_t i = 1; i < len; ++i)
      if (memo[arg].prob < memo[i].prob) arg = i;
    const TransProb lowp = memo[arg];
    memo[arg].prob = -INFINITY;
    current_probability = lowp.prob;
    if (current_probability < best_probability) break;
    if (current_probability == -INFINITY && best_probability == -INFINITY) {
      /* nothing was ever scored (empty rotation window): the reference would
       * spin; report failure instead */
      free(memo); free(rotated);
      return -2;
    }
    oracle_rotate(a, na, (double)lowp.theta, rotated);
    const double x_min = fmax((double)lowp.x, -trans_range);
    const double x_max = fmin((double)lowp.x + low_res, trans_range);
    const double y_min = fmax((double)lowp.y, -trans_range);
    const double y_max = fmin((double)lowp.y + low_res, trans_range);
    ++n_pop;
    const TransProb hp = prob_and_transformation(rotated, na, hi, w, high_res, x_min,
                                                 x_max, y_min, y_max,
                                                 (double)lowp.theta, &n_hi);
    if (hp.prob > best_probability) {
      best_probability = hp.prob;
      best = hp;
    }
  }
  free(memo);
  free(rotated);
  *score = best_probability;
  *x = best.x; *y = best.y; *theta = best.theta;
  if (stats) { stats[0] = n_low; stats[1] = n_hi; stats[2] = n_pop; }
  return 0;
}

/* GetTransformation, 4-arg form: CorrelativeScanMatcher.cc:145-245 */
int oracle_match(double range, double trans_range, double low_res, double high_res,
                 const float* a, int64_t na, const float* b, int64_t nb,
                 double rot_min, double rot_max, double rot_step, double* score,
                 float* x, float* y, float* theta, int64_t* stats) {
  const uint64_t w = oracle_table_dim(range, high_res);
  const uint64_t wl = oracle_table_dim(range, low_res);
  double* hi = (double*)malloc(sizeof(double) * (w > 0 ? w * w : 1));
  double* low = (double*)malloc(sizeof(double) * (wl > 0 ? wl * wl : 1));
  oracle_highres_table(range, high_res, b, nb, hi);
  oracle_lowres_table(range, low_res, hi, w, w, high_res, low);
  int rc = oracle_search(trans_range, low_res, high_res, a, na, hi, w, low, wl,
                         rot_min, rot_max, rot_step, score, x, y, theta, stats);
  free(hi);
  free(low);
  return rc;
}

/* math_util::AngleMod / AngleDiff, math_util.h:79-88 and the window of the
 * 5-arg GetTransformation, CorrelativeScanMatcher.cc:247-264 */
void oracle_rotation_window(double rot_a, double rot_b, double restriction,
                            double* rot_min, double* rot_max) {
  const double two_pi = 2.0 * M_PI;
  double d = rot_a - rot_b;
  d -= two_pi * rint(d / two_pi);
  double lo = d - restriction / 2;
  lo -= two_pi * rint(lo / two_pi);
  double hi = d + restriction / 2;
  hi -= two_pi * rint(hi / two_pi);
  if (hi < lo) hi += 2 * M_PI;
  *rot_min = lo;
  *rot_max = hi;
}
