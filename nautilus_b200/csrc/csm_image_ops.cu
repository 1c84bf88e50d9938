// LookupTable::GaussianBlur and LookupTable::normalize on a CALLER-MADE table
// (CorrelativeScanMatcher.h:79-87), for the drop-in C++ header: a table a caller
// filled by hand with SetPointValue has arbitrary fp64 content, so the batch
// kernels of csm_tables.cu (bitmap input, cropped rectangle) do not apply.  These
// run the same arithmetic on a dense image, one thread per line:
//   CImg::blur(sigma, sigma, 0, true, true)  -> vanvliet 'x' then 'y'
//                                               (CImg.h:49337-49351, 49248-49314,
//                                                _cimg_recursive_apply 49062-49115)
//   CImg::normalize(0, 1)                       (CImg.h:39499-39517)
// Same operations in the same order as the reference => bit-identical results
// (tests/test_gpu_shim.py against oracle/_ref).  Not a hot path: the reference only
// calls these from GetLookupTable, which the batch kernels replace.
#include <math.h>

#include "csm_handle.h"

namespace {

// One line of _cimg_recursive_apply, order 0, Neumann boundary, in place.
__global__ void __launch_bounds__(128) blur_lines_kernel(double* __restrict__ data, int n_lines,
                                                         int n, size_t line_stride, size_t off,
                                                         VanVliet f) {
  const int line = blockIdx.x * blockDim.x + threadIdx.x;
  if (line >= n_lines) return;
  double* p = data + (size_t)line * line_stride;
  const double a1 = f.a1, a2 = f.a2, a3 = f.a3, sum = __dmul_rn(f.B, f.B);
  const double iplus = p[(size_t)(n - 1) * off];             // read before the causal pass
  double v1 = __ddiv_rn(p[0], f.B), v2 = v1, v3 = v1;        // val[1..3] = *data / sumsq
  for (int k = 0; k < n; ++k) {
    const double v0 = vv_step(p[(size_t)k * off], v1, v2, v3, a1, a2, a3);
    p[(size_t)k * off] = v0;
    v3 = v2; v2 = v1; v1 = v0;
  }
  double w0, w1, w2;
  vv_triggs(f, iplus, v1, v2, v3, w0, w1, w2);
  p[(size_t)(n - 1) * off] = w0;
  v1 = w0; v2 = w1; v3 = w2;                                 // after the shift of CImg.h:49100
  for (int k = n - 2; k >= 0; --k) {
    const double v0 = vv_step(__dmul_rn(p[(size_t)k * off], sum), v1, v2, v3, a1, a2, a3);
    p[(size_t)k * off] = v0;
    v3 = v2; v2 = v1; v1 = v0;
  }
}

__global__ void __launch_bounds__(256) minmax_kernel(const double* __restrict__ v, size_t n,
                                                     unsigned long long* __restrict__ keys) {
  double mx = -INFINITY, mn = INFINITY;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const double x = v[i];
    mx = x > mx ? x : mx;
    mn = x < mn ? x : mn;
  }
  for (int o = 16; o > 0; o >>= 1) {
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  if ((threadIdx.x & 31) == 0 && mx >= mn) {
    atomicMax(&keys[0], order_key(mx));
    atomicMin(&keys[1], order_key(mn));
  }
}

// (v - m) / (M - m) * (b - a) + a with a = 0, b = 1; fill with a when m == M; untouched
// when the range already is [0, 1] (CImg.h:39505-39516)
__global__ void __launch_bounds__(256) rescale_kernel(double* __restrict__ v, size_t n,
                                                      const unsigned long long* __restrict__ keys) {
  const double M = key_to_double(keys[0]), m = key_to_double(keys[1]);
  if (m == 0.0 && M == 1.0) return;
  const double range = __dsub_rn(M, m);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    v[i] = (m == M) ? 0.0 : __dadd_rn(__dmul_rn(__ddiv_rn(__dsub_rn(v[i], m), range), 1.0), 0.0);
}

struct DeviceImage {
  double* p = nullptr;
  unsigned long long* keys = nullptr;
  cudaStream_t st = nullptr;
  ~DeviceImage() {
    cudaFree(p); cudaFree(keys);
    if (st) cudaStreamDestroy(st);
  }
};

int open_image(DeviceImage& d, const double* values, uint64_t w, uint64_t h, int32_t device) {
  if (!values || w == 0 || h == 0 || w > 30000 || h > 30000) return fail(NAUTILUS_EINVAL, "bad table");
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0)
    return fail(NAUTILUS_ECUDA, "no CUDA device (there is no CPU fallback)");
  if (device < 0 || device >= n_dev) return fail(NAUTILUS_EINVAL, "bad device ordinal");
  CU_TRY(cudaSetDevice(device));
  CU_TRY(cudaStreamCreateWithFlags(&d.st, cudaStreamNonBlocking));
  CU_TRY(cudaMalloc(&d.p, sizeof(double) * w * h));
  CU_TRY(cudaMalloc(&d.keys, 2 * sizeof(unsigned long long)));
  CU_TRY(cudaMemcpyAsync(d.p, values, sizeof(double) * w * h, cudaMemcpyHostToDevice, d.st));
  return NAUTILUS_OK;
}

int close_image(DeviceImage& d, double* values, uint64_t w, uint64_t h) {
  CU_TRY(cudaGetLastError());
  CU_TRY(cudaMemcpyAsync(values, d.p, sizeof(double) * w * h, cudaMemcpyDeviceToHost, d.st));
  CU_TRY(cudaStreamSynchronize(d.st));
  return NAUTILUS_OK;
}

// ---- self-check of the table kernels' arithmetic helpers ------------------------------
// csm_log against log() in units in the last place, div_by against __ddiv_rn bit for bit,
// on seeded samples of the ranges the table kernels use (values in (1e-10, 1], ranges in
// [1e-3, 1e3]).
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z += 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
__global__ void __launch_bounds__(256) math_check_kernel(long long n, unsigned long long seed,
                                                         unsigned long long* __restrict__ out) {
  unsigned long long worst = 0, bad_div = 0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const unsigned long long a = mix64(seed + 2 * (unsigned long long)i), b = mix64(seed + 2 * (unsigned long long)i + 1);
    const double u = (double)(a >> 11) * 0x1p-53, v = (double)(b >> 11) * 0x1p-53;
    const double x = exp(-23.02585092994046 * u);           // log-uniform in (1e-10, 1]
    const double range = exp(13.815510557964274 * (v - 0.5));  // log-uniform in [1e-3, 1e3]
    const double want = log(x), got = csm_log(x);
    const long long wa = __double_as_longlong(want), ga = __double_as_longlong(got);
    const unsigned long long ulps = (unsigned long long)(wa > ga ? wa - ga : ga - wa);  // same sign
    worst = ulps > worst ? ulps : worst;
    const double d = __dmul_rn(x, range);
    if (__double_as_longlong(div_by(d, range, __ddiv_rn(1.0, range))) != __double_as_longlong(__ddiv_rn(d, range)))
      ++bad_div;
  }
  atomicMax(&out[0], worst);
  atomicAdd(&out[1], bad_div);
}

// ---- allocation-free normalize (no counterpart in the reference) ---------------------
// For callers that cannot cudaMalloc (graph capture, serving loops): the min/max
// scratch of normalize lives in a file-scope device array instead of a per-call
// allocation.  One (max, min) pair per block of NA_CELLS_PER_BLOCK cells, so the array
// bounds the table: NA_MAX_BLOCKS * NA_CELLS_PER_BLOCK = 67 108 864 cells (a 6000 x 6000
// solver table is 36 000 000).  256 KB + 16 B per GPU that loads the library, held
// until exit.  Every slot a launch reads was written by the same call (the pass-1 grid
// writes slots [0, blocks), pass 2 reads exactly those), so nothing has to start at
// zero and a smaller table after a larger one never sees the larger one's partials.
constexpr int NA_CELLS_PER_BLOCK = 4096;
constexpr int NA_MAX_BLOCKS = 16384;
constexpr unsigned long long NA_MAX_CELLS = (unsigned long long)NA_CELLS_PER_BLOCK * NA_MAX_BLOCKS;
// [2b] = max, [2b+1] = min of block b; the last two slots hold the table's (max, min)
static __device__ __align__(256) double na_scratch[2 * NA_MAX_BLOCKS + 2];

__global__ void __launch_bounds__(256) na_partial_kernel(const double* __restrict__ v, size_t n,
                                                         double* __restrict__ part) {
  __shared__ double s_mx[8], s_mn[8];
  const size_t base = (size_t)blockIdx.x * NA_CELLS_PER_BLOCK;
  const size_t end = base + NA_CELLS_PER_BLOCK < n ? base + NA_CELLS_PER_BLOCK : n;
  double mx = -INFINITY, mn = INFINITY;
  for (size_t i = base + threadIdx.x; i < end; i += 256) {
    const double x = v[i];
    mx = x > mx ? x : mx;
    mn = x < mn ? x : mn;
  }
  for (int o = 16; o > 0; o >>= 1) {
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  if ((threadIdx.x & 31) == 0) { s_mx[threadIdx.x >> 5] = mx; s_mn[threadIdx.x >> 5] = mn; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < 8; ++k) { mx = fmax(mx, s_mx[k]); mn = fmin(mn, s_mn[k]); }
    part[2 * blockIdx.x] = mx;
    part[2 * blockIdx.x + 1] = mn;
  }
}

// one block folds the partials of THIS call into the last two slots
__global__ void __launch_bounds__(1024) na_final_kernel(double* __restrict__ part, int blocks) {
  __shared__ double s_mx[32], s_mn[32];
  double mx = -INFINITY, mn = INFINITY;
  for (int b = threadIdx.x; b < blocks; b += 1024) {
    mx = fmax(mx, part[2 * b]);
    mn = fmin(mn, part[2 * b + 1]);
  }
  for (int o = 16; o > 0; o >>= 1) {
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  if ((threadIdx.x & 31) == 0) { s_mx[threadIdx.x >> 5] = mx; s_mn[threadIdx.x >> 5] = mn; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < 32; ++k) { mx = fmax(mx, s_mx[k]); mn = fmin(mn, s_mn[k]); }
    part[2 * NA_MAX_BLOCKS] = mx;
    part[2 * NA_MAX_BLOCKS + 1] = mn;
  }
}

// rescale_kernel's arithmetic, out of place (out may alias in)
__global__ void __launch_bounds__(256) na_rescale_kernel(const double* __restrict__ in,
                                                         double* __restrict__ out, size_t n,
                                                         const double* __restrict__ part) {
  const double M = part[2 * NA_MAX_BLOCKS], m = part[2 * NA_MAX_BLOCKS + 1];
  const bool keep = (m == 0.0 && M == 1.0);
  const double range = __dsub_rn(M, m);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const double x = in[i];
    out[i] = keep ? x
                  : (m == M) ? 0.0 : __dadd_rn(__dmul_rn(__ddiv_rn(__dsub_rn(x, m), range), 1.0), 0.0);
  }
}

// Device address of na_scratch, looked up once per device with that device current
// (in host code `na_scratch` is only the symbol's host-side stand-in).
int na_scratch_ptr(int device, double** out) {
  static std::mutex mu;
  static double* ptr[64] = {};
  if (device < 0 || device >= 64) return fail(NAUTILUS_EINVAL, "bad device ordinal");
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[device]) {
    double* p = nullptr;
    CU_TRY(cudaGetSymbolAddress((void**)&p, na_scratch));
    if (!p) return fail(NAUTILUS_ECUDA, "cudaGetSymbolAddress returned a null scratch address");
    ptr[device] = p;
  }
  *out = ptr[device];
  return NAUTILUS_OK;
}

}  // namespace

extern "C" {

int nautilus_csm_table_normalize_noalloc(const double* values_dev, double* out_dev, uint64_t width,
                                         uint64_t height, int32_t device, void* cuda_stream) {
  if (!values_dev || !out_dev || width == 0 || height == 0) return fail(NAUTILUS_EINVAL, "bad table");
  if (width > NA_MAX_CELLS || height > NA_MAX_CELLS || width * height > NA_MAX_CELLS)
    return fail(NAUTILUS_EINVAL,
                "nautilus_csm_table_normalize_noalloc: table of " + std::to_string(width) + " x " +
                    std::to_string(height) + " cells exceeds the static scratch's maximum of " +
                    std::to_string(NA_MAX_CELLS) + " cells; use nautilus_csm_table_normalize");
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0)
    return fail(NAUTILUS_ECUDA, "no CUDA device (there is no CPU fallback)");
  if (device < 0 || device >= n_dev) return fail(NAUTILUS_EINVAL, "bad device ordinal");
  CU_TRY(cudaSetDevice(device));
  double* scratch = nullptr;
  int rc = na_scratch_ptr(device, &scratch);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)cuda_stream;
  const size_t n = (size_t)width * height;
  const int blocks = (int)((n + NA_CELLS_PER_BLOCK - 1) / NA_CELLS_PER_BLOCK);
  na_partial_kernel<<<blocks, 256, 0, st>>>(values_dev, n, scratch);
  na_final_kernel<<<1, 1024, 0, st>>>(scratch, blocks);
  const int rblocks = (int)std::min<size_t>(1184, (n + 255) / 256);
  na_rescale_kernel<<<rblocks, 256, 0, st>>>(values_dev, out_dev, n, scratch);
  CU_TRY(cudaGetLastError());
  return NAUTILUS_OK;
}

int nautilus_csm_math_check(int64_t n_samples, uint64_t seed, int32_t device, int64_t* log_max_ulps,
                            int64_t* div_mismatches) {
  if (n_samples <= 0 || !log_max_ulps || !div_mismatches) return fail(NAUTILUS_EINVAL, "bad arguments");
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0)
    return fail(NAUTILUS_ECUDA, "no CUDA device (there is no CPU fallback)");
  if (device < 0 || device >= n_dev) return fail(NAUTILUS_EINVAL, "bad device ordinal");
  CU_TRY(cudaSetDevice(device));
  unsigned long long* out = nullptr;
  CU_TRY(cudaMalloc(&out, 2 * sizeof(unsigned long long)));
  cudaMemset(out, 0, 2 * sizeof(unsigned long long));
  math_check_kernel<<<1184, 256>>>((long long)n_samples, (unsigned long long)seed, out);
  unsigned long long host[2] = {0, 0};
  const cudaError_t e = cudaMemcpy(host, out, sizeof(host), cudaMemcpyDeviceToHost);
  cudaFree(out);
  if (e != cudaSuccess) return fail(NAUTILUS_ECUDA, std::string("math check: ") + cudaGetErrorString(e));
  *log_max_ulps = (int64_t)host[0];
  *div_mismatches = (int64_t)host[1];
  return NAUTILUS_OK;
}

int nautilus_csm_table_blur(double* values, uint64_t width, uint64_t height, double sigma,
                            int32_t device) {
  // blur() takes float sigmas; vanvliet() returns the image untouched for sigma < 0.5
  const float fs = (float)sigma;
  if (!(fs >= 0.5f)) {
    if (fs < 0.f) return fail(NAUTILUS_EINVAL, "negative sigma (CImg: percentage of the size) is not supported");
    return values ? NAUTILUS_OK : fail(NAUTILUS_EINVAL, "null table");
  }
  DeviceImage d;
  int rc = open_image(d, values, width, height, device);
  if (rc) return rc;
  VanVliet f;
  vanvliet_coeffs((double)fs, &f);
  const int w = (int)width, h = (int)height;
  if (w > 1) blur_lines_kernel<<<(h + 127) / 128, 128, 0, d.st>>>(d.p, h, w, (size_t)w, 1, f);
  if (h > 1) blur_lines_kernel<<<(w + 127) / 128, 128, 0, d.st>>>(d.p, w, h, 1, (size_t)w, f);
  return close_image(d, values, width, height);
}

int nautilus_csm_table_normalize(double* values, uint64_t width, uint64_t height, int32_t device) {
  DeviceImage d;
  int rc = open_image(d, values, width, height, device);
  if (rc) return rc;
  const unsigned long long init[2] = {0ULL, ~0ULL};
  CU_TRY(cudaMemcpyAsync(d.keys, init, sizeof(init), cudaMemcpyHostToDevice, d.st));
  const size_t n = (size_t)width * height;
  const int blocks = (int)std::min<size_t>(1184, (n + 255) / 256);
  minmax_kernel<<<blocks, 256, 0, d.st>>>(d.p, n, d.keys);
  rescale_kernel<<<blocks, 256, 0, d.st>>>(d.p, n, d.keys);
  return close_image(d, values, width, height);
}

}  // extern "C"
