// Microbenchmark: what HBM sustains for the access pattern of the table y-pass.
// A "table" is a rectangle of ROWS x COLS doubles with a row pitch of PITCH doubles inside
// a slot of SLOT doubles (the tables of a wave lie 32 MB apart).  A thread owns VEC
// adjacent columns and walks the rows top to bottom, 16 rows of loads in flight per chunk,
// like ypass_kernel: a warp's load is a granule of 256 * VEC bytes and consecutive granules
// of a warp lie one row pitch apart.  Prints GB/s for VEC = 1, 2, 4 and for a fully
// coalesced stream over the same bytes.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a scripts/micro/granule_bench.cu -o scripts/micro/granule_bench
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ROWS = 800, COLS = 640, PITCH = 640;
constexpr size_t SLOT = 4000000;  // doubles between tables (2000 x 2000 grid)

template <int VEC>
__global__ void __launch_bounds__(128) column_walk(const double* __restrict__ base, double* out) {
  const int table = blockIdx.y;
  const int col = (blockIdx.x * 128 + threadIdx.x) * VEC;
  if (col >= COLS) return;
  const double* p = base + (size_t)table * SLOT + col;
  double acc = 0.0;
  for (int r0 = 0; r0 < ROWS; r0 += 16) {
    double v[16][VEC];
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      const double* q = p + (size_t)(r0 + u) * PITCH;
      if (VEC == 1) v[u][0] = __ldg(q);
      if (VEC == 2) { const double2 t = __ldg(reinterpret_cast<const double2*>(q)); v[u][0] = t.x; v[u][VEC - 1] = t.y; }
      if (VEC == 4) {
        const double2 t0 = __ldg(reinterpret_cast<const double2*>(q)), t1 = __ldg(reinterpret_cast<const double2*>(q) + 1);
        v[u][0] = t0.x; v[u][1 % VEC] = t0.y; v[u][2 % VEC] = t1.x; v[u][3 % VEC] = t1.y;
      }
    }
#pragma unroll
    for (int u = 0; u < 16; ++u)
#pragma unroll
      for (int k = 0; k < VEC; ++k) acc += v[u][k];
  }
  if (acc == 12345.678) out[0] = acc;  // keep the loads
}

// The same walk with the y-pass's work per chunk: after a chunk's 16 loads, 16 steps of the
// dependent recursion v0 = ((x + v1 a1) + v2 a2) + v3 a3 (separately rounded, like vv_step).
// Dynamic shared memory only limits the resident blocks per SM.
template <int SWEEPS>
__global__ void __launch_bounds__(128) column_walk_chain(const double* __restrict__ base, double* out,
                                                         double a1, double a2, double a3) {
  extern __shared__ double pad[];
  const int table = blockIdx.y;
  const int col = blockIdx.x * 128 + threadIdx.x;
  if (col >= COLS) return;
  const double* p = base + (size_t)table * SLOT + col;
  double v1 = 0.0, v2 = 0.0, v3 = 0.0;
  for (int r0 = 0; r0 < ROWS; r0 += 16) {
    double v[16];
#pragma unroll
    for (int u = 0; u < 16; ++u) v[u] = __ldg(p + (size_t)(r0 + u) * PITCH);
    // SWEEPS passes over the chunk, each feeding the next (like causal rebuild -> anti-causal)
#pragma unroll
    for (int s = 0; s < SWEEPS; ++s) {
#pragma unroll
      for (int u = 0; u < 16; ++u) {
        double v0 = __dadd_rn(v[u], __dmul_rn(v1, a1));
        v0 = __dadd_rn(v0, __dmul_rn(v2, a2));
        v0 = __dadd_rn(v0, __dmul_rn(v3, a3));
        v3 = v2; v2 = v1; v1 = v0;
        v[u] = v0;
      }
    }
  }
  if (v1 == 12345.678) out[0] = v1 + pad[0];
}

__global__ void __launch_bounds__(256) stream_read(const double* __restrict__ base, double* out, int n_tables) {
  double acc = 0.0;
  const size_t per = (size_t)ROWS * PITCH;
  for (int t = blockIdx.y; t < n_tables; t += gridDim.y) {
    const double* p = base + (size_t)t * SLOT;
    for (size_t i = blockIdx.x * 256 + threadIdx.x; i < per; i += (size_t)gridDim.x * 256) acc += __ldg(p + i);
  }
  if (acc == 12345.678) out[0] = acc;
}

template <int VEC>
float time_walk(const double* d, double* out, int n_tables) {
  dim3 grid((COLS / VEC + 127) / 128, n_tables);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  column_walk<VEC><<<grid, 128>>>(d, out);
  cudaEventRecord(e0);
  for (int i = 0; i < 3; ++i) column_walk<VEC><<<grid, 128>>>(d, out);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  return ms / 3;
}

int main() {
  const int n_tables = 2000;
  double *d = nullptr, *out = nullptr;
  if (cudaMalloc(&d, sizeof(double) * SLOT * n_tables) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMalloc(&out, 64);
  cudaMemset(d, 0, sizeof(double) * SLOT * n_tables);
  const double gb = (double)n_tables * ROWS * COLS * 8 / 1e9;
  printf("bytes per pass: %.2f GB (%d tables of %d x %d doubles, pitch %d, slots %.0f MB apart)\n", gb,
         n_tables, ROWS, COLS, PITCH, SLOT * 8 / 1e6);
  printf("column walk, 256 B granules (1 double per lane): %.0f GB/s\n", gb / (time_walk<1>(d, out, n_tables) * 1e-3));
  printf("column walk, 512 B granules (2 doubles per lane): %.0f GB/s\n", gb / (time_walk<2>(d, out, n_tables) * 1e-3));
  printf("column walk, 1 KB granules (4 doubles per lane): %.0f GB/s\n", gb / (time_walk<4>(d, out, n_tables) * 1e-3));
  auto chain = [&](auto kernel, int sweeps) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int blocks : {4, 6, 8, 12}) {
      const size_t smem = (size_t)(227 * 1024 / blocks - 1024) & ~(size_t)255;
      dim3 grid((COLS + 127) / 128, n_tables);
      cudaEvent_t a, b;
      cudaEventCreate(&a); cudaEventCreate(&b);
      kernel<<<grid, 128, smem>>>(d, out, 0.5, 0.25, 0.125);
      cudaEventRecord(a);
      for (int i = 0; i < 3; ++i) kernel<<<grid, 128, smem>>>(d, out, 0.5, 0.25, 0.125);
      cudaEventRecord(b);
      cudaEventSynchronize(b);
      float t = 0.f;
      cudaEventElapsedTime(&t, a, b);
      printf("column walk + %d x 16-step fp64 recursion per chunk, %2d blocks of 128 per SM: %.0f GB/s, %.3f ms per pass, %.2f G steps/ms\n",
             sweeps, blocks, gb / (t / 3 * 1e-3), t / 3, gb / 8 * sweeps / (t / 3));
    }
  };
  chain(column_walk_chain<1>, 1);
  chain(column_walk_chain<2>, 2);
  chain(column_walk_chain<3>, 3);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  stream_read<<<dim3(64, 148), 256>>>(d, out, n_tables);
  cudaEventRecord(e0);
  for (int i = 0; i < 3; ++i) stream_read<<<dim3(64, 148), 256>>>(d, out, n_tables);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  printf("coalesced stream over the same bytes: %.0f GB/s\n", gb / (ms / 3 * 1e-3));
  printf("last error: %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
