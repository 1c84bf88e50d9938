// Microbenchmark: DADD dependent-chain latency and pipe throughput on sm_100a.
// chains = independent accumulators per thread, warps = warps per block (1 block per SM).
#include <cstdio>
#include <cuda_runtime.h>
template <int CH>
__global__ void k(double* out, int iters, double t) {
  double acc[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c) acc[c] = threadIdx.x + c;
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int c = 0; c < CH; ++c) acc[c] = __dadd_rn(acc[c], t);
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int c = 0; c < CH; ++c) s += acc[c];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (double)(t1 - t0);
}
template <int CH>
void run(int warps) {
  double* d; cudaMalloc(&d, 8 * 148 * 1024);
  int iters = 2000;
  k<CH><<<148, warps * 32>>>(d, iters, 1.5);
  cudaDeviceSynchronize();
  k<CH><<<148, warps * 32>>>(d, iters, 1.5);
  double h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  double per_group = h / (iters * 8.0);  // cycles per round of CH DADDs (per warp)
  double warp_instr_per_clk_sm = warps * CH / per_group;
  printf("chains %2d warps/SM %2d : %.1f clk per round, %.2f DADD warp-instr/clk/SM\n", CH, warps, per_group, warp_instr_per_clk_sm);
  cudaFree(d);
}
int main() {
  for (int w : {1, 4, 8, 16, 24, 32}) { run<1>(w); run<2>(w); run<4>(w); run<7>(w); run<14>(w); }
  return 0;
}
