// BASELINE configs[0] the way the reference's callers run it: T host threads share ONE
// CorrelativeScanMatcher (the drop-in header) and each calls the single-pair
// GetTransformation in a loop (bag_avg_matcher.cc:104-141, CorrelativeScanMatcher.cc:
// 381-386).  Prints single-caller latency, the concurrent rate, and how many batches
// the library's combining queue formed.  argv[1]: clouds (int32 n_pairs, then per pair
// int32 na, nb + floats), argv[2]: threads, argv[3]: calls per thread.
#include <omp.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "CorrelativeScanMatcher.h"

using Eigen::Vector2f;
typedef std::chrono::steady_clock Clock;

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  FILE* f = std::fopen(argv[1], "rb");
  int n_pairs = 0;
  if (!f || std::fread(&n_pairs, sizeof(int), 1, f) != 1) return 2;
  std::vector<std::vector<Vector2f>> A(n_pairs), B(n_pairs);
  for (int k = 0; k < n_pairs; ++k) {
    int n[2];
    if (std::fread(n, sizeof(int), 2, f) != 2) return 2;
    A[k].resize(n[0]); B[k].resize(n[1]);
    if (std::fread(A[k].data(), sizeof(Vector2f), A[k].size(), f) != A[k].size()) return 2;
    if (std::fread(B[k].data(), sizeof(Vector2f), B[k].size(), f) != B[k].size()) return 2;
  }
  std::fclose(f);
  const int threads = std::atoi(argv[2]), calls = std::atoi(argv[3]);
  CorrelativeScanMatcher matcher(30, 2, 0.3, 0.03);
  // warm-up + the serial results every concurrent call must reproduce
  std::vector<double> want(n_pairs);
  for (int k = 0; k < n_pairs; ++k) want[k] = matcher.GetTransformation(A[k], B[k], -0.5, 0.5).first;
  // single caller: latency of one call, one after the other
  const int n_serial = threads * calls < 256 ? threads * calls : 256;
  auto t0 = Clock::now();
  for (int i = 0; i < n_serial; ++i) matcher.GetTransformation(A[i % n_pairs], B[i % n_pairs], -0.5, 0.5);
  const double serial_s = std::chrono::duration<double>(Clock::now() - t0).count();
  int64_t req0 = 0, bat0 = 0, req1 = 0, bat1 = 0;
  nautilus_csm_single_queue_stats(matcher.Handle(), &req0, &bat0);
  int bad = 0;
  t0 = Clock::now();
#pragma omp parallel for num_threads(threads) schedule(static) reduction(+ : bad)
  for (int i = 0; i < threads * calls; ++i) {
    const int k = i % n_pairs;
    bad += matcher.GetTransformation(A[k], B[k], -0.5, 0.5).first != want[k];
  }
  const double conc_s = std::chrono::duration<double>(Clock::now() - t0).count();
  nautilus_csm_single_queue_stats(matcher.Handle(), &req1, &bat1);
  std::printf("{\"threads\": %d, \"calls_per_thread\": %d, \"single_caller_ms_per_pair\": %.4f, "
              "\"single_caller_pairs_per_sec\": %.1f, \"concurrent_pairs_per_sec\": %.1f, "
              "\"speedup_vs_single_caller\": %.2f, \"queue_requests\": %lld, \"queue_batches\": %lld, "
              "\"mean_batch\": %.2f, \"mismatches\": %d}\n",
              threads, calls, 1e3 * serial_s / n_serial, n_serial / serial_s,
              threads * calls / conc_s, (threads * calls / conc_s) / (n_serial / serial_s),
              (long long)(req1 - req0), (long long)(bat1 - bat0),
              (double)(req1 - req0) / (double)(bat1 - bat0 > 0 ? bat1 - bat0 : 1), bad);
  return bad ? 1 : 0;
}
