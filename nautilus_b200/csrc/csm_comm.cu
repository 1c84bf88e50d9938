// Multi-GPU plumbing of the batch path (include/nautilus_csm.h, "multi-GPU"):
// partition of one batch over the ranks of a box, NCCL bootstrap, and the one
// collective of the path - an ncclAllGather of the 48-byte PairResult records on
// the search stream.  The reference has no distributed
// runtime (its only parallelism is OpenMP, CMakeLists.txt:16,20); pairs are
// independent (CorrelativeScanMatcher.cc:145-245), so nothing else is exchanged:
// a rank that needs a target's table builds it from the 8.6 KB scan.
//
// NCCL is bound at run time (dlopen of libnccl.so.2): a process that already holds
// an NCCL - torch's bundled one under torchrun - shares that copy instead of
// loading a second one, a plain C++ host gets the system library, and a single-GPU
// caller never touches it.
#include <dlfcn.h>
#include <math.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <vector>

#include <thread>

#include "csm_handle.h"

namespace {

// The few NCCL entry points used, with the types of nccl.h (stable since 2.0:
// ncclUniqueId is 128 bytes, ncclResult_t / ncclDataType_t are ints, ncclUint8 = 1).
struct NcclId { char internal[NAUTILUS_COMM_ID_BYTES]; };
typedef void* NcclComm;
struct NcclApi {
  int (*GetUniqueId)(NcclId*) = nullptr;
  int (*CommInitRank)(NcclComm*, int, NcclId, int) = nullptr;
  int (*CommDestroy)(NcclComm) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  int (*GetVersion)(int*) = nullptr;
  void* lib = nullptr;
  std::string error;
};
constexpr int kNcclUint8 = 1;

NcclApi* nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (api.lib) break;
    }
    if (!api.lib) { api.error = std::string("dlopen(libnccl.so.2): ") + dlerror(); return; }
    auto sym = [&](const char* name) {
      void* p = dlsym(api.lib, name);
      if (!p && api.error.empty()) api.error = std::string("libnccl lacks ") + name;
      return p;
    };
    api.GetUniqueId = reinterpret_cast<int (*)(NcclId*)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<int (*)(NcclComm*, int, NcclId, int)>(sym("ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<int (*)(NcclComm)>(sym("ncclCommDestroy"));
    api.AllGather = reinterpret_cast<int (*)(const void*, void*, size_t, int, NcclComm, cudaStream_t)>(
        sym("ncclAllGather"));
    api.GetErrorString = reinterpret_cast<const char* (*)(int)>(sym("ncclGetErrorString"));
    api.GetVersion = reinterpret_cast<int (*)(int*)>(sym("ncclGetVersion"));
  });
  return &api;
}

int nccl_fail(const char* what, int rc) {
  NcclApi* api = nccl_api();
  return fail(NAUTILUS_ECUDA, std::string(what) + ": " +
                                  (api->GetErrorString ? api->GetErrorString(rc) : "NCCL error"));
}

// Trip count of `for (r = rot_min; r < rot_max; r += step)` (cc:92-93), by the same
// accumulation when it is short, by the quotient when it is long (a weight only
// needs to be the same on every rank, which either form is).
double rotation_count(double rot_min, double rot_max, double step) {
  if (!(rot_min < rot_max) || !(step > 0)) return 0.0;
  const double est = (rot_max - rot_min) / step;
  if (est > 100000.0) return ceil(est);
  double n = 0;
  for (double r = rot_min; r < rot_max; r += step) n += 1.0;
  return n;
}

}  // namespace

extern "C" {

int nautilus_csm_partition(const int32_t* b_ids, const double* weights, int32_t n_pairs,
                           int32_t world_size, int32_t* order, int64_t* bounds) {
  if (n_pairs < 0 || world_size < 1 || !bounds || (n_pairs > 0 && (!b_ids || !order)))
    return fail(NAUTILUS_EINVAL, "bad partition arguments");
  // stable sort by target: counting sort when the ids are dense, std::stable_sort otherwise
  int32_t lo = 0, hi = -1;
  for (int32_t p = 0; p < n_pairs; ++p) {
    if (p == 0 || b_ids[p] < lo) lo = b_ids[p];
    if (p == 0 || b_ids[p] > hi) hi = b_ids[p];
  }
  const long long span = n_pairs ? (long long)hi - lo + 1 : 0;
  if (n_pairs && span <= 4LL * n_pairs + 1024) {
    std::vector<int32_t> first((size_t)span + 1, 0);
    for (int32_t p = 0; p < n_pairs; ++p) ++first[b_ids[p] - lo + 1];
    for (long long t = 0; t < span; ++t) first[t + 1] += first[t];
    for (int32_t p = 0; p < n_pairs; ++p) order[first[b_ids[p] - lo]++] = p;
  } else {
    for (int32_t p = 0; p < n_pairs; ++p) order[p] = p;
    std::stable_sort(order, order + n_pairs, [&](int32_t u, int32_t v) { return b_ids[u] < b_ids[v]; });
  }
  // cuts at the first position whose running weight reaches r / world of the total
  double total = 0.0;
  for (int32_t k = 0; k < n_pairs; ++k) total += weights ? weights[order[k]] : 1.0;
  bounds[0] = 0;
  double cum = 0.0;
  int32_t k = 0;
  for (int32_t r = 1; r < world_size; ++r) {
    const double want = total * r / world_size;
    while (k < n_pairs && !(cum + (weights ? weights[order[k]] : 1.0) >= want)) {
      cum += weights ? weights[order[k]] : 1.0;
      ++k;
    }
    bounds[r] = k;
  }
  bounds[world_size] = n_pairs;
  return NAUTILUS_OK;
}

int nautilus_csm_pair_weights(const NautilusCsmParams* p, const int32_t* a_ids,
                              const double* rot_min, const double* rot_max,
                              const int64_t* scan_offsets, int32_t n_pairs, double* weights) {
  if (!p || n_pairs < 0 || (n_pairs > 0 && (!a_ids || !rot_min || !rot_max || !weights)))
    return fail(NAUTILUS_EINVAL, "bad pair-weight arguments");
  const double step = p->rot_step > 0 ? p->rot_step : M_PI / 180;
  double n_t = 0;  // trip count of cc:96-99
  for (double t = -p->trans_range + CSM_EPSILON; t < p->trans_range; t += p->low_res) {
    n_t += 1.0;
    if (n_t > 100000) break;
  }
  double last_min = 0, last_max = 0, last_n = -1;  // batches mostly share one window
  for (int32_t k = 0; k < n_pairs; ++k) {
    if (last_n < 0 || rot_min[k] != last_min || rot_max[k] != last_max) {
      last_min = rot_min[k]; last_max = rot_max[k];
      last_n = rotation_count(last_min, last_max, step);
    }
    const double pts = scan_offsets ? (double)(scan_offsets[a_ids[k] + 1] - scan_offsets[a_ids[k]]) : 1.0;
    weights[k] = last_n * n_t * n_t * pts;
  }
  return NAUTILUS_OK;
}

int nautilus_csm_comm_unique_id(void* id_out) {
  if (!id_out) return fail(NAUTILUS_EINVAL, "null argument");
  NcclApi* api = nccl_api();
  if (!api->error.empty()) return fail(NAUTILUS_ECUDA, api->error);
  NcclId id;
  const int rc = api->GetUniqueId(&id);
  if (rc) return nccl_fail("ncclGetUniqueId", rc);
  memcpy(id_out, &id, sizeof(id));
  return NAUTILUS_OK;
}

int nautilus_csm_comm_init(NautilusCsm* h, const void* id_bytes) {
  if (!h || !id_bytes) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  if (h->world_size <= 1) return NAUTILUS_OK;  // single GPU: nothing to bootstrap
  NcclApi* api = nccl_api();
  if (!api->error.empty()) return fail(NAUTILUS_ECUDA, api->error);
  nautilus_internal_comm_release(h);
  NcclId id;
  memcpy(&id, id_bytes, sizeof(id));
  NcclComm comm = nullptr;
  const int rc = api->CommInitRank(&comm, h->world_size, id, h->world_rank);
  if (rc) return nccl_fail("ncclCommInitRank", rc);
  h->nccl_comm = comm;
  h->comm_owned = true;
  return NAUTILUS_OK;
}

int nautilus_csm_comm_attach(NautilusCsm* h, void* nccl_comm, int32_t world_rank,
                             int32_t world_size) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  if (world_size < 1 || world_rank < 0 || world_rank >= world_size || (world_size > 1 && !nccl_comm))
    return fail(NAUTILUS_EINVAL, "bad communicator");
  NcclApi* api = nccl_api();
  if (world_size > 1 && !api->error.empty()) return fail(NAUTILUS_ECUDA, api->error);
  nautilus_internal_comm_release(h);
  h->nccl_comm = world_size > 1 ? nccl_comm : nullptr;
  h->comm_owned = false;
  h->world_rank = world_rank;
  h->world_size = world_size;
  return NAUTILUS_OK;
}

}  // extern "C"

void nautilus_internal_comm_release(NautilusCsm* h) {
  if (h->nccl_comm && h->comm_owned) {
    NcclApi* api = nccl_api();
    if (api->CommDestroy) api->CommDestroy(h->nccl_comm);
  }
  h->nccl_comm = nullptr;
  h->comm_owned = false;
}

namespace {

// All-gather of the launched batch's records (this rank's block, already in
// h->results_dev, padded to the largest block) and unpacking of every rank's block
// into caller order.  Synchronises the stream.
int gather_impl(NautilusCsm* h, const int32_t* order, const int64_t* bounds, double* score,
                float* x, float* y, float* theta, NautilusCsmStats* stats, float ms_plan_extra) {
  const int world = h->world_size, rank = h->world_rank;
  float ms_wait = 0.f;
  const long long n_total = bounds[world];
  long long max_count = 0;
  for (int r = 0; r < world; ++r) {
    if (bounds[r + 1] < bounds[r]) return fail(NAUTILUS_EINVAL, "bounds must be non-decreasing");
    max_count = std::max<long long>(max_count, bounds[r + 1] - bounds[r]);
  }
  const long long n_local = bounds[rank + 1] - bounds[rank];
  if (n_local != h->pending.n_pairs)
    return fail(NAUTILUS_ESTATE, "the last match on this handle is not this rank's block");
  if (n_total == 0) return NAUTILUS_OK;
  if (world > 1 && !h->nccl_comm) return fail(NAUTILUS_ESTATE, "no communicator: call nautilus_csm_comm_init");
  int rc;
  cudaStream_t st = h->stream;
  const PairResult* src_dev = h->results_dev.p;
  if (world > 1) {
    // The all-gather sends max_count records from every rank.  A block matched through
    // nautilus_csm_match_batch reserved only its own size: grow the buffer WITHOUT losing
    // the records (DevBuf::reserve frees before it allocates).
    if (h->results_dev.cap < (size_t)max_count) {
      PairResult* bigger = nullptr;
      const size_t want = std::max<size_t>((size_t)max_count, 16);
      CU_TRY(cudaMalloc(&bigger, sizeof(PairResult) * want));
      CU_TRY(cudaMemsetAsync(bigger, 0, sizeof(PairResult) * want, st));
      if (n_local > 0)
        CU_TRY(cudaMemcpyAsync(bigger, h->results_dev.p, sizeof(PairResult) * (size_t)n_local,
                               cudaMemcpyDeviceToDevice, st));
      CU_TRY(cudaStreamSynchronize(st));  // (growth path only) before the old buffer goes
      cudaFree(h->results_dev.p);
      h->results_dev.p = bigger;
      h->results_dev.cap = want;
    }
    if ((rc = h->gather_dev.reserve((size_t)max_count * world))) return rc;
    NcclApi* api = nccl_api();
    cudaEvent_t e0, e1;
    // events 1+5w+k belong to the waves; take two beyond them
    {
      const size_t base = 1 + 5 * (size_t)std::max(1, h->pending.n_waves);
      while (h->events.size() <= base + 1) { cudaEvent_t e; cudaEventCreate(&e); h->events.push_back(e); }
      e0 = h->events[base]; e1 = h->events[base + 1];
    }
    CU_TRY(cudaEventRecord(e0, st));
    const int nrc = api->AllGather(h->results_dev.p, h->gather_dev.p,
                                   (size_t)max_count * sizeof(PairResult), kNcclUint8,
                                   h->nccl_comm, st);
    if (nrc) return nccl_fail("ncclAllGather", nrc);
    CU_TRY(cudaEventRecord(e1, st));
    src_dev = h->gather_dev.p;
    if ((rc = h->gather_pin.reserve((size_t)max_count * world))) return rc;
    CU_TRY(cudaMemcpyAsync(h->gather_pin.p, src_dev, sizeof(PairResult) * (size_t)max_count * world,
                           cudaMemcpyDeviceToHost, st));
    int* spec_counts = nullptr;
    if ((rc = nautilus_internal_queue_spec_counts(h, &spec_counts))) return rc;
    const HostTimer wait_timer;
    CU_TRY(cudaStreamSynchronize(st));
    ms_wait = wait_timer.ms();
    nautilus_internal_match_timings(h, spec_counts);
    cudaEventElapsedTime(&h->timings.ms_gather, e0, e1);
  } else {
    if ((rc = h->gather_pin.reserve((size_t)max_count))) return rc;
    CU_TRY(cudaMemcpyAsync(h->gather_pin.p, src_dev, sizeof(PairResult) * (size_t)max_count,
                           cudaMemcpyDeviceToHost, st));
    int* spec_counts = nullptr;
    if ((rc = nautilus_internal_queue_spec_counts(h, &spec_counts))) return rc;
    const HostTimer wait_timer;
    CU_TRY(cudaStreamSynchronize(st));
    ms_wait = wait_timer.ms();
    nautilus_internal_match_timings(h, spec_counts);
  }
  const HostTimer collect_timer;
  h->timings.ms_host_plan += ms_plan_extra;
  h->timings.ms_host_wait = ms_wait;
  // records -> the caller's arrays, in caller order (a scatter: every record lands in a
  // place of its own, so large gathers are split over a few host threads by rank block)
  auto scatter_blocks = [&](int r0, int r1) -> int {
    int bad_ = 0;
    for (int r = r0; r < r1; ++r) {
      const PairResult* blk = h->gather_pin.p + (size_t)r * max_count;
      for (long long k = bounds[r]; k < bounds[r + 1]; ++k) {
        const PairResult& v = blk[k - bounds[r]];
        const int p = order ? order[k] : (int)k;
        score[p] = v.score; x[p] = v.x; y[p] = v.y; theta[p] = v.theta;
        if (stats) { stats[p].n_low = v.n_low; stats[p].n_high = v.n_high; stats[p].n_pops = v.n_pops; }
        bad_ |= v.pad;
      }
    }
    return bad_;
  };
  int bad = 0;
  // (20 000 records: 0.08 ms on one thread, 0.13 on two - the threads cost more than they save
  // below a few times that)
  const int n_thr = (world >= 2 && n_total >= 60000) ? std::min(world, 4) : 1;
  if (n_thr == 1) {
    bad = scatter_blocks(0, world);
  } else {
    int bads[4] = {0, 0, 0, 0};
    std::thread th[3];
    for (int t = 1; t < n_thr; ++t)
      th[t - 1] = std::thread([&, t] { bads[t] = scatter_blocks(world * t / n_thr, world * (t + 1) / n_thr); });
    bads[0] = scatter_blocks(0, world / n_thr);
    for (int t = 1; t < n_thr; ++t) th[t - 1].join();
    bad = bads[0] | bads[1] | bads[2] | bads[3];
  }
  h->timings.ms_host_collect = collect_timer.ms();
  if (bad) return fail(NAUTILUS_EEMPTY, "a pair had nothing to score");
  return NAUTILUS_OK;
}

}  // namespace

extern "C" {

int nautilus_csm_gather(NautilusCsm* h, const int32_t* order, const int64_t* bounds, double* score,
                        float* x, float* y, float* theta, NautilusCsmStats* stats) {
  if (!h || !bounds || !score || !x || !y || !theta) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  return gather_impl(h, order, bounds, score, x, y, theta, stats, 0.f);
}

int nautilus_csm_match_batch_sharded(NautilusCsm* h, const int32_t* a_ids, const int32_t* b_ids,
                                     const double* rot_min, const double* rot_max, int32_t n_pairs,
                                     double* score, float* x, float* y, float* theta,
                                     NautilusCsmStats* stats) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  const HostTimer shard_timer;
  if (n_pairs < 0) return fail(NAUTILUS_EINVAL, "n_pairs < 0");
  if (n_pairs == 0) { h->timings = NautilusCsmTimings{}; return NAUTILUS_OK; }
  if (!a_ids || !b_ids || !rot_min || !rot_max || !score || !x || !y || !theta)
    return fail(NAUTILUS_EINVAL, "null argument");
  ScanSet& ss = h->user_scans;
  if (ss.n == 0) return fail(NAUTILUS_ESTATE, "no scans uploaded");
  for (int p = 0; p < n_pairs; ++p)
    if (a_ids[p] < 0 || a_ids[p] >= ss.n || b_ids[p] < 0 || b_ids[p] >= ss.n)
      return fail(NAUTILUS_EINVAL, "scan id out of range");
  const int world = h->world_size, rank = h->world_rank;
  // partition: the same inputs give the same cuts on every rank
  h->shard_w.resize(n_pairs);
  h->shard_order.resize(n_pairs);
  h->shard_bounds.assign(world + 1, 0);
  static_assert(sizeof(long long) == sizeof(int64_t), "offset types");
  int rc = nautilus_csm_pair_weights(&h->params, a_ids, rot_min, rot_max,
                                     reinterpret_cast<const int64_t*>(ss.host_offsets.data()),
                                     n_pairs, h->shard_w.data());
  if (rc) return rc;
  rc = nautilus_csm_partition(b_ids, h->shard_w.data(), n_pairs, world, h->shard_order.data(),
                              reinterpret_cast<int64_t*>(h->shard_bounds.data()));
  if (rc) return rc;
  const long long k0 = h->shard_bounds[rank], k1 = h->shard_bounds[rank + 1];
  const int n_local = (int)(k1 - k0);
  h->shard_a.resize(n_local); h->shard_b.resize(n_local);
  h->shard_rmin.resize(n_local); h->shard_rmax.resize(n_local);
  for (int k = 0; k < n_local; ++k) {
    const int p = h->shard_order[k0 + k];
    h->shard_a[k] = a_ids[p]; h->shard_b[k] = b_ids[p];
    h->shard_rmin[k] = rot_min[p]; h->shard_rmax[k] = rot_max[p];
  }
  long long max_count = 0;
  for (int r = 0; r < world; ++r)
    max_count = std::max<long long>(max_count, h->shard_bounds[r + 1] - h->shard_bounds[r]);
  const float ms_shard_plan = shard_timer.ms();
  rc = nautilus_internal_match_launch(h, ss, h->shard_a.data(), h->shard_b.data(),
                                      h->shard_rmin.data(), h->shard_rmax.data(), n_local,
                                      (size_t)max_count);
  // a rank whose launch failed must still enter the collective, or the others hang:
  // it contributes whatever its result buffer holds and reports its own error after
  const std::string launch_error = rc ? error_text() : std::string();
  if (rc && h->results_dev.reserve((size_t)std::max<long long>(max_count, 1))) return rc;
  const int grc = gather_impl(h, h->shard_order.data(),
                              reinterpret_cast<const int64_t*>(h->shard_bounds.data()), score, x, y,
                              theta, stats, ms_shard_plan);
  if (rc) return fail(rc, launch_error);
  return grc;
}

}  // extern "C"
