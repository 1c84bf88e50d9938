// C ABI of the nearest-neighbour correspondence path (include/nautilus_csm.h,
// "point-to-point matching").  Host work here is what the reference also does on
// the host, once per scan at load time (slam_types.h:57-75): building the
// KD-tree.  The tree is rebuilt with the reference's own rules so that the
// device can (1) mark the leaves its walk skips and (2) replay the walk for the
// rare ambiguous queries (nn_kernels.cu).  All searching runs on the device.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <thread>
#include <vector>

#include "csm_handle.h"

using namespace csm;
using namespace csmh;

namespace csmh {
void NnState::release() {
  cudaFree(pad_off); cudaFree(sorted_pts); cudaFree(skip_val); cudaFree(skip_dim);
  cudaFree(bucket_start); cudaFree(nodes); cudaFree(self_idx); cudaFree(pairs);
  cudaFree(match_idx); cudaFree(match_dist); cudaFree(counters);
  cudaFree(normals); cudaFree(corr_cnt); cudaFree(corr_off); cudaFree(corr);
  normals = nullptr; corr_cnt = nullptr; corr_off = nullptr; corr = nullptr;
  corr_pairs_cap = corr_cap = 0; corr_pairs = 0;
  pad_off = nullptr; sorted_pts = nullptr; skip_val = nullptr; skip_dim = nullptr;
  bucket_start = nullptr; nodes = nullptr; self_idx = nullptr; pairs = nullptr;
  match_idx = nullptr; match_dist = nullptr; counters = nullptr;
  pairs_cap = match_cap = 0;
  built = false;
}
}  // namespace csmh

namespace {

struct KdItem {
  float p[2];
  int idx;
};

// GetSplittingPlane (kdtree.cc:46-77): float mean, float sum of squared
// deviations, first dimension with the strictly greatest spread.
int splitting_plane(const KdItem* v, size_t n) {
  float mean[2] = {0.f, 0.f}, dev[2] = {0.f, 0.f};
  for (size_t i = 0; i < n; ++i) {
    mean[0] = mean[0] + v[i].p[0];
    mean[1] = mean[1] + v[i].p[1];
  }
  const float fn = static_cast<float>(n);
  mean[0] = mean[0] / fn;
  mean[1] = mean[1] / fn;
  for (size_t i = 0; i < n; ++i)
    for (int j = 0; j < 2; ++j) {
      const float d = v[i].p[j] - mean[j];
      const float sq = d * d;
      dev[j] = dev[j] + sq;
    }
  int plane = 0;
  float best = 0.0f;
  for (int j = 0; j < 2; ++j)
    if (dev[j] > best) { plane = j; best = dev[j]; }
  return plane;
}

// KDTree::BuildKDTree (kdtree.cc:114-150).  Nodes are numbered in pre-order.
// std::sort is libstdc++'s introsort in both builds; its sequence of moves
// depends only on the comparison results, so equal keys end up in the same
// order as in the reference's vector<KDNodeValue>.
int build_node(KdItem* v, size_t n, KdNode* nodes, int* counter, float* skip_val, int* skip_dim) {
  const int id = (*counter)++;
  const int dim = splitting_plane(v, n);
  std::sort(v, v + n, [dim](const KdItem& a, const KdItem& b) { return a.p[dim] < b.p[dim]; });
  const size_t ind = n / 2;
  nodes[id].point = v[ind].idx;
  nodes[id].dim = dim;
  nodes[id].left = -1;
  nodes[id].right = -1;
  const float split = v[ind].p[dim];
  if (ind > 0) {
    // the reference copies the left part into a new vector; sorting it in place is
    // equivalent because the parent never looks at it again
    const int l = build_node(v, ind, nodes, counter, skip_val, skip_dim);
    nodes[id].left = l;
  }
  if (ind < n - 1) {
    const int r = build_node(v + ind + 1, n - ind - 1, nodes, counter, skip_val, skip_dim);
    nodes[id].right = r;
  } else if (ind > 0) {
    // two values: a single (left, leaf) child that FindNearestPoint never visits
    // for queries on the right of this node (cc:270-293)
    const int leaf_point = nodes[nodes[id].left].point;
    skip_dim[leaf_point] = dim + 1;
    skip_val[leaf_point] = split;
  }
  return id;
}

void build_trees(const float* xy, const std::vector<long long>& offs, int s0, int s1,
                 KdNode* nodes, float* skip_val, int* skip_dim) {
  std::vector<KdItem> items;
  for (int s = s0; s < s1; ++s) {
    const long long p0 = offs[s], n = offs[s + 1] - p0;
    if (n == 0) continue;
    items.resize(n);
    for (long long i = 0; i < n; ++i) {
      items[i].p[0] = xy[2 * (p0 + i)];
      items[i].p[1] = xy[2 * (p0 + i) + 1];
      items[i].idx = (int)i;
    }
    int counter = 0;
    build_node(items.data(), (size_t)n, nodes + p0, &counter, skip_val + p0, skip_dim + p0);
  }
}

int reserve_out(NnState& nn, size_t n_pairs, size_t n_queries) {
  if (n_pairs > nn.pairs_cap) {
    cudaFree(nn.pairs); nn.pairs = nullptr; nn.pairs_cap = 0;
    CU_TRY(cudaMalloc(&nn.pairs, sizeof(NnPair) * std::max<size_t>(n_pairs, 64)));
    nn.pairs_cap = std::max<size_t>(n_pairs, 64);
  }
  if (n_queries > nn.match_cap) {
    cudaFree(nn.match_idx); cudaFree(nn.match_dist);
    nn.match_idx = nullptr; nn.match_dist = nullptr; nn.match_cap = 0;
    const size_t want = std::max<size_t>(n_queries, 1024);
    CU_TRY(cudaMalloc(&nn.match_idx, sizeof(int) * want));
    CU_TRY(cudaMalloc(&nn.match_dist, sizeof(float) * want));
    nn.match_cap = want;
  }
  return NAUTILUS_OK;
}

NnIndex make_index(const NautilusCsm* h) {
  NnIndex ix;
  ix.points = h->user_scans.points;
  ix.offsets = h->user_scans.offsets;
  ix.pad_off = h->nn.pad_off;
  ix.sorted_pts = h->nn.sorted_pts;
  ix.bucket_start = h->nn.bucket_start;
  ix.nodes = h->nn.nodes;
  ix.n_buckets = h->nn.n_buckets;
  ix.inv_cell = 1.0 / h->nn.cell;
  return ix;
}

}  // namespace

extern "C" {

void nautilus_nn_pose_transform(const double* source_pose, const double* target_pose, float* m) {
  // GetPoseAsAffine<float> (slam_util.h:20-40): Translation2d * Rotation2Dd, cast
  // to float; then target_to_world.inverse() * source_to_world (solver.cc:159)
  // with Eigen 3.3.7's Affine inverse / product rules.
  float sl[4], st[2], tl[4], tt[2];
  auto to_affine = [](const double* pose, float* l, float* t) {
    const double c = cos(pose[2]), s = sin(pose[2]);
    l[0] = (float)c; l[1] = (float)-s; l[2] = (float)s; l[3] = (float)c;  // row-major
    t[0] = (float)pose[0]; t[1] = (float)pose[1];
  };
  to_affine(source_pose, sl, st);
  to_affine(target_pose, tl, tt);
  // linear().inverse(): compute_inverse_size2 -> invdet = 1/det, adjugate * invdet
  const float det = tl[0] * tl[3] - tl[2] * tl[1];
  const float invdet = 1.0f / det;
  float il[4];
  il[0] = tl[3] * invdet;
  il[2] = -tl[2] * invdet;
  il[1] = -tl[1] * invdet;
  il[3] = tl[0] * invdet;
  // translation() = -linear_inv * t
  float it[2];
  it[0] = (-il[0]) * tt[0] + (-il[1]) * tt[1];
  it[1] = (-il[2]) * tt[0] + (-il[3]) * tt[1];
  // Affine * Affine: linear = A.l * B.l ; translation = A.l * B.t + A.t
  m[0] = il[0] * sl[0] + il[1] * sl[2];
  m[1] = il[0] * sl[1] + il[1] * sl[3];
  m[3] = il[2] * sl[0] + il[3] * sl[2];
  m[4] = il[2] * sl[1] + il[3] * sl[3];
  m[2] = (il[0] * st[0] + il[1] * st[1]) + it[0];
  m[5] = (il[2] * st[0] + il[3] * st[1]) + it[1];
}

}  // extern "C"

// Host half of the index build: the reference's trees (kdtree.cc:114-150) for every
// scan, scans spread over threads.  nautilus_nn_build_begin runs it on a library
// thread so that it overlaps whatever the caller does next on the GPU (the CSM batch
// does not touch it); nautilus_nn_build joins it, or runs it inline.
struct NnHostBuild {
  std::vector<KdNode> nodes;
  std::vector<float> skip_val;
  std::vector<int> skip_dim;
  double cell_size = 0.0;
  std::thread worker;
  bool running = false;
  void run(const float* xy_host, const std::vector<long long>* offs, int n_scans) {
    const long long total = (*offs)[n_scans];
    nodes.assign((size_t)std::max<long long>(total, 1), KdNode());
    skip_val.assign((size_t)std::max<long long>(total, 1), 0.f);
    skip_dim.assign((size_t)std::max<long long>(total, 1), 0);
    unsigned nt = std::max(1u, std::min(std::thread::hardware_concurrency(), 32u));
    nt = std::min<unsigned>(nt, (unsigned)std::max(1, n_scans));
    std::vector<std::thread> pool;
    for (unsigned t = 0; t < nt; ++t) {
      const int s0 = (int)((long long)n_scans * t / nt), s1 = (int)((long long)n_scans * (t + 1) / nt);
      pool.emplace_back(build_trees, xy_host, std::cref(*offs), s0, s1, nodes.data(),
                        skip_val.data(), skip_dim.data());
    }
    for (auto& th : pool) th.join();
  }
};

void nautilus_internal_nn_cancel(NautilusCsm* h) {
  NnHostBuild* hb = static_cast<NnHostBuild*>(h->nn_host_build);
  if (!hb) return;
  if (hb->running) hb->worker.join();
  delete hb;
  h->nn_host_build = nullptr;
}

extern "C" {

int nautilus_nn_build_begin(NautilusCsm* h, double cell_size) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  if (!(cell_size > 0)) return fail(NAUTILUS_EINVAL, "cell_size must be positive");
  ScanSet& ss = h->user_scans;
  if (ss.n == 0) return fail(NAUTILUS_ESTATE, "no scans uploaded");
  CU_TRY(cudaSetDevice(h->device));
  nautilus_internal_nn_cancel(h);
  const int rc_xy = nautilus_internal_host_xy(h);  // clouds made on the device are fetched once
  if (rc_xy) return rc_xy;
  NnHostBuild* hb = new NnHostBuild();
  hb->cell_size = cell_size;
  hb->running = true;
  // the scan set may not be replaced while this runs: upload_scans cancels (joins) first
  hb->worker = std::thread(&NnHostBuild::run, hb, ss.host_xy.data(), &ss.host_offsets, ss.n);
  h->nn_host_build = hb;
  return NAUTILUS_OK;
}

int nautilus_nn_build(NautilusCsm* h, double cell_size) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  if (!(cell_size > 0)) return fail(NAUTILUS_EINVAL, "cell_size must be positive");
  ScanSet& ss = h->user_scans;
  if (ss.n == 0) return fail(NAUTILUS_ESTATE, "no scans uploaded");
  CU_TRY(cudaSetDevice(h->device));
  NnState& nn = h->nn;
  nn.release();
  const long long total = ss.host_offsets[ss.n];
  NnHostBuild* hb = static_cast<NnHostBuild*>(h->nn_host_build);
  if (hb && hb->cell_size != cell_size) {  // a begun build for another threshold: drop it
    nautilus_internal_nn_cancel(h);
    hb = nullptr;
  }
  if (hb) {
    if (hb->running) { hb->worker.join(); hb->running = false; }
  } else {
    const int rc_xy = nautilus_internal_host_xy(h);
    if (rc_xy) return rc_xy;
    hb = new NnHostBuild();
    hb->cell_size = cell_size;
    h->nn_host_build = hb;
    hb->run(ss.host_xy.data(), &ss.host_offsets, ss.n);
  }
  std::vector<KdNode>& nodes = hb->nodes;
  std::vector<float>& skip_val = hb->skip_val;
  std::vector<int>& skip_dim = hb->skip_dim;
  int max_pts = 0;
  for (int s = 0; s < ss.n; ++s)
    max_pts = std::max<long long>(max_pts, ss.host_offsets[s + 1] - ss.host_offsets[s]);
  if (max_pts >= (1 << 28)) return fail(NAUTILUS_EINVAL, "cloud too large");
  int nb = 64;
  while (nb < 2 * max_pts && nb < 16384) nb <<= 1;
  nn.n_buckets = nb;
  nn.max_points = max_pts;
  // a hair above the threshold so that float rounding never separates two points
  // closer than the threshold by more than one cell
  nn.cell = cell_size * (1.0 + 1e-6);
  nn.pad_off_host = ss.host_offsets;  // GridPoint is 16 bytes: no padding needed

  const size_t tot = (size_t)std::max<long long>(total, 1);
  CU_TRY(cudaMalloc(&nn.pad_off, sizeof(long long) * (ss.n + 1)));
  CU_TRY(cudaMalloc(&nn.sorted_pts, sizeof(GridPoint) * tot));
  CU_TRY(cudaMalloc(&nn.skip_val, sizeof(float) * tot));
  CU_TRY(cudaMalloc(&nn.skip_dim, sizeof(int) * tot));
  CU_TRY(cudaMalloc(&nn.bucket_start, sizeof(int) * (size_t)ss.n * (nb + 4)));
  CU_TRY(cudaMalloc(&nn.nodes, sizeof(KdNode) * tot));
  CU_TRY(cudaMalloc(&nn.self_idx, sizeof(int) * tot));
  CU_TRY(cudaMalloc(&nn.counters, sizeof(unsigned long long) * 4));
  cudaStream_t st = h->stream;
  CU_TRY(cudaMemcpyAsync(nn.pad_off, nn.pad_off_host.data(), sizeof(long long) * (ss.n + 1),
                         cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(nn.skip_val, skip_val.data(), sizeof(float) * tot, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(nn.skip_dim, skip_dim.data(), sizeof(int) * tot, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(nn.nodes, nodes.data(), sizeof(KdNode) * tot, cudaMemcpyHostToDevice, st));
  launch_nn_build(ss.points, ss.offsets, nn.pad_off, nn.skip_val, nn.skip_dim, ss.n, nb,
                  1.0 / nn.cell, nn.sorted_pts, nn.bucket_start, st);
  CU_TRY(cudaGetLastError());

  // ---- GetPointNormal's lookup (solver.cc:67-77): every point against its own
  // cloud with threshold 1e-5 => the index whose stored normal the reference returns
  std::vector<NnPair> self(ss.n);
  for (int s = 0; s < ss.n; ++s) {
    self[s].src = s; self[s].tgt = s;
    const float id[6] = {1.f, 0.f, 0.f, 0.f, 1.f, 0.f};
    memcpy(self[s].m, id, sizeof(id));
    self[s].out = ss.host_offsets[s];
  }
  int rc = reserve_out(nn, ss.n, 0);
  if (rc) return rc;
  CU_TRY(cudaMemcpyAsync(nn.pairs, self.data(), sizeof(NnPair) * ss.n, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemsetAsync(nn.counters, 0, sizeof(unsigned long long) * 4, st));
  nn.built = true;
  launch_nn_match(make_index(h), nn.pairs, ss.n, nn.max_points, 0.00001f, 0.00001, nn.self_idx,
                  nullptr, nn.counters, st);
  CU_TRY(cudaGetLastError());
  CU_TRY(cudaStreamSynchronize(st));
  nautilus_internal_nn_cancel(h);  // the host arrays are on the device now
  return NAUTILUS_OK;
}

int nautilus_nn_self_index(NautilusCsm* h, int32_t* out) {
  if (!h || !out) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  if (!h->nn.built) return fail(NAUTILUS_ESTATE, "nautilus_nn_build has not run for this scan set");
  CU_TRY(cudaSetDevice(h->device));
  const long long total = h->user_scans.host_offsets[h->user_scans.n];
  CU_TRY(cudaMemcpyAsync(out, h->nn.self_idx, sizeof(int) * total, cudaMemcpyDeviceToHost, h->stream));
  CU_TRY(cudaStreamSynchronize(h->stream));
  return NAUTILUS_OK;
}

// Validates a batch, uploads its work list and launches the search; the indices
// stay in nn.match_idx (one per source point, pair p at offsets[p]).
static int nn_match_launch(NautilusCsm* h, const int32_t* src_ids, const int32_t* tgt_ids,
                           const float* src_to_tgt, double threshold, int32_t n_pairs,
                           bool want_dist, int64_t* offsets, long long* n_queries) {
  NnState& nn = h->nn;
  const ScanSet& ss = h->user_scans;
  std::vector<NnPair> pairs(n_pairs);
  long long run = 0;
  for (int p = 0; p < n_pairs; ++p) {
    if (src_ids[p] < 0 || src_ids[p] >= ss.n || tgt_ids[p] < 0 || tgt_ids[p] >= ss.n)
      return fail(NAUTILUS_EINVAL, "scan id out of range");
    pairs[p].src = src_ids[p];
    pairs[p].tgt = tgt_ids[p];
    memcpy(pairs[p].m, src_to_tgt + 6 * (size_t)p, sizeof(float) * 6);
    pairs[p].out = run;
    run += ss.host_offsets[src_ids[p] + 1] - ss.host_offsets[src_ids[p]];
    if (offsets) offsets[p + 1] = run;
  }
  int rc = reserve_out(nn, n_pairs, (size_t)run);
  if (rc) return rc;
  cudaStream_t st = h->stream;
  CU_TRY(cudaMemcpyAsync(nn.pairs, pairs.data(), sizeof(NnPair) * n_pairs, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemsetAsync(nn.counters, 0, sizeof(unsigned long long) * 4, st));
  while (h->events.size() < 2) {
    cudaEvent_t e;
    CU_TRY(cudaEventCreate(&e));
    h->events.push_back(e);
  }
  CU_TRY(cudaEventRecord(h->events[0], st));
  launch_nn_match(make_index(h), nn.pairs, n_pairs, nn.max_points, (float)threshold, threshold,
                  nn.match_idx, want_dist ? nn.match_dist : nullptr, nn.counters, st);
  CU_TRY(cudaEventRecord(h->events[1], st));
  CU_TRY(cudaGetLastError());
  *n_queries = run;
  return NAUTILUS_OK;
}

int nautilus_nn_match_batch(NautilusCsm* h, const int32_t* src_ids, const int32_t* tgt_ids,
                            const float* src_to_tgt, double threshold, int32_t n_pairs,
                            int32_t* match_idx, float* match_dist, int64_t* offsets) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  if (n_pairs < 0) return fail(NAUTILUS_EINVAL, "n_pairs < 0");
  NnState& nn = h->nn;
  if (!nn.built) return fail(NAUTILUS_ESTATE, "nautilus_nn_build has not run for this scan set");
  if (!(threshold > 0) || threshold > nn.cell)
    return fail(NAUTILUS_EINVAL, "threshold must be in (0, cell_size of nautilus_nn_build]");
  if (offsets) offsets[0] = 0;
  nn.ms_last = 0; nn.walked_last = 0;
  if (n_pairs == 0) return NAUTILUS_OK;
  if (!src_ids || !tgt_ids || !src_to_tgt || !match_idx || !offsets)
    return fail(NAUTILUS_EINVAL, "null argument");
  CU_TRY(cudaSetDevice(h->device));
  long long run = 0;
  int rc = nn_match_launch(h, src_ids, tgt_ids, src_to_tgt, threshold, n_pairs, match_dist != nullptr,
                           offsets, &run);
  if (rc) return rc;
  cudaStream_t st = h->stream;
  if (run > 0) {
    CU_TRY(cudaMemcpyAsync(match_idx, nn.match_idx, sizeof(int) * run, cudaMemcpyDeviceToHost, st));
    if (match_dist)
      CU_TRY(cudaMemcpyAsync(match_dist, nn.match_dist, sizeof(float) * run, cudaMemcpyDeviceToHost, st));
  }
  CU_TRY(cudaMemcpyAsync(&nn.walked_last, nn.counters, sizeof(unsigned long long),
                         cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  cudaEventElapsedTime(&nn.ms_last, h->events[0], h->events[1]);
  return NAUTILUS_OK;
}

int nautilus_nn_upload_normals(NautilusCsm* h, const float* normals_xy) {
  if (!h || !normals_xy) return fail(NAUTILUS_EINVAL, "null argument");
  NAUTILUS_GUARD(h);
  NnState& nn = h->nn;
  if (!nn.built) return fail(NAUTILUS_ESTATE, "nautilus_nn_build has not run for this scan set");
  const long long total = h->user_scans.host_offsets[h->user_scans.n];
  cudaFree(nn.normals);
  nn.normals = nullptr;
  CU_TRY(cudaMalloc(&nn.normals, sizeof(float2) * std::max<long long>(total, 1)));
  CU_TRY(cudaMemcpyAsync(nn.normals, normals_xy, sizeof(float2) * total, cudaMemcpyHostToDevice,
                         h->stream));
  CU_TRY(cudaStreamSynchronize(h->stream));
  return NAUTILUS_OK;
}

int nautilus_nn_match_resident(NautilusCsm* h, const int32_t* src_ids, const int32_t* tgt_ids,
                               const float* src_to_tgt, double threshold, int32_t n_pairs,
                               int64_t* corr_offsets) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  if (n_pairs < 0) return fail(NAUTILUS_EINVAL, "n_pairs < 0");
  NnState& nn = h->nn;
  if (!nn.built) return fail(NAUTILUS_ESTATE, "nautilus_nn_build has not run for this scan set");
  if (!(threshold > 0) || threshold > nn.cell)
    return fail(NAUTILUS_EINVAL, "threshold must be in (0, cell_size of nautilus_nn_build]");
  if (!corr_offsets) return fail(NAUTILUS_EINVAL, "null argument");
  corr_offsets[0] = 0;
  nn.ms_last = 0; nn.walked_last = 0; nn.corr_pairs = 0;
  if (n_pairs == 0) return NAUTILUS_OK;
  if (!src_ids || !tgt_ids || !src_to_tgt) return fail(NAUTILUS_EINVAL, "null argument");
  if (h->user_scans.host_offsets[h->user_scans.n] > 2147483647LL)
    return fail(NAUTILUS_EINVAL, "scan set too large for 32-bit correspondence indices");
  long long run = 0;
  int rc = nn_match_launch(h, src_ids, tgt_ids, src_to_tgt, threshold, n_pairs, false, nullptr, &run);
  if (rc) return rc;
  cudaStream_t st = h->stream;
  if ((size_t)n_pairs > nn.corr_pairs_cap) {
    cudaFree(nn.corr_cnt); cudaFree(nn.corr_off);
    nn.corr_cnt = nullptr; nn.corr_off = nullptr; nn.corr_pairs_cap = 0;
    const size_t want = std::max<size_t>(n_pairs, 64);
    CU_TRY(cudaMalloc(&nn.corr_cnt, sizeof(int) * want));
    CU_TRY(cudaMalloc(&nn.corr_off, sizeof(long long) * (want + 1)));
    nn.corr_pairs_cap = want;
  }
  const NnIndex ix = make_index(h);
  launch_corr_count(ix, nn.pairs, n_pairs, nn.match_idx, nn.corr_cnt, st);
  std::vector<int> cnt(n_pairs);
  CU_TRY(cudaMemcpyAsync(cnt.data(), nn.corr_cnt, sizeof(int) * n_pairs, cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaMemcpyAsync(&nn.walked_last, nn.counters, sizeof(unsigned long long),
                         cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  nn.corr_off_host.assign((size_t)n_pairs + 1, 0);
  for (int p = 0; p < n_pairs; ++p) nn.corr_off_host[p + 1] = nn.corr_off_host[p] + cnt[p];
  const long long n_corr = nn.corr_off_host[n_pairs];
  if ((size_t)n_corr > nn.corr_cap) {
    cudaFree(nn.corr);
    nn.corr = nullptr; nn.corr_cap = 0;
    const size_t want = std::max<size_t>((size_t)n_corr, 1024);
    CU_TRY(cudaMalloc(&nn.corr, sizeof(int4) * want));
    nn.corr_cap = want;
  }
  CU_TRY(cudaMemcpyAsync(nn.corr_off, nn.corr_off_host.data(), sizeof(long long) * ((size_t)n_pairs + 1),
                         cudaMemcpyHostToDevice, st));
  launch_corr_write(ix, nn.pairs, n_pairs, nn.match_idx, nn.self_idx, nn.corr_off, nn.corr, st);
  CU_TRY(cudaGetLastError());
  CU_TRY(cudaStreamSynchronize(st));
  cudaEventElapsedTime(&nn.ms_last, h->events[0], h->events[1]);
  for (int p = 0; p <= n_pairs; ++p) corr_offsets[p] = nn.corr_off_host[p];
  nn.corr_pairs = n_pairs;
  return NAUTILUS_OK;
}

int nautilus_nn_last_stats(const NautilusCsm* h, float* ms_kernel, uint64_t* n_walked) {
  if (!h) return fail(NAUTILUS_EINVAL, "null handle");
  NAUTILUS_GUARD(h);
  if (ms_kernel) *ms_kernel = h->nn.ms_last;
  if (n_walked) *n_walked = h->nn.walked_last;
  return NAUTILUS_OK;
}

}  // extern "C"
