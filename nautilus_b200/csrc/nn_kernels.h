// Host-side launch interface of the nearest-neighbour correspondence kernels
// (nn_kernels.cu), driven by nn_capi.cu.  Replaces the search side of
//   Solver::GetPointToPointMatching     (src/optimization/solver.cc:132-172)
//   FindClosestPoint / GetPointNormal   (src/optimization/solver.cc:67-89)
//   KDTree<float,2>::FindNearestPoint   (src/util/kdtree.cc:253-305)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

namespace csm {

// One node of a scan's KD-tree exactly as KDTree::BuildKDTree lays it out
// (kdtree.cc:114-150); children are node indices local to the scan, -1 = none.
// Only the rare ambiguous queries walk it (see nn_kernels.cu).
struct KdNode {
  int point;  // index of value_ in the scan
  int left, right;
  int dim;    // splitting_dimension_
};

// A target point in bucket order.  idx_dim = index within the scan | code << 30,
// code 0: plain point; 1 / 2: the point is the only child of a KD-tree node
// that splits on x / y at skip_val, and the walk skips it for queries whose
// coordinate on that axis is greater than skip_val (see nn_kernels.cu).
struct __align__(16) GridPoint {
  float x, y, skip_val;
  int idx_dim;
};

// Device view of the per-scan search structures of a scan set.
struct NnIndex {
  const float2* points;        // the scan set's clouds (caller order)
  const long long* offsets;    // [n_scans+1] first point of each scan
  const long long* pad_off;    // [n_scans+1] first entry of each scan in sorted_pts
  const GridPoint* sorted_pts; // points of a scan ordered by hash bucket
  const int* bucket_start;     // [n_scans][n_buckets + 4] exclusive prefix sums
  const KdNode* nodes;         // [total points] KD-tree nodes, root = first node of a scan
  int n_buckets;               // power of two
  double inv_cell;             // 1 / grid cell size (cell size >= every threshold used)
};

// Work item of a batch: source cloud, target cloud, source->target transform
// (row-major 2x3: [l00 l01 t0; l10 l11 t1]) and first output slot.
struct NnPair {
  int src, tgt;
  float m[6];
  long long out;
};

void launch_nn_build(const float2* points, const long long* offsets, const long long* pad_off,
                     const float* skip_val, const int* skip_dim, int n_scans, int n_buckets,
                     double inv_cell, GridPoint* sorted_pts, int* bucket_start,
                     cudaStream_t stream);
// match_idx[out + i] = index in the target cloud of the nearest neighbour of
// source point i, or -1; match_dist (optional) = its distance.  counters[0]
// counts queries that took the exact KD-tree walk.
void launch_nn_match(const NnIndex& ix, const NnPair* pairs, int n_pairs, int max_tgt_points,
                     float thr_f, double thr_d, int* match_idx, float* match_dist,
                     unsigned long long* counters, cudaStream_t stream);
size_t nn_match_smem_bytes(int n_buckets, int max_points);

// Resident correspondence set (AddCorrespondence order, data_structures.h:89-99):
// per matched source point {source point, target point, source normal, target
// normal} as indices into the scan set's point / normal arrays.
// count: cnt[p] = matches of pair p.  write: corr[corr_off[p] ..] in source order.
void launch_corr_count(const NnIndex& ix, const NnPair* pairs, int n_pairs, const int* match_idx,
                       int* cnt, cudaStream_t stream);
void launch_corr_write(const NnIndex& ix, const NnPair* pairs, int n_pairs, const int* match_idx,
                       const int* self_idx, const long long* corr_off, int4* corr,
                       cudaStream_t stream);

}  // namespace csm

namespace csmh {

// NN state of a handle: built by nautilus_nn_build for the uploaded scan set.
struct NnState {
  bool built = false;
  double cell = 0;
  int n_buckets = 0;
  int max_points = 0;
  std::vector<long long> pad_off_host;
  long long* pad_off = nullptr;
  csm::GridPoint* sorted_pts = nullptr;
  float* skip_val = nullptr;
  int* skip_dim = nullptr;
  int* bucket_start = nullptr;
  csm::KdNode* nodes = nullptr;
  int* self_idx = nullptr;          // canonical index of every point (GetPointNormal)
  csm::NnPair* pairs = nullptr;
  size_t pairs_cap = 0;
  int* match_idx = nullptr;
  float* match_dist = nullptr;
  size_t match_cap = 0;
  unsigned long long* counters = nullptr;
  // resident chain: normals of the scan set, compacted correspondences of the last
  // nautilus_nn_match_resident
  float2* normals = nullptr;
  int* corr_cnt = nullptr;
  long long* corr_off = nullptr;
  size_t corr_pairs_cap = 0;
  int4* corr = nullptr;
  size_t corr_cap = 0;
  int corr_pairs = 0;                       // pairs of the resident set, 0 = none
  std::vector<long long> corr_off_host;
  float ms_last = 0;
  unsigned long long walked_last = 0;
  void release();
};

}  // namespace csmh
