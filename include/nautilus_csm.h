/* nautilus_csm.h -- C ABI of libnautilus_csm_b200.so
 *
 * B200 (sm_100a) implementation of the correlative-scan-matching hot path of
 * ut-amrl/nautilus.  The reference has no FFI layer for this path: its
 * "operator API" is the C++ class in
 *   third_party/csm/src/CorrelativeScanMatcher.h:27-152
 * linked statically as libcsm (third_party/csm/CMakeLists.txt:32-39).  The
 * entry points below are what a binding of that class has to reach; each one
 * names the reference member it replaces.  A header-compatible C++ shim that
 * forwards the reference class to these calls lives in
 * nautilus_b200/cpp/CorrelativeScanMatcher.h, and INTEGRATION.md shows the
 * reference-side change.
 *
 * Conventions
 *   - every function returns 0 on success, a negative NAUTILUS_E* code on
 *     failure; nautilus_csm_last_error() gives the text (thread-local).
 *     There is NO CPU fallback: without a usable CUDA device create() fails.
 *   - point clouds are contiguous float (x, y) pairs, exactly the memory of a
 *     std::vector<Eigen::Vector2f>; never retained after the call returns
 *     unless uploaded with nautilus_csm_upload_scans.
 *   - "a" is the query cloud, "b" the cloud that gets rasterised; the result
 *     is the transform that moves a onto b (CorrelativeScanMatcher.cc:145-245).
 *   - tables are double, x fastest: values[x + y * width]
 *     (CImg<double>, CorrelativeScanMatcher.h:31-39).
 *   - all pointers are HOST pointers unless a name ends in _dev.
 *   - threading: the reference object is stateless and is called from OpenMP
 *     threads (CorrelativeScanMatcher.cc:381-386, bag_avg_matcher.cc:104-141).
 *     Every call that takes a handle locks it for the duration of the call and
 *     selects the handle's device for the calling thread: concurrent callers of
 *     one handle are serialised, different handles run concurrently.
 */
#ifndef NAUTILUS_CSM_H_
#define NAUTILUS_CSM_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NAUTILUS_OK 0
#define NAUTILUS_EINVAL (-1)   /* bad argument */
#define NAUTILUS_ECUDA (-2)    /* CUDA runtime error / no device */
#define NAUTILUS_ENOMEM (-3)   /* device or host allocation failed */
#define NAUTILUS_EEMPTY (-4)   /* empty query cloud or empty rotation window: the
                                  reference never terminates on these inputs
                                  (CorrelativeScanMatcher.cc:170-181) */
#define NAUTILUS_ESTATE (-5)   /* call order violated (e.g. match before upload) */

typedef struct NautilusCsm NautilusCsm;

/* Constructor arguments of CorrelativeScanMatcher (CorrelativeScanMatcher.h:114-119)
 * plus the knobs the reference hard-codes. */
typedef struct NautilusCsmParams {
  double scanner_range; /* range_       : table half extent, metres            */
  double trans_range;   /* trans_range_ : translation search half extent       */
  double low_res;       /* low_res_     : coarse cell size                     */
  double high_res;      /* high_res_    : fine cell size                       */
  double rot_step;      /* 0 => M_PI/180, the reference's constant (cc:92)     */
  double blur_sigma;    /* 0 => 4, DEFAULT_GAUSSIAN_SIGMA (h:19)               */
  int32_t device;       /* CUDA device ordinal                                 */
  int32_t margin_cells; /* table crop margin in hi-res cells around the scan's
                           occupied box.  <0 => never crop (every table is
                           computed over the full grid, bit-identical to the
                           reference); 0 => default (128: cells above the lookup
                           floor differ from the uncropped table by at most
                           5e-14 relative, scripts/margin_study.py)            */
  int32_t max_tables;   /* resident hi-res table slots per wave; 0 => auto     */
  int32_t stage_cells;  /* shared-memory budget of the low-res scorer's staged table,
                           in doubles; 0 => default                             */
  int32_t world_rank;   /* multi-GPU (one process per GPU): this process's rank ...  */
  int32_t world_size;   /* ... of world_size; 0 or 1 => single GPU                 */
} NautilusCsmParams;

/* Per-pair search statistics, counted by the same rule as the oracle:
 * low-res candidates scored, hi-res candidates scored, windows popped. */
typedef struct NautilusCsmStats {
  int64_t n_low;
  int64_t n_high;
  int64_t n_pops;
} NautilusCsmStats;

/* Device-side timings of the last batch call, CUDA events on the handle's stream. */
typedef struct NautilusCsmTimings {
  float ms_total;    /* first launch to last launch of the call                */
  float ms_tables;   /* rasterise + blur + normalise + pool                    */
  float ms_low;      /* low-res exhaustive scoring                             */
  float ms_refine;   /* memo reduction + best-first hi-res refinement          */
  int32_t launches;  /* kernels launched by the call                           */
  int32_t waves;     /* table waves the call was split into                    */
  int32_t spec_windows; /* hi-res windows scored speculatively (K6 queue length) */
  float ms_gather;   /* NCCL all-gather of the result records (sharded calls)  */
  float ms_host_plan;   /* host wall time from call entry to the last kernel launch */
  float ms_host_collect; /* host wall time from the stream synchronise to return  */
  float ms_host_wait;   /* host wall time blocked in the stream synchronise       */
} NautilusCsmTimings;

const char* nautilus_csm_last_error(void);
/* Library build tag, e.g. "nautilus_csm_b200 sm_100a". */
const char* nautilus_csm_version(void);

/* CorrelativeScanMatcher::CorrelativeScanMatcher (h:114-119). */
int nautilus_csm_create(const NautilusCsmParams* params, NautilusCsm** out);
void nautilus_csm_destroy(NautilusCsm* h);

/* Run the handle's work on a caller-owned cudaStream_t (e.g. torch's current
 * stream) instead of its own.  NULL restores the handle's stream. */
int nautilus_csm_set_stream(NautilusCsm* h, void* cuda_stream);

/* LookupTable::width of the hi- and low-res tables (h:32-35). */
int nautilus_csm_table_dims(const NautilusCsm* h, uint64_t* w_high, uint64_t* w_low);

/* LookupTable::convertX / AbsCoords (h:43-49,104-109) for a table of the given
 * resolution; exposed so host code and tests can check the index maps. */
int64_t nautilus_csm_cell(double resolution, uint64_t dim, float v);
/* The same map as the device evaluates it (float thresholds, no division);
 * returns -1 outside [0, dim).  Host twin of the kernels' cell_fast, for tests. */
int64_t nautilus_csm_cell_fast(double resolution, uint64_t dim, float v);

/* Test hook for the kernels' "regular cells" shortcut (host code, no launches): the
 * search kernels derive the cells of consecutive translations from the first one
 * when its float quotient lies clear of a cell edge (margins proven at create time
 * from the actual translation lists).  Replays that rule on n_samples random and
 * cell-edge-hugging coordinates against the exact cell map for the low-res grid
 * (MemoizeLowRes, cc:96-99) and every hi-res window (cc:126-129); writes how many
 * predictions were made and how many disagreed (must be 0). */
int nautilus_csm_regular_check(NautilusCsm* h, int64_t n_samples, uint64_t seed,
                               int64_t* n_checked, int64_t* n_bad_low, int64_t* n_bad_high);

/* --- scan residency -------------------------------------------------------
 * Copies n_scans clouds to the device (one H2D); scan ids are 0..n_scans-1 and
 * replace any previous set.  offsets has n_scans+1 entries, in points. */
int nautilus_csm_upload_scans(NautilusCsm* h, const float* xy, const int64_t* offsets,
                              int32_t n_scans);

/* --- tables ---------------------------------------------------------------
 * GetLookupTableHighRes (cc:50-53 -> GetLookupTable cc:21-30): rasterise,
 * Van Vliet blur, normalise.  out_high: w_high^2 doubles; out_low (optional):
 * w_low^2 doubles = GetLookupTableLowRes of the same table. */
int nautilus_csm_highres_table(NautilusCsm* h, const float* xy, int64_t n_points,
                               double* out_high, double* out_low);

/* GetLookupTableLowRes(const LookupTable&) (cc:32-48) on a caller-made table of
 * dimensions width x height and cell size `resolution`. out_low: w_low^2. */
int nautilus_csm_lowres_from_table(NautilusCsm* h, const double* high, uint64_t width,
                                   uint64_t height, double resolution, double* out_low);

/* LookupTable::GaussianBlur(sigma) (h:81-87 -> CImg::blur(sigma, sigma, 0, true, true),
 * CImg.h:49337-49351) and LookupTable::normalize() (h:79 -> CImg::normalize(0, 1),
 * CImg.h:39499-39517) on a caller-made table of width x height doubles, in place, on
 * CUDA device `device`.  Same operations in the same order as CImg: bit-identical.
 * For the drop-in header's LookupTable members; the batch path never calls them. */
int nautilus_csm_table_blur(double* values, uint64_t width, uint64_t height, double sigma,
                            int32_t device);
int nautilus_csm_table_normalize(double* values, uint64_t width, uint64_t height, int32_t device);

/* Self-check of the two arithmetic helpers of the table kernels (csm_common.cuh), like
 * nautilus_csm_regular_check for the cell rule: the table-driven logarithm the kernels
 * take of every cell above the lookup floor (cc:72 takes log() per lookup) against
 * CUDA's log(), worst distance in units in the last place, and the reciprocal-based
 * division of the normalisation (CImg.h:39515) against the IEEE division, number of
 * quotients that differ, over n_samples seeded samples of the tables' value range. */
int nautilus_csm_math_check(int64_t n_samples, uint64_t seed, int32_t device, int64_t* log_max_ulps,
                            int64_t* div_mismatches);

/* No counterpart in the reference: added for
 * callers that cannot allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  nautilus_csm_table_normalize's arithmetic on
 * DEVICE buffers: values_dev -> out_dev (width x height doubles each, out_dev may equal
 * values_dev), launched on `cuda_stream` (NULL = the default stream) of CUDA device
 * `device`; asynchronous, the caller synchronises the stream.  Calls no cudaMalloc /
 * cudaFree / stream creation: the min/max scratch is a static __device__ array
 * (256 KB per GPU that loads the library, held until exit), which bounds the table at
 * NAUTILUS_NOALLOC_MAX_CELLS cells; larger tables return NAUTILUS_EINVAL with a message
 * naming the limit.  Every call on a device shares that array: calls that overlap on
 * the GPU (different streams, no ordering between them) race; calls ordered on one
 * stream, or by events, are safe, and nothing in the array has to be cleared between
 * them (each call writes every slot it later reads). */
#define NAUTILUS_NOALLOC_MAX_CELLS 67108864ULL
int nautilus_csm_table_normalize_noalloc(const double* values_dev, double* out_dev, uint64_t width,
                                         uint64_t height, int32_t device, void* cuda_stream);

/* Builds (and discards) the tables of the given uploaded scans; the batch
 * form of the two calls above, used for the table-build configuration.
 * checksum (optional) receives the sum over all low-res tables. */
int nautilus_csm_build_tables(NautilusCsm* h, const int32_t* scan_ids, int32_t n,
                              double* checksum);

/* --- matching -------------------------------------------------------------
 * GetTransformation(a, b, rotation_min, rotation_max) (cc:145-245) for n_pairs
 * independent pairs of uploaded scans.  Outputs have n_pairs entries; stats is
 * optional.  Pairs may come in any order; they are grouped by target internally
 * so each target's tables are built once per call. */
int nautilus_csm_match_batch(NautilusCsm* h, const int32_t* a_ids, const int32_t* b_ids,
                             const double* rot_min, const double* rot_max,
                             int32_t n_pairs, double* score, float* x, float* y,
                             float* theta, NautilusCsmStats* stats);

/* Device-side copy of the last batch's results for a collective that follows on
 * the same stream (multi-GPU gather): n_pairs records of 48 bytes in caller
 * order, each { double score; float x, y, theta; int32 status; int64 n_low,
 * n_high, n_pops; }.  dst_dev is a DEVICE pointer; the copy is stream-ordered. */
int nautilus_csm_copy_results_dev(NautilusCsm* h, void* dst_dev, int32_t n_pairs);

/* --- multi-GPU: one batch partitioned over the GPUs of a box ---------------------
 * The reference has no distributed runtime; pairs are independent
 * (CorrelativeScanMatcher.cc:145-245 keeps no cross-pair state), so a batch is cut
 * into world_size blocks, each rank matches its block, and the only exchange is one
 * ncclAllGather of the 48-byte result records on the search stream.  One process per
 * GPU; libnccl.so.2 is bound at run time (dlopen), so a process that already holds
 * an NCCL (e.g. torch's) shares it and a single-GPU caller needs none.
 *
 * Partition (pure host code, no handle, no GPU): pairs sorted by target scan id
 * (stable), cut into world_size contiguous blocks of near-equal total weight, so the
 * pairs that share a lookup table land on the rank that builds it; a target whose
 * pairs straddle a cut is rasterised on both sides (8.6 KB of points against a 32 MB
 * table).  order: permutation of 0..n_pairs-1; rank r owns
 * order[bounds[r] .. bounds[r+1]).  weights == NULL => equal weights. */
int nautilus_csm_partition(const int32_t* b_ids, const double* weights, int32_t n_pairs,
                           int32_t world_size, int32_t* order, int64_t* bounds);
/* Low-res work of every pair, N_theta * N_t^2 * N (rotation steps of cc:92-93,
 * translations of cc:96-99, points of the query scan); scan_offsets = the offsets
 * array of nautilus_csm_upload_scans. */
int nautilus_csm_pair_weights(const NautilusCsmParams* params, const int32_t* a_ids,
                              const double* rot_min, const double* rot_max,
                              const int64_t* scan_offsets, int32_t n_pairs, double* weights);

/* Bootstrap.  Rank 0 calls comm_unique_id and hands the 128 bytes to every rank by
 * whatever channel the host program has (the Nautilus node would use its launcher's
 * environment or a ROS parameter; bench.py broadcasts it); every rank then calls
 * comm_init with the handle created with its world_rank / world_size.  Collective. */
#define NAUTILUS_COMM_ID_BYTES 128
int nautilus_csm_comm_unique_id(void* id_out);
int nautilus_csm_comm_init(NautilusCsm* h, const void* id);
/* Alternative: use a caller-owned ncclComm_t of the same world (not destroyed by the
 * library). */
int nautilus_csm_comm_attach(NautilusCsm* h, void* nccl_comm, int32_t world_rank,
                             int32_t world_size);

/* After this rank matched ITS block with nautilus_csm_match_batch (pairs
 * order[bounds[rank] .. bounds[rank+1]) in that order): all-gathers every rank's
 * records and writes all n_pairs = bounds[world_size] results in caller order.
 * Collective; every rank receives everything. */
int nautilus_csm_gather(NautilusCsm* h, const int32_t* order, const int64_t* bounds,
                        double* score, float* x, float* y, float* theta,
                        NautilusCsmStats* stats);

/* nautilus_csm_match_batch for one batch split over the world: every rank passes the
 * SAME arguments (the whole batch; the scan set resident on every rank), the library
 * partitions, matches this rank's block, gathers, and every rank returns all results.
 * With world_size 1 it is nautilus_csm_match_batch. */
int nautilus_csm_match_batch_sharded(NautilusCsm* h, const int32_t* a_ids, const int32_t* b_ids,
                                     const double* rot_min, const double* rot_max,
                                     int32_t n_pairs, double* score, float* x, float* y,
                                     float* theta, NautilusCsmStats* stats);

/* Single pair from host clouds (upload + tables + search).  Thread-safe and
 * COMBINING: the reference object is called from OpenMP threads
 * (bag_avg_matcher.cc:104-141, cc:381-386) and one pair cannot fill a B200, so calls
 * that arrive while a batch is in flight are queued and matched together as the next
 * batch by whichever caller is leading; every caller gets exactly the result a lone
 * call returns.  Clouds passed by the same pointer and size by several queued calls
 * are uploaded (and their tables built) once. */
int nautilus_csm_match(NautilusCsm* h, const float* a, int64_t na, const float* b,
                       int64_t nb, double rot_min, double rot_max, double* score,
                       float* x, float* y, float* theta, NautilusCsmStats* stats);

/* How many nautilus_csm_match calls the handle has served and in how many batches. */
int nautilus_csm_single_queue_stats(const NautilusCsm* h, int64_t* n_requests, int64_t* n_batches);

/* GetUncertaintyMatrix: dims == 3 -> the 3x3 form over all rotations
 * (cc:267-319, `rotation` ignored); dims == 2 -> the 2x2 form at `rotation`
 * (cc:322-373).  out: dims*dims floats, row-major.  Float moment sums in the
 * reference's loop order; expect agreement to ~1e-4 relative (exp / log differ
 * from glibc's by ulps), not bit parity. */
int nautilus_csm_uncertainty(NautilusCsm* h, const float* a, int64_t na, const float* b,
                             int64_t nb, int32_t dims, double rotation, float* out);

/* The 2x2 form for a batch of pairs of uploaded scans, pair p at rotation[p]: what
 * GetLocalUncertaintyStats (cc:375-398) asks for after its GetTransformation calls,
 * for all comparison clouds at once - and for all base clouds of a bag at once, the way
 * bag_avg_matcher.cc:104-141 loops.  Tables are built once per distinct target.
 * out: 4 floats per pair, row-major. */
int nautilus_csm_uncertainty_batch(NautilusCsm* h, const int32_t* a_ids, const int32_t* b_ids,
                                   const double* rotation, int32_t n_pairs, float* out);

/* The device part of GetLocalUncertaintyStats (cc:375-398) for host clouds: every
 * comparison cloud is matched against the target cloud over the full rotation
 * (GetTransformation(targetCloud, comparisonCloud), cc:384-385) and the 2x2 matrix is
 * taken at the matched rotation (cc:386-387), all comparison clouds in two batched
 * passes.  comp_offsets: n_comp + 1 entries, in points.  u_out: 4 floats per
 * comparison cloud; theta_out (optional): the matched rotations.  The eigenvalue
 * averages of cc:388-396 are left to the caller (the C++ shim uses Eigen like the
 * reference).  Does not touch the uploaded scan set. */
int nautilus_csm_local_uncertainty(NautilusCsm* h, const float* target_xy, int64_t n_target,
                                   const float* comp_xy, const int64_t* comp_offsets,
                                   int32_t n_comp, float* u_out, float* theta_out);

/* The rotation window of the 5-argument GetTransformation (cc:247-264). */
void nautilus_csm_rotation_window(double rot_a, double rot_b, double restriction,
                                  double* rot_min, double* rot_max);

int nautilus_csm_last_timings(const NautilusCsm* h, NautilusCsmTimings* out);

/* --- front end ----------------------------------------------------------------
 * Range -> point conversion and keyframe scoring on the device, so that a bag's
 * scans cross PCIe once as ranges (4 B per beam):
 *   nautilus::LaserScanToPointCloud            (src/input/pointcloud_helpers.cc:28-48)   variant 0
 *   pointcloud_helpers::LaserScanToPointCloud  (third_party/csm/src/pointcloud_helpers.cc:44-66,
 *                                               with FilterGlancing :21-42)              variant 1
 *   ComputeScatterMatrixScore / GetLCCandidates (src/loop_closure/lc_candidate_filter.cc:23-81) */
typedef struct NautilusLaserSpec {
  float angle_min, angle_max, angle_increment, range_min; /* sensor_msgs/LaserScan fields */
  double max_range;       /* the functions' max_range argument                      */
  int32_t variant;        /* 0: nautilus (angle = min + inc * index); 1: csm
                             (angle accumulated in float, glancing filter)           */
  int32_t truncate_ends;  /* csm variant: drop pi/12 at either end                   */
} NautilusLaserSpec;

/* Converts n_scans scans of n_beams ranges each (row-major) and makes the
 * resulting clouds the handle's scan set, exactly as if they had been passed to
 * nautilus_csm_upload_scans (scan ids 0..n_scans-1). */
int nautilus_csm_upload_ranges(NautilusCsm* h, const float* ranges, int32_t n_scans,
                               int32_t n_beams, const NautilusLaserSpec* spec);
/* Device time of the kernels of the last nautilus_csm_upload_ranges (CUDA events:
 * count pass + offsets + write pass). */
int nautilus_csm_frontend_last_ms(const NautilusCsm* h, float* kernel_ms);
/* Reads the resident scan set back: offsets[n_scans + 1] and, when xy is not NULL,
 * the points (cap_points >= offsets[n_scans]). */
int nautilus_csm_download_scans(NautilusCsm* h, float* xy, int64_t cap_points, int64_t* offsets);
/* ComputeScatterMatrixScore of every resident scan (min / max eigenvalue of the
 * scatter matrix; float sums in point order like the reference). */
int nautilus_lc_scatter_scores(NautilusCsm* h, double* score);
/* GetLCCandidates: sequential keyframe selection over the scores (host; the
 * reference's constants are min_dist = 5, min_score = 0.70).  pose_xy = n x 2
 * floats (GetPoseTranslation).  out has room for n indices. */
int nautilus_lc_candidates(const double* score, const float* pose_xy, int32_t n, double min_dist,
                           double min_score, int64_t* out, int32_t* n_out);

/* --- residual blocks ------------------------------------------------------------
 * The step after the point-to-point matching: the lidar residual functors that
 * Solver::AddLidarResiduals hands to Ceres (src/optimization/solver.cc:275-296),
 *   LIDARPointResidual  (src/optimization/slam_residuals.h:123-145)  kind 0
 *   LIDARNormalResidual (src/optimization/slam_residuals.h:64-90)    kind 1
 * evaluated for a batch of residual blocks with analytic Jacobians -- the values
 * ceres::AutoDiffCostFunction<..., DYNAMIC, 3, 3> computes by automatic
 * differentiation.  Block b owns correspondences [offsets[b], offsets[b+1]) of the
 * four parallel float (x, y) arrays (PointCorrespondences, data_structures.h:62-100)
 * and the poses source_poses[3b..], target_poses[3b..] = (x, y, theta).
 * residuals: 2 doubles per correspondence.  jac_source / jac_target (optional):
 * per correspondence 2 rows x 3 columns, row-major, i.e. a block's slice is Ceres'
 * jacobians[0] / jacobians[1] of that block. */
int nautilus_residuals_eval(NautilusCsm* h, int32_t kind, const double* source_poses,
                            const double* target_poses, const int64_t* offsets, int32_t n_blocks,
                            const float* source_points, const float* target_points,
                            const float* source_normals, const float* target_normals,
                            double* residuals, double* jac_source, double* jac_target);
/* Device time of the last nautilus_residuals_eval kernel (CUDA events). */
int nautilus_residuals_last_ms(const NautilusCsm* h, float* kernel_ms);

/* --- point-to-point matching -------------------------------------------------
 * The nearest-neighbour correspondence search that feeds the Ceres residuals:
 *   Solver::GetPointToPointMatching   (src/optimization/solver.cc:132-172)
 *   FindClosestPoint / GetPointNormal (src/optimization/solver.cc:67-89)
 *   KDTree<float,2>::FindNearestPoint (src/util/kdtree.cc:253-305)
 * over the clouds uploaded with nautilus_csm_upload_scans.  Results are the
 * reference KD-tree's, index for index (see nn_kernels.cu for why that holds). */

/* Builds the per-scan search structures (hashed grid on the device; the
 * reference's KD-tree layout, kdtree.cc:114-150, on the host -- the reference
 * builds its trees on the host at load time too, slam_types.h:57-75).
 * cell_size = the largest threshold later calls will use, e.g.
 * outlier_threshold (config/default_config.lua:66). */
int nautilus_nn_build(NautilusCsm* h, double cell_size);
/* Starts the host half of nautilus_nn_build (the KD-tree layout, the slow part: 0.4 s
 * for 20 000 scans) on library threads and returns at once, so that it overlaps what
 * the caller does next on the GPU - typically the CSM batch, which does not need the
 * index.  A later nautilus_nn_build with the same cell_size joins it and finishes on
 * the device.  Replacing the scan set cancels it. */
int nautilus_nn_build_begin(NautilusCsm* h, double cell_size);

/* target_to_world.inverse() * source_to_world as a row-major 2x3 float matrix
 * from two (x, y, theta) poses: GetPoseAsAffine<float> (src/util/slam_util.h:
 * 20-40) and solver.cc:159.  A caller that has Eigen should build the matrix
 * with Eigen itself, as the reference does. */
void nautilus_nn_pose_transform(const double* source_pose, const double* target_pose,
                                float* m_2x3);

/* For every pair p and source point i: match_idx[offsets[p] + i] = index in the
 * target cloud of the point FindClosestPoint returns for (m_p * source_i), or -1
 * when it returns nullopt (distance >= threshold, compared in double as
 * solver.cc:85 does).  src_to_tgt: 6 floats per pair.  offsets: n_pairs + 1
 * entries, filled by the call (prefix sums of the source cloud sizes).
 * match_dist (optional): the distance FindNearestPoint returned where matched. */
int nautilus_nn_match_batch(NautilusCsm* h, const int32_t* src_ids, const int32_t* tgt_ids,
                            const float* src_to_tgt, double threshold, int32_t n_pairs,
                            int32_t* match_idx, float* match_dist, int64_t* offsets);

/* GetPointNormal (solver.cc:67-77): for every uploaded point, the index (within
 * its scan) of the tree node the reference would take the normal from, i.e. the
 * point's nearest neighbour in its own cloud within 1e-5 (itself, or the first
 * visited duplicate).  out: one entry per uploaded point, scan by scan. */
int nautilus_nn_self_index(NautilusCsm* h, int32_t* out);

/* ---- the matching -> residual chain without a host round trip ---------------
 * Solver::AddLidarResiduals (src/optimization/solver.cc:297-318) calls
 * GetPointToPointMatching for a pair and hands the PointCorrespondences to a
 * LIDAR*Residual cost function, which Ceres then evaluates many times while the
 * poses move.  The three calls below keep the correspondences on the device
 * between those steps: only counts come back from the search, and every
 * evaluation moves poses in, residuals / Jacobians out. */

/* Normals of the uploaded scan set, one (x, y) per point in upload order: the
 * KDNodeValue::normal of every tree node, which the reference fills at load time
 * (src/util/kdtree.cc:29, NormalComputation::GetNormals; slam_types.h:57-75). */
int nautilus_nn_upload_normals(NautilusCsm* h, const float* normals_xy);

/* nautilus_nn_match_batch whose result stays on the device as the resident
 * correspondence set: for every pair the matched source points in source order
 * (the order of AddCorrespondence, data_structures.h:89-99), each with its target
 * point and the two normals GetPointNormal would return (solver.cc:67-77).
 * corr_offsets: n_pairs + 1 entries, filled by the call (correspondences per
 * pair, prefix sums).  A later call replaces the set. */
int nautilus_nn_match_resident(NautilusCsm* h, const int32_t* src_ids, const int32_t* tgt_ids,
                               const float* src_to_tgt, double threshold, int32_t n_pairs,
                               int64_t* corr_offsets);

/* nautilus_residuals_eval over the resident correspondence set: one residual
 * block per pair of the last nautilus_nn_match_resident, source_poses /
 * target_poses = 3 doubles (x, y, theta) per pair.  Output layout as
 * nautilus_residuals_eval; jac_* may be null.  Repeatable with new poses. */
int nautilus_residuals_eval_resident(NautilusCsm* h, int32_t kind, const double* source_poses,
                                     const double* target_poses, double* residuals,
                                     double* jac_source, double* jac_target);

/* Page-locked host memory for large result arrays (residuals / Jacobians of
 * nautilus_residuals_eval*, match indices): device -> host copies into it run at
 * PCIe speed instead of through the driver's staging of pageable memory.  Any
 * host pointer argument of this library may point into such a buffer. */
int nautilus_host_alloc(uint64_t bytes, void** out);
int nautilus_host_free(void* p);

/* Kernel time of the last nautilus_nn_match_batch (CUDA events) and how many of
 * its queries took the exact tree walk instead of the grid. */
int nautilus_nn_last_stats(const NautilusCsm* h, float* ms_kernel, uint64_t* n_walked);

#ifdef __cplusplus
}
#endif
#endif /* NAUTILUS_CSM_H_ */
