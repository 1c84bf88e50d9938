// Host-side launch interface of the CSM kernels (implemented in csm_tables.cu
// and csm_search.cu, driven by csm_capi.cu).
#pragma once
#include "csm_common.cuh"

namespace csm {

// 1-D pooling window list of GetLookupTableLowRes (CorrelativeScanMatcher.cc:37-45),
// device arrays of n entries: hi-res cell window [s, e] (inclusive, after
// get_crop's swap, may reach outside the grid), destination low-res cell (or
// -1 when SetPointValue would drop it) and "last writer of its destination".
struct PoolListDev {
  int* s;
  int* e;
  int* dst;
  unsigned char* win;
  int n;
  int two_run_ok;  // windows tile the axis, every one >= 8 cells, n <= 1024: ypass2_kernel applies
};

void launch_table_wave(const Geometry& g, const VanVliet& f, const float2* points,
                       const long long* offsets, const int* scan_ids, TableSlot* slots,
                       TableMeta* metas, const PoolListDev& pool, int n_tables, bool store_log,
                       cudaStream_t stream, int* launches, bool has_ck = false);
// Must run once over freshly allocated slots: the raster kernel only clears what the
// slot's previous table wrote into the low-res arrays.
void launch_slot_init(const Geometry& g, const TableSlot* slots, TableMeta* metas, int n_slots,
                      cudaStream_t stream);
// grids the checkpointed (and therefore the two-run) table kernels cover
bool table_two_run_supported(int w);
void launch_pool_generic(const double* hi, int w, int h, const PoolListDev& px,
                         const PoolListDev& py, double* low, int wl, cudaStream_t stream);
void launch_expand(const Geometry& g, const TableSlot* slots, const TableMeta* metas, int t,
                   double* dense, cudaStream_t stream);
void launch_low_checksum(const Geometry& g, const TableSlot* slots, int n_tables, double* out,
                         cudaStream_t stream);

// ---- search ----------------------------------------------------------------

// Search-grid description shared by every pair of a batch; all lists are built
// on the host with the reference's accumulating-double loops.
struct SearchGrid {
  int n_t;               // low-res translations per axis (cc:96-99 trip count)
  const float* t_low;    // [n_t] (float)x_trans of the low-res grid
  const int* slot_rank;  // [n_t*n_t] index i*n_t+j -> rank of AbsCoords(x_i,y_j) among slots
  int n_h_max;           // max hi-res steps per window axis
  const int* n_h;        // [n_t] hi-res steps of the window opened at low index i
  const float* t_high;   // [n_t*n_h_max] (float)x_trans of the hi-res steps (cc:126-129)
  // Low-res scorer shortcut: if the float estimate e of (p + t_low[k0]) / low_res
  // has |e| < reg_qmax and a fractional part in (reg_eps, 1 - reg_eps), then the
  // cells of t_low[k0 + k], k < 16, are floor(e) + k (proved on the host from the
  // actual t_low values and float error bounds, build_search_grid).
  float reg_eps, reg_qmax;
  // The same for the hi-res windows (refine): cells of t_high[i*n_h_max + k] are
  // floor(e) + k for the estimate e of (p + t_high[i*n_h_max]) / high_res.
  float reg_eps_h, reg_qmax_h;
};

// Per-pair work item.
struct PairWork {
  int a_scan;      // query scan id
  int table;       // table slot (in the current wave) of the target scan
  int rot_begin;   // first entry of this pair's rotation list
  int n_rot;       // rotations in [rot_min, rot_max) (cc:92 trip count)
  int out_index;   // position in the caller's output arrays
  int pad;
};

struct PairResult {
  double score;
  float x, y, theta;
  int pad;
  long long n_low, n_high, n_pops;
};

// Rotation list entry: angle as stored ((float)rotation), cosf/sinf from libm.
struct Rotation {
  float theta, c, s, pad;
};

// Speculative refinement (csm_search.cu, K6): state of a pair after its first
// pop, one queue entry per hi-res window that the reference may still open, and
// the window's result.
// Listed windows per pair: every slot but the first pop when the translation grid has
// at most RF_SPEC_CAP_MAX + 1 slots (the reference's 14 x 14 grid: 195), else the pair
// finishes serially beyond the cap.
#ifndef RF_SPEC_CAP_MAX_N
#define RF_SPEC_CAP_MAX_N 255
#endif
constexpr int RF_SPEC_CAP_MAX = RF_SPEC_CAP_MAX_N;
inline int spec_cap_for(int n_tr) { return n_tr - 1 < 1 ? 1 : (n_tr - 1 < RF_SPEC_CAP_MAX ? n_tr - 1 : RF_SPEC_CAP_MAX); }
struct SpecPair {
  double best_prob;   // best1: result of the first pop
  float x, y, theta;
  int status;         // 1: nothing to score
  int n_list;         // windows listed (<= SearchBuffers::spec_cap)
  int n_more;         // slots with bound >= best1 beyond the list
  int item_base;      // first queue entry of the pair
  int first_tr;       // slot of the first pop (-1: none)
  long long n_high;   // candidates of the first pop
};
struct SpecItem {
  int pair, tr, rot, pad;  // pair (wave-relative), slot i*n_t+j, rotation index within the pair
  double memo;             // the slot's low-res bound
};
struct SpecOut {
  double score;
  int idx, pad;
};

struct SearchBuffers {
  double* chunk_best;   // [n_pairs][n_chunks_max][n_t*n_t] best score per theta chunk
  int* chunk_rot;       // same shape: rotation index (within the pair) of that best
  int chunks_per_pair;  // theta chunks per pair (uniform upper bound)
  int rot_per_chunk;
  SpecPair* spec_pairs;  // [n_pairs]
  SpecItem* spec_items;  // [n_pairs * spec_cap]
  SpecOut* spec_out;     // [n_pairs * spec_cap]
  int spec_cap;          // listed windows per pair (spec_cap_for(n_t * n_t))
  int* spec_counters;    // [0] queue entries written, [1] next entry to fetch
};

void launch_low_score(const Geometry& g, const SearchGrid& sg, const CellMap& cm_low,
                      const float2* points, const long long* offsets, const TableSlot* slots,
                      const TableMeta* metas, const PairWork* pairs, const Rotation* rots,
                      int n_pairs, const SearchBuffers& sb, int table_cells, cudaStream_t stream);
// Returns the number of kernels launched.  serial: one block per pair runs the
// whole best-first loop; otherwise first pop / speculative windows / replay.
int launch_refine(const Geometry& g, const SearchGrid& sg, const CellMap& cm_high,
                  const float2* points, const long long* offsets, const TableSlot* slots,
                  const TableMeta* metas, const PairWork* pairs, const Rotation* rots,
                  int n_pairs, const SearchBuffers& sb, PairResult* results, bool serial,
                  int sm_count, cudaStream_t stream);
size_t low_score_smem_bytes(int n_t, int wl, int table_cells);
int low_score_warps();  // warps of a low_score_kernel block = rotations it scores per round
// GetUncertaintyMatrix (cc:267-373) after launch_low_score filled sb for the pair:
// costs[n_rot][n_t*n_t] scratch, out = dims x dims floats (row-major).
void launch_uncertainty(const SearchGrid& sg, const CellMap& cm_high, const float2* points,
                        const long long* offsets, const TableSlot* slots, const TableMeta* metas,
                        const PairWork& pw, const Rotation* rots, const SearchBuffers& sb,
                        double useless_threshold, double* costs, int dims, float* out,
                        cudaStream_t stream);

// The 2x2 form for a batch of pairs, one rotation each (pairs[k].rot_begin), after
// launch_low_score filled sb with chunks_per_pair chunks per pair: costs[n_pairs][n_t*n_t]
// scratch, out = 4 floats per pair at pairs[k].out_index.
void launch_uncertainty_batch(const SearchGrid& sg, const CellMap& cm_high, const float2* points,
                              const long long* offsets, const TableSlot* slots,
                              const TableMeta* metas, const PairWork* pairs, const Rotation* rots,
                              int n_pairs, const SearchBuffers& sb, double useless_threshold,
                              double* costs, float* out, cudaStream_t stream);

}  // namespace csm
