"""Host-side mirror of the reference's CSM interface over the C ABI.

Names, argument meaning and error behaviour follow
/root/reference/third_party/csm/src/CorrelativeScanMatcher.h:27-152 so that the
parity tests read like the reference's own tests
(third_party/csm/test/correlative_scan_matcher_tests.cc).  All compute goes
through libnautilus_csm_b200.so; nothing here falls back to the CPU.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import _lib
from ._lib import CsmError, Params, Stats, Timings, check, ptr

MIN_VALUE_FOR_LOOKUP = 1e-10  # CorrelativeScanMatcher.h:20


def _cloud(a) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a.reshape(-1, 2)


class LookupTable:
    """struct LookupTable (CorrelativeScanMatcher.h:27-110): a host copy of a table
    with the reference's world<->cell maps.  Index math is the library's
    (nautilus_csm_cell), values are a numpy array [height, width] (x fastest)."""

    def __init__(self, range_: float, resolution: float, values: np.ndarray | None = None):
        self.resolution = float(resolution)
        dim = int(math.floor((range_ * 2.0) / resolution))  # h:32-35
        self.width = self.height = dim
        if values is None:
            values = np.zeros((dim, dim), dtype=np.float64)
        self.values = np.ascontiguousarray(values, dtype=np.float64).reshape(self.height, self.width)

    def convertX(self, x) -> int:
        return int(_lib.lib().nautilus_csm_cell(self.resolution, self.width, np.float32(x)))

    def convertY(self, y) -> int:
        return int(_lib.lib().nautilus_csm_cell(self.resolution, self.height, np.float32(y)))

    def IsInside(self, point) -> bool:
        x, y = self.convertX(point[0]), self.convertY(point[1])
        return 0 <= x < self.width and 0 <= y < self.height

    def GetPointValue(self, point) -> float:
        x, y = self.convertX(point[0]), self.convertY(point[1])
        if not (0 <= x < self.width and 0 <= y < self.height) or \
                self.values[y, x] <= MIN_VALUE_FOR_LOOKUP:
            return MIN_VALUE_FOR_LOOKUP
        return float(self.values[y, x])

    def SetPointValue(self, point, value: float) -> None:
        x, y = self.convertX(point[0]), self.convertY(point[1])
        if 0 <= x < self.width and 0 <= y < self.height:
            self.values[y, x] = value

    def GetDebugImage(self) -> np.ndarray:
        return self.values

    def AbsCoords(self, x: float, y: float) -> int:
        row = int((self.height // 2 + _round_half_away(y / self.resolution)) * self.width)
        col = int(self.width // 2 + _round_half_away(x / self.resolution))
        return row + col


def _round_half_away(v: float) -> float:
    return math.copysign(math.floor(abs(v) + 0.5), v)  # C round()


class CorrelativeScanMatcher:
    """class CorrelativeScanMatcher (CorrelativeScanMatcher.h:112-152) on one B200.

    Extra keyword arguments expose what the reference hard-codes: rot_step
    (M_PI/180, cc:92), device, margin_cells (see include/nautilus_csm.h)."""

    def __init__(self, scanner_range: float, trans_range: float, low_res: float,
                 high_res: float, *, rot_step: float = 0.0, device: int = 0,
                 margin_cells: int = 0, max_tables: int = 0, blur_sigma: float = 0.0,
                 stage_cells: int = 0, world_rank: int = 0, world_size: int = 1):
        self._L = _lib.lib()
        self.range_, self.trans_range_ = float(scanner_range), float(trans_range)
        self.low_res_, self.high_res_ = float(low_res), float(high_res)
        p = Params(scanner_range, trans_range, low_res, high_res, rot_step, blur_sigma,
                   device, margin_cells, max_tables, stage_cells, world_rank, world_size)
        self._params = p
        self.world_rank, self.world_size = (world_rank, world_size) if world_size > 1 else (0, 1)
        h = C.c_void_p()
        check(self._L.nautilus_csm_create(C.byref(p), C.byref(h)))
        self._h = h
        wh, wl = C.c_uint64(), C.c_uint64()
        check(self._L.nautilus_csm_table_dims(self._h, C.byref(wh), C.byref(wl)))
        self.w_high, self.w_low = int(wh.value), int(wl.value)
        self.n_scans = 0
        self._scan_sizes = np.zeros(0, np.int64)

    def close(self):
        if getattr(self, "_h", None):
            self._L.nautilus_csm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream_ptr: int | None):
        check(self._L.nautilus_csm_set_stream(self._h, C.c_void_p(cuda_stream_ptr or 0)))

    # ---- tables ------------------------------------------------------------
    def GetLookupTableHighRes(self, pointcloud) -> LookupTable:
        pc = _cloud(pointcloud)
        out = np.empty((self.w_high, self.w_high), dtype=np.float64)
        check(self._L.nautilus_csm_highres_table(self._h, ptr(pc), len(pc), ptr(out), None))
        return LookupTable(self.range_, self.high_res_, out)

    def GetLookupTables(self, pointcloud) -> tuple[LookupTable, LookupTable]:
        """Hi- and low-res tables of one cloud in one device pass."""
        pc = _cloud(pointcloud)
        hi = np.empty((self.w_high, self.w_high), dtype=np.float64)
        lo = np.empty((self.w_low, self.w_low), dtype=np.float64)
        check(self._L.nautilus_csm_highres_table(self._h, ptr(pc), len(pc), ptr(hi), ptr(lo)))
        return (LookupTable(self.range_, self.high_res_, hi),
                LookupTable(self.range_, self.low_res_, lo))

    def GetLookupTableLowRes(self, high_res_table: LookupTable) -> LookupTable:
        hv = np.ascontiguousarray(high_res_table.values, dtype=np.float64)
        out = np.empty((self.w_low, self.w_low), dtype=np.float64)
        check(self._L.nautilus_csm_lowres_from_table(
            self._h, ptr(hv), high_res_table.width, high_res_table.height,
            high_res_table.resolution, ptr(out)))
        return LookupTable(self.range_, self.low_res_, out)

    # ---- matching ----------------------------------------------------------
    def GetTransformation(self, pointcloud_a, pointcloud_b, rotation_min: float = 0.0,
                          rotation_max: float = 2 * math.pi, *, with_stats: bool = False):
        """-> (prob, ((x, y), theta)) like the reference's pair<double, pair<Vector2f, float>>."""
        a, b = _cloud(pointcloud_a), _cloud(pointcloud_b)
        s, x, y, t = C.c_double(), C.c_float(), C.c_float(), C.c_float()
        st = Stats()
        check(self._L.nautilus_csm_match(self._h, ptr(a), len(a), ptr(b), len(b), rotation_min,
                                         rotation_max, C.byref(s), C.byref(x), C.byref(y),
                                         C.byref(t), C.byref(st)))
        r = (s.value, ((np.float32(x.value), np.float32(y.value)), np.float32(t.value)))
        return (r, (st.n_low, st.n_high, st.n_pops)) if with_stats else r

    def GetTransformationRestricted(self, pointcloud_a, pointcloud_b, rotation_a: float,
                                    rotation_b: float, rotation_restriction: float, **kw):
        """The 5-argument overload (cc:247-264)."""
        lo, hi = self.rotation_window(rotation_a, rotation_b, rotation_restriction)
        return self.GetTransformation(pointcloud_a, pointcloud_b, lo, hi, **kw)

    def rotation_window(self, rotation_a, rotation_b, rotation_restriction):
        lo, hi = C.c_double(), C.c_double()
        self._L.nautilus_csm_rotation_window(rotation_a, rotation_b, rotation_restriction,
                                             C.byref(lo), C.byref(hi))
        return lo.value, hi.value

    def regular_check(self, n_samples: int = 200000, seed: int = 1):
        """Test hook: replay the kernels' regular-cell rule on the host against the
        exact cell map; returns (predictions checked, low-res misses, hi-res misses)."""
        n, bl, bh = C.c_int64(), C.c_int64(), C.c_int64()
        check(self._L.nautilus_csm_regular_check(self._h, n_samples, seed, C.byref(n), C.byref(bl),
                                                 C.byref(bh)))
        return n.value, bl.value, bh.value

    def GetUncertaintyMatrix(self, pointcloud_a, pointcloud_b, rotation: float | None = None):
        """Olson covariance: 3x3 over all rotations (cc:267-319) or, with `rotation`,
        2x2 over translations (cc:322-373).  float32 matrix like the Eigen result."""
        a, b = _cloud(pointcloud_a), _cloud(pointcloud_b)
        dims = 3 if rotation is None else 2
        out = np.zeros(dims * dims, np.float32)
        check(self._L.nautilus_csm_uncertainty(self._h, ptr(a), len(a), ptr(b), len(b), dims,
                                               0.0 if rotation is None else float(rotation),
                                               ptr(out)))
        return out.reshape(dims, dims)

    def uncertainty_batch(self, a_ids, b_ids, rotation) -> np.ndarray:
        """2x2 GetUncertaintyMatrix of many pairs of uploaded scans, pair p at rotation[p]
        (nautilus_csm_uncertainty_batch).  -> float32 [n, 2, 2]."""
        a_ids = np.ascontiguousarray(a_ids, dtype=np.int32)
        b_ids = np.ascontiguousarray(b_ids, dtype=np.int32)
        rot = np.ascontiguousarray(rotation, dtype=np.float64)
        out = np.zeros((len(a_ids), 2, 2), np.float32)
        check(self._L.nautilus_csm_uncertainty_batch(self._h, ptr(a_ids), ptr(b_ids), ptr(rot),
                                                     len(a_ids), ptr(out)))
        return out

    def GetLocalUncertaintyStats(self, comparison_clouds, target_cloud):
        """cc:375-398: mean condition number and mean largest eigenvalue of the 2x2
        covariance of target_cloud matched against every comparison cloud.  Two batch
        calls (all matches, then all covariances) instead of the reference's loop."""
        tgt = _cloud(target_cloud)
        comps = [_cloud(c) for c in comparison_clouds]
        n = len(comps)
        offs = np.zeros(n + 1, np.int64)
        offs[1:] = np.cumsum([len(c) for c in comps])
        cxy = np.ascontiguousarray(np.concatenate(comps)) if n else np.zeros((0, 2), np.float32)
        us = np.zeros((n, 2, 2), np.float32)
        check(self._L.nautilus_csm_local_uncertainty(self._h, ptr(tgt), len(tgt), ptr(cxy), ptr(offs),
                                                     n, ptr(us), None))
        cond, scale = 0.0, 0.0
        for u in us.astype(np.float64):
            tr, det = u[0, 0] + u[1, 1], u[0, 0] * u[1, 1] - u[0, 1] * u[1, 0]
            root = math.sqrt(max(tr * tr / 4.0 - det, 0.0))
            ev = np.asarray([tr / 2.0 + root, tr / 2.0 - root], np.float32)
            cond += float(ev.max() / ev.min())
            scale += float(ev.max())
        return cond / n, scale / n

    # ---- batch API (new; what a batched SolveAutoLC would call) -------------
    def upload_scans(self, xy, offsets) -> None:
        xy = _cloud(xy)
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        check(self._L.nautilus_csm_upload_scans(self._h, ptr(xy), ptr(offsets), len(offsets) - 1))
        self.n_scans = len(offsets) - 1
        self._scan_sizes = (offsets[1:] - offsets[:-1]).astype(np.int64)

    # ---- front end (include/nautilus_csm.h, "front end") -----------------------
    def upload_ranges(self, ranges, angle_min: float, angle_max: float, angle_increment: float,
                      range_min: float, max_range: float, *, variant: int = 0,
                      truncate_ends: bool = False) -> None:
        """LaserScanToPointCloud for a batch of scans (rows of `ranges`) on the device;
        the clouds become the resident scan set.  variant 0:
        src/input/pointcloud_helpers.cc:28-48, variant 1 (csm, glancing filter):
        third_party/csm/src/pointcloud_helpers.cc:21-66."""
        r = np.ascontiguousarray(ranges, dtype=np.float32)
        if r.ndim != 2:
            raise ValueError("ranges must be [n_scans, n_beams]")
        spec = _lib.LaserSpec(angle_min, angle_max, angle_increment, range_min, max_range,
                              int(variant), int(bool(truncate_ends)))
        check(self._L.nautilus_csm_upload_ranges(self._h, ptr(r), r.shape[0], r.shape[1],
                                                 C.byref(spec)))
        self.n_scans = r.shape[0]
        offs = np.zeros(self.n_scans + 1, np.int64)
        check(self._L.nautilus_csm_download_scans(self._h, None, 0, ptr(offs)))
        self._scan_sizes = (offs[1:] - offs[:-1]).astype(np.int64)

    def frontend_last_ms(self) -> float:
        ms = C.c_float()
        check(self._L.nautilus_csm_frontend_last_ms(self._h, C.byref(ms)))
        return float(ms.value)

    def download_scans(self):
        """(xy float32 [n, 2], offsets int64 [n_scans + 1]) of the resident scan set."""
        offs = np.zeros(self.n_scans + 1, np.int64)
        check(self._L.nautilus_csm_download_scans(self._h, None, 0, ptr(offs)))
        xy = np.empty((int(offs[-1]), 2), np.float32)
        check(self._L.nautilus_csm_download_scans(self._h, ptr(xy), len(xy), ptr(offs)))
        return xy, offs

    def scatter_scores(self) -> np.ndarray:
        """ComputeScatterMatrixScore (lc_candidate_filter.cc:34-52) of every resident scan."""
        out = np.empty(self.n_scans, np.float64)
        check(self._L.nautilus_lc_scatter_scores(self._h, ptr(out)))
        return out

    def GetLCCandidates(self, pose_xy, *, min_dist: float = 5.0, min_score: float = 0.70):
        """LCCandidateFilter::GetLCCandidates (lc_candidate_filter.cc:66-81) over the
        resident scans: device scatter scores + the reference's sequential selection."""
        pose_xy = np.ascontiguousarray(pose_xy, np.float32).reshape(-1, 2)
        if len(pose_xy) != self.n_scans:
            raise ValueError("one pose per resident scan")
        score = self.scatter_scores()
        out = np.empty(self.n_scans, np.int64)
        n = C.c_int32()
        check(self._L.nautilus_lc_candidates(ptr(score), ptr(pose_xy), self.n_scans, min_dist,
                                             min_score, ptr(out), C.byref(n)))
        return out[:n.value].copy()

    def match_batch(self, a_ids, b_ids, rot_min, rot_max, *, with_stats: bool = False):
        a_ids = np.ascontiguousarray(a_ids, dtype=np.int32)
        b_ids = np.ascontiguousarray(b_ids, dtype=np.int32)
        n = len(a_ids)
        rot_min = np.ascontiguousarray(np.broadcast_to(np.asarray(rot_min, np.float64), (n,)))
        rot_max = np.ascontiguousarray(np.broadcast_to(np.asarray(rot_max, np.float64), (n,)))
        score = np.empty(n, np.float64)
        x, y, th = (np.empty(n, np.float32) for _ in range(3))
        stats = np.zeros((n, 3), np.int64) if with_stats else None
        check(self._L.nautilus_csm_match_batch(
            self._h, ptr(a_ids), ptr(b_ids), ptr(rot_min), ptr(rot_max), n, ptr(score), ptr(x),
            ptr(y), ptr(th), ptr(stats) if with_stats else None))
        return (score, x, y, th, stats) if with_stats else (score, x, y, th)

    # ---- multi-GPU: one batch partitioned over the ranks (include/nautilus_csm.h) ----
    @staticmethod
    def comm_unique_id() -> bytes:
        """ncclGetUniqueId through the library; rank 0 calls it and distributes the bytes."""
        buf = C.create_string_buffer(128)
        check(_lib.lib().nautilus_csm_comm_unique_id(buf))
        return buf.raw

    def comm_init(self, unique_id: bytes) -> None:
        """ncclCommInitRank for this handle's (world_rank, world_size).  Collective."""
        assert len(unique_id) == 128
        check(self._L.nautilus_csm_comm_init(self._h, C.c_char_p(unique_id)))

    def pair_weights(self, a_ids, rot_min, rot_max, offsets) -> np.ndarray:
        from . import distributed
        return distributed.pair_weights_c(self._params, a_ids, rot_min, rot_max, offsets)

    def gather(self, order, bounds, *, with_stats: bool = False):
        """After match_batch on this rank's block: all ranks' results in caller order."""
        order = np.ascontiguousarray(order, dtype=np.int32)
        bounds = np.ascontiguousarray(bounds, dtype=np.int64)
        n = int(bounds[-1])
        score = np.empty(n, np.float64)
        x, y, th = (np.empty(n, np.float32) for _ in range(3))
        stats = np.zeros((n, 3), np.int64) if with_stats else None
        check(self._L.nautilus_csm_gather(self._h, ptr(order), ptr(bounds), ptr(score), ptr(x),
                                          ptr(y), ptr(th), ptr(stats) if with_stats else None))
        return (score, x, y, th, stats) if with_stats else (score, x, y, th)

    def match_batch_sharded(self, a_ids, b_ids, rot_min, rot_max, *, with_stats: bool = False):
        """match_batch of ONE batch split over the world: every rank passes the whole
        batch and gets every result (nautilus_csm_match_batch_sharded)."""
        a_ids = np.ascontiguousarray(a_ids, dtype=np.int32)
        b_ids = np.ascontiguousarray(b_ids, dtype=np.int32)
        n = len(a_ids)
        rot_min = np.ascontiguousarray(np.broadcast_to(np.asarray(rot_min, np.float64), (n,)))
        rot_max = np.ascontiguousarray(np.broadcast_to(np.asarray(rot_max, np.float64), (n,)))
        score = np.empty(n, np.float64)
        x, y, th = (np.empty(n, np.float32) for _ in range(3))
        stats = np.zeros((n, 3), np.int64) if with_stats else None
        check(self._L.nautilus_csm_match_batch_sharded(
            self._h, ptr(a_ids), ptr(b_ids), ptr(rot_min), ptr(rot_max), n, ptr(score), ptr(x),
            ptr(y), ptr(th), ptr(stats) if with_stats else None))
        return (score, x, y, th, stats) if with_stats else (score, x, y, th)

    RESULT_BYTES = 48  # record size of nautilus_csm_copy_results_dev

    def copy_results_dev(self, dst_dev_ptr: int, n_pairs: int) -> None:
        """Stream-ordered device copy of the last batch's result records."""
        check(self._L.nautilus_csm_copy_results_dev(self._h, C.c_void_p(dst_dev_ptr), n_pairs))

    def build_tables(self, scan_ids, *, checksum: bool = True) -> float:
        ids = np.ascontiguousarray(scan_ids, dtype=np.int32)
        cs = C.c_double()
        check(self._L.nautilus_csm_build_tables(self._h, ptr(ids), len(ids),
                                                C.byref(cs) if checksum else None))
        return cs.value

    # ---- point-to-point matching (solver.cc:67-172 over kdtree.cc) -----------
    def nn_build(self, cell_size: float) -> None:
        """Per-scan search structures for the uploaded clouds; cell_size = the
        largest threshold later calls use (outlier_threshold)."""
        check(self._L.nautilus_nn_build(self._h, float(cell_size)))

    def nn_build_begin(self, cell_size: float) -> None:
        """Start the host KD-tree layout of nn_build on library threads (returns at once);
        a later nn_build(cell_size) joins it."""
        check(self._L.nautilus_nn_build_begin(self._h, float(cell_size)))

    def nn_match_batch(self, src_ids, tgt_ids, transforms, threshold: float, *,
                       with_dist: bool = False):
        """-> (match_idx int32 [sum of source sizes], offsets int64 [n_pairs+1][, dist])."""
        src_ids = np.ascontiguousarray(src_ids, dtype=np.int32)
        tgt_ids = np.ascontiguousarray(tgt_ids, dtype=np.int32)
        n = len(src_ids)
        m = np.ascontiguousarray(transforms, dtype=np.float32).reshape(n, 6)
        sizes = self._scan_sizes[src_ids] if n else np.zeros(0, np.int64)
        total = int(sizes.sum())
        idx = np.empty(max(total, 1), np.int32)
        dist = np.empty(max(total, 1), np.float32) if with_dist else None
        offsets = np.zeros(n + 1, np.int64)
        check(self._L.nautilus_nn_match_batch(
            self._h, ptr(src_ids), ptr(tgt_ids), ptr(m), float(threshold), n, ptr(idx),
            ptr(dist) if with_dist else None, ptr(offsets)))
        idx = idx[:total]
        return (idx, offsets, dist[:total]) if with_dist else (idx, offsets)

    def nn_upload_normals(self, normals) -> None:
        """One normal per uploaded point (KDNodeValue::normal of the reference's tree nodes)."""
        nrm = np.ascontiguousarray(normals, dtype=np.float32).reshape(-1, 2)
        if len(nrm) != int(self._scan_sizes.sum()):
            raise ValueError("one normal per uploaded point")
        check(self._L.nautilus_nn_upload_normals(self._h, ptr(nrm)))

    def nn_match_resident(self, src_ids, tgt_ids, transforms, threshold: float) -> np.ndarray:
        """The search of nn_match_batch with the correspondences kept on the device
        (source order, as GetPointToPointMatching appends them); -> correspondence
        offsets int64 [n_pairs + 1]."""
        src_ids = np.ascontiguousarray(src_ids, dtype=np.int32)
        tgt_ids = np.ascontiguousarray(tgt_ids, dtype=np.int32)
        n = len(src_ids)
        m = np.ascontiguousarray(transforms, dtype=np.float32).reshape(n, 6)
        offsets = np.zeros(n + 1, np.int64)
        check(self._L.nautilus_nn_match_resident(self._h, ptr(src_ids), ptr(tgt_ids), ptr(m),
                                                 float(threshold), n, ptr(offsets)))
        return offsets

    def nn_self_index(self) -> np.ndarray:
        out = np.empty(max(int(self._scan_sizes.sum()), 1), np.int32)
        check(self._L.nautilus_nn_self_index(self._h, ptr(out)))
        return out[:int(self._scan_sizes.sum())]

    def nn_last_stats(self) -> dict:
        ms, walked = C.c_float(), C.c_uint64()
        check(self._L.nautilus_nn_last_stats(self._h, C.byref(ms), C.byref(walked)))
        return {"ms_kernel": ms.value, "n_walked": int(walked.value)}

    def single_queue_stats(self) -> tuple[int, int]:
        """(single-pair calls served, batches they were combined into)."""
        n, b = C.c_int64(), C.c_int64()
        check(self._L.nautilus_csm_single_queue_stats(self._h, C.byref(n), C.byref(b)))
        return n.value, b.value

    def last_timings(self) -> dict:
        t = Timings()
        check(self._L.nautilus_csm_last_timings(self._h, C.byref(t)))
        return {k: getattr(t, k) for k, _ in Timings._fields_}
