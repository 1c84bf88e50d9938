"""ctypes binding of libnautilus_csm_b200.so (include/nautilus_csm.h).

The library is the product; this module only loads it.  There is no CPU
fallback: if the shared library is missing, or no B200 is visible, calls fail
loudly (RuntimeError / CsmError).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_DIR = os.path.join(_HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libnautilus_csm_b200.so")
SOURCES = ["csm_tables.cu", "csm_search.cu", "csm_capi.cu", "csm_comm.cu", "csm_image_ops.cu", "nn_kernels.cu", "nn_capi.cu", "csm_uncertainty.cu", "frontend.cu", "residuals.cu"]
HEADERS = ["csm_common.cuh", "csm_kernels.h", "csm_handle.h", "nn_kernels.h", os.path.join("..", "..", "include", "nautilus_csm.h"), "log_table.inc"]

NVCC_FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "--fmad=false", "-Xcompiler", "-fPIC", "-shared"]


class CsmError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"nautilus_csm error {code}: {msg}")
        self.code = code


class Params(C.Structure):
    _fields_ = [("scanner_range", C.c_double), ("trans_range", C.c_double),
                ("low_res", C.c_double), ("high_res", C.c_double),
                ("rot_step", C.c_double), ("blur_sigma", C.c_double),
                ("device", C.c_int32), ("margin_cells", C.c_int32),
                ("max_tables", C.c_int32), ("stage_cells", C.c_int32),
                ("world_rank", C.c_int32), ("world_size", C.c_int32)]


class LaserSpec(C.Structure):
    _fields_ = [("angle_min", C.c_float), ("angle_max", C.c_float), ("angle_increment", C.c_float),
                ("range_min", C.c_float), ("max_range", C.c_double), ("variant", C.c_int32),
                ("truncate_ends", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("n_low", C.c_int64), ("n_high", C.c_int64), ("n_pops", C.c_int64)]


class Timings(C.Structure):
    _fields_ = [("ms_total", C.c_float), ("ms_tables", C.c_float), ("ms_low", C.c_float),
                ("ms_refine", C.c_float), ("launches", C.c_int32), ("waves", C.c_int32),
                ("spec_windows", C.c_int32), ("ms_gather", C.c_float),
                ("ms_host_plan", C.c_float), ("ms_host_collect", C.c_float), ("ms_host_wait", C.c_float)]


EXPORTS = [
    "nautilus_csm_last_error", "nautilus_csm_version", "nautilus_csm_create",
    "nautilus_csm_destroy", "nautilus_csm_set_stream", "nautilus_csm_table_dims",
    "nautilus_csm_cell", "nautilus_csm_cell_fast", "nautilus_csm_upload_scans", "nautilus_csm_highres_table",
    "nautilus_csm_lowres_from_table", "nautilus_csm_build_tables",
    "nautilus_csm_match_batch", "nautilus_csm_copy_results_dev", "nautilus_csm_match", "nautilus_csm_rotation_window",
    "nautilus_csm_last_timings", "nautilus_csm_uncertainty", "nautilus_csm_uncertainty_batch", "nautilus_csm_local_uncertainty", "nautilus_csm_regular_check",
    "nautilus_nn_build", "nautilus_nn_build_begin", "nautilus_nn_pose_transform", "nautilus_nn_match_batch",
    "nautilus_nn_self_index", "nautilus_nn_last_stats",
    "nautilus_host_alloc", "nautilus_host_free",
    "nautilus_nn_upload_normals", "nautilus_nn_match_resident", "nautilus_residuals_eval_resident",
    "nautilus_csm_single_queue_stats", "nautilus_csm_table_blur", "nautilus_csm_table_normalize",
    "nautilus_csm_table_normalize_noalloc", "nautilus_csm_math_check",
    "nautilus_csm_partition", "nautilus_csm_pair_weights", "nautilus_csm_comm_unique_id",
    "nautilus_csm_comm_init", "nautilus_csm_comm_attach", "nautilus_csm_gather",
    "nautilus_csm_match_batch_sharded",
    "nautilus_csm_upload_ranges", "nautilus_csm_download_scans", "nautilus_lc_scatter_scores",
    "nautilus_lc_candidates", "nautilus_csm_frontend_last_ms", "nautilus_residuals_eval", "nautilus_residuals_last_ms",
]


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def _stale(obj: str, src: str) -> bool:
    if not os.path.exists(obj):
        return True
    t = os.path.getmtime(obj)
    deps = [src] + [os.path.join(CSRC, h) for h in HEADERS]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile the CUDA library for sm_100a in-tree (nvcc cross-compiles without a GPU).
    One object per translation unit, compiled in parallel, then one link."""
    if not force and not needs_build():
        return LIB_PATH
    from concurrent.futures import ThreadPoolExecutor
    os.makedirs(LIB_DIR, exist_ok=True)
    obj_dir = os.path.join(LIB_DIR, "obj")
    os.makedirs(obj_dir, exist_ok=True)
    cflags = [f for f in NVCC_FLAGS if f != "-shared"]

    def compile_one(name):
        src = os.path.join(CSRC, name)
        obj = os.path.join(obj_dir, name.replace(".cu", ".o"))
        if not force and not _stale(obj, src):
            return obj, ""
        cmd = ["nvcc", *cflags, *(["-Xptxas", "-v"] if verbose else []), "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        return obj, r.stderr

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        res = list(ex.map(compile_one, SOURCES))
    if verbose:
        print("".join(log for _, log in res))
    cmd = ["nvcc", "-shared", "-gencode", "arch=compute_100a,code=sm_100a",
           *[o for o, _ in res], "-o", LIB_PATH, "-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; "
                           "g.build()'` (there is no CPU fallback)")
    # NAUTILUS_CSM_LIB: a tuning variant built by scripts/ab_build.sh (A/B runs only)
    L = C.CDLL(os.environ.get("NAUTILUS_CSM_LIB") or LIB_PATH)
    vp, d, f = C.c_void_p, C.c_double, C.c_float
    i32, i64, u64 = C.c_int32, C.c_int64, C.c_uint64
    L.nautilus_csm_last_error.restype = C.c_char_p
    L.nautilus_csm_version.restype = C.c_char_p
    L.nautilus_csm_create.argtypes = [C.POINTER(Params), C.POINTER(vp)]
    L.nautilus_csm_destroy.argtypes = [vp]
    L.nautilus_csm_destroy.restype = None
    L.nautilus_csm_set_stream.argtypes = [vp, vp]
    L.nautilus_csm_table_dims.argtypes = [vp, C.POINTER(u64), C.POINTER(u64)]
    L.nautilus_csm_cell.restype = i64
    L.nautilus_csm_cell.argtypes = [d, u64, f]
    L.nautilus_csm_cell_fast.restype = i64
    L.nautilus_csm_cell_fast.argtypes = [d, u64, f]
    L.nautilus_csm_upload_scans.argtypes = [vp, vp, vp, i32]
    L.nautilus_csm_highres_table.argtypes = [vp, vp, i64, vp, vp]
    L.nautilus_csm_lowres_from_table.argtypes = [vp, vp, u64, u64, d, vp]
    L.nautilus_csm_build_tables.argtypes = [vp, vp, i32, C.POINTER(d)]
    L.nautilus_csm_match_batch.argtypes = [vp, vp, vp, vp, vp, i32, vp, vp, vp, vp, vp]
    L.nautilus_csm_copy_results_dev.argtypes = [vp, vp, i32]
    L.nautilus_csm_single_queue_stats.argtypes = [vp, C.POINTER(i64), C.POINTER(i64)]
    L.nautilus_csm_table_blur.argtypes = [vp, u64, u64, d, i32]
    L.nautilus_csm_table_normalize.argtypes = [vp, u64, u64, i32]
    L.nautilus_csm_table_normalize_noalloc.argtypes = [vp, vp, u64, u64, i32, vp]
    L.nautilus_csm_math_check.argtypes = [i64, u64, i32, C.POINTER(i64), C.POINTER(i64)]
    L.nautilus_csm_partition.argtypes = [vp, vp, i32, i32, vp, vp]
    L.nautilus_csm_pair_weights.argtypes = [C.POINTER(Params), vp, vp, vp, vp, i32, vp]
    L.nautilus_csm_comm_unique_id.argtypes = [vp]
    L.nautilus_csm_comm_init.argtypes = [vp, vp]
    L.nautilus_csm_comm_attach.argtypes = [vp, vp, i32, i32]
    L.nautilus_csm_gather.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp]
    L.nautilus_csm_match_batch_sharded.argtypes = [vp, vp, vp, vp, vp, i32, vp, vp, vp, vp, vp]
    L.nautilus_csm_match.argtypes = [vp, vp, i64, vp, i64, d, d, C.POINTER(d), C.POINTER(f),
                                     C.POINTER(f), C.POINTER(f), C.POINTER(Stats)]
    L.nautilus_csm_rotation_window.restype = None
    L.nautilus_csm_rotation_window.argtypes = [d, d, d, C.POINTER(d), C.POINTER(d)]
    L.nautilus_csm_last_timings.argtypes = [vp, C.POINTER(Timings)]
    L.nautilus_csm_uncertainty.argtypes = [vp, vp, i64, vp, i64, i32, d, vp]
    L.nautilus_csm_uncertainty_batch.argtypes = [vp, vp, vp, vp, i32, vp]
    L.nautilus_csm_local_uncertainty.argtypes = [vp, vp, i64, vp, vp, i32, vp, vp]
    L.nautilus_csm_regular_check.argtypes = [vp, i64, u64, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]
    L.nautilus_host_alloc.argtypes = [u64, C.POINTER(vp)]
    L.nautilus_host_free.argtypes = [vp]
    L.nautilus_nn_upload_normals.argtypes = [vp, vp]
    L.nautilus_nn_match_resident.argtypes = [vp, vp, vp, vp, d, i32, vp]
    L.nautilus_residuals_eval_resident.argtypes = [vp, i32, vp, vp, vp, vp, vp]
    L.nautilus_nn_build.argtypes = [vp, d]
    L.nautilus_nn_build_begin.argtypes = [vp, d]
    L.nautilus_nn_pose_transform.restype = None
    L.nautilus_nn_pose_transform.argtypes = [vp, vp, vp]
    L.nautilus_nn_match_batch.argtypes = [vp, vp, vp, vp, d, i32, vp, vp, vp]
    L.nautilus_nn_self_index.argtypes = [vp, vp]
    L.nautilus_nn_last_stats.argtypes = [vp, C.POINTER(f), C.POINTER(u64)]
    L.nautilus_csm_upload_ranges.argtypes = [vp, vp, i32, i32, C.POINTER(LaserSpec)]
    L.nautilus_csm_download_scans.argtypes = [vp, vp, i64, vp]
    L.nautilus_lc_scatter_scores.argtypes = [vp, vp]
    L.nautilus_lc_candidates.argtypes = [vp, vp, i32, d, d, vp, C.POINTER(i32)]
    L.nautilus_residuals_eval.argtypes = [vp, i32, vp, vp, vp, i32, vp, vp, vp, vp, vp, vp, vp]
    L.nautilus_residuals_last_ms.argtypes = [vp, C.POINTER(f)]
    L.nautilus_csm_frontend_last_ms.argtypes = [vp, C.POINTER(f)]
    _lib = L
    return L


def check(rc: int):
    if rc != 0:
        raise CsmError(rc, lib().nautilus_csm_last_error().decode())


def ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class _PinnedBlock:
    """Owner of one nautilus_host_alloc block; freed when the last array over it goes."""

    def __init__(self, nbytes: int):
        p = C.c_void_p()
        check(lib().nautilus_host_alloc(int(nbytes), C.byref(p)))
        self.p, self.nbytes = p, int(nbytes)

    def __del__(self):
        try:
            if self.p:
                lib().nautilus_host_free(self.p)
        except Exception:  # interpreter shutdown
            pass


def pinned_empty(shape, dtype) -> np.ndarray:
    """np.empty over page-locked host memory (nautilus_host_alloc)."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) if np.ndim(shape) else int(shape)
    if n == 0:
        return np.empty(shape, dtype)
    blk = _PinnedBlock(n * dtype.itemsize)
    buf = (C.c_char * blk.nbytes).from_address(blk.p.value)
    buf._owner = blk  # the ctypes object (the array's base) keeps the block alive
    return np.frombuffer(buf, dtype=dtype, count=n).reshape(shape)
