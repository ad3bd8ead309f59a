"""ctypes binding of libmdi_b200.so (the C ABI of include/mdi_b200.h).

There is no CPU fallback: if the library is missing or the device is not an sm_100 part the
calls raise.  PyTorch is only used by callers for device memory and streams; this module
passes raw pointers.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

PKG_DIR = Path(__file__).resolve().parent
LIB_PATH = PKG_DIR / "libmdi_b200.so"

# ---- constants mirrored from include/mdi_b200.h ------------------------------------------------
ABI_VERSION = 5
TILE_PX = 4096
TILE_WORDS = TILE_PX // 32
PATCH_STRIDE = 2048
INTERP_STATIC_MAX_PX = 2 * 1024 * 1024
FACTOR_TABLE_STRIDE = 40
STATUS_OK, STATUS_LOW_CONFIDENCE, STATUS_EMPTY_STATS, STATUS_STATS_RETRY = 0, 1, 2, 4
STATS_EXACT, STATS_RETRY_FLAGGED, STATS_RAW = 64, 128, 512
IMAGE_F32_UNIT, IMAGE_F32_255, IMAGE_U8 = 0, 1, 2
SUBSAMPLE_STATIC, SUBSAMPLE_ADAPTIVE = 0, 1
KEEP_PRED_MASK, KEEP_SENTINEL, KEEP_NONNEG, KEEP_PATCH, KEEP_GRAD, KEEP_SUBSAMPLE = 1, 2, 4, 8, 16, 32
KEEP_PREZEROED = 256
LOSS_RANSAC, LOSS_MSAC = 0, 1

# struct MdiView (align=True reproduces the C layout; checked against mdi_sizeof_view)
VIEW_DTYPE = np.dtype(
    [
        ("depth", np.uint64), ("mask", np.uint64), ("image", np.uint64),
        ("H", np.int32), ("W", np.int32), ("sfm_begin", np.int32), ("sfm_count", np.int32),
        ("px_begin", np.int64), ("tile_begin", np.int64), ("n_tiles", np.int32),
        ("patch_h", np.int32), ("patch_w", np.int32), ("grid_h", np.int32), ("grid_w", np.int32),
        ("_pad", np.int32),
        ("P", np.float32, (12,)), ("Kinv", np.float32, (9,)), ("c2w", np.float32, (12,)),
    ],
    align=True,
)


class RansacConfigC(C.Structure):
    _fields_ = [
        ("inlier_threshold", C.c_float), ("max_iters", C.c_int32), ("confidence", C.c_double),
        ("sample_size", C.c_int32), ("min_iters", C.c_int32), ("loss_kind", C.c_int32),
        ("view_base", C.c_int32), ("seed", C.c_uint64),
    ]


class SubsampleConfigC(C.Structure):
    _fields_ = [
        ("mode", C.c_int32), ("static_factor", C.c_int32), ("factor_min", C.c_int32),
        ("factor_max", C.c_int32), ("flags", C.c_int32), ("grad_thresh", C.c_float),
    ]


class MdiError(RuntimeError):
    pass


_P = C.c_void_p
_I64 = C.c_int64
_INT = C.c_int

_SIGNATURES = {
    "mdi_last_error": (C.c_char_p, []),
    "mdi_abi_version": (_INT, []),
    "mdi_check_device": (_INT, [_INT]),
    "mdi_sizeof_view": (_INT, []),
    "mdi_profile_enable": (None, [_INT]),
    "mdi_profile_collect": (_INT, [_P, _INT, _P, _P, _INT]),
    "mdi_setup_cameras": (_INT, [_P, _INT, _P, _P, _P, _P, _INT, _P]),
    "mdi_project_filter": (_INT, [_P, _INT, _P, _P, _I64, _P, _P, _P, _P, _P, _P]),
    "mdi_setup_project_filter": (_INT, [_P, _INT, _P, _P, _P, _P, _INT, _P, _P, _I64, _P, _P, _P, _P, _P, _P]),
    "mdi_gather_fit": (_INT, [_P, _INT, _I64, _P, _P, _P, _P, _P, _P, _INT, _INT, _P, _P]),
    "mdi_ransac": (_INT, [_P, _INT, _P, _P, _P, _P, _P, C.POINTER(RansacConfigC), _INT, _P, _P, _P, _P, _P]),
    "mdi_apply_scale_shift": (_INT, [_P, _I64, _P, _P, _P]),
    "mdi_apply_region_scale_shift": (_INT, [_P, _P, _I64, _P, _P, _INT, _P, _P]),
    "mdi_depth_range_workspace_bytes": (_I64, [_INT, _I64]),
    "mdi_depth_range": (_INT, [_P, _INT, _I64, _P, _P, _INT, _P, _P, _P, _P, _P]),
    "mdi_depth_range_map": (_INT, [_INT, _I64, _P, _P, _INT, _P, _P, _P, _P, _P]),
    "mdi_grad_range": (_INT, [_P, _INT, _I64, _P, _P, _P, _P]),
    "mdi_factor_table": (_INT, [_INT, _P, _INT, _INT, _P, _P]),
    "mdi_keep_mask": (_INT, [_P, _INT, _I64, _I64, _INT, _INT, _P, _P, _P, _P, _P, _P, C.POINTER(SubsampleConfigC), _P, _P, _P, _P]),
    "mdi_scan_tiles": (_INT, [_P, _INT, _I64, _P, _P, _P, _P, _P, _P]),
    "mdi_emit_points": (_INT, [_P, _INT, _I64, _P, _P, _P, _P, _INT, _P, _P, _P]),
    "mdi_upload_table": (_INT, [_P, _P, _I64, _P]),
    "mdi_emit_points_capped": (_INT, [_P, _INT, _I64, _P, _P, _P, _P, _INT, _P, _P, _P, _I64, _P]),
    "mdi_emit_points_multi": (_INT, [_P, _INT, _I64, _P, _P, _P, _P, _INT, _INT, _P, _P, _I64, _P]),
    "mdi_emit_points_multicast": (_INT, [_P, _INT, _I64, _P, _P, _P, _P, _INT, _P, _P, _I64, _P]),
    "mdi_emit_points_exchange": (_INT, [_P, _INT, _I64, _P, _P, _P, _P, _INT, _INT, _P, _P, _P, _P, _P, _I64, _P, _INT, _P]),
    "mdi_expand_rgb8": (_INT, [_P, _P, _P, _INT, _I64, _P]),
    "mdi_exchange_counts": (_INT, [_P, _INT, _INT, C.c_uint32, _P, _P, _P]),
    "mdi_min_extents": (_INT, [_P, _I64, _P, _P, _P, _INT, _P, _P]),
    "mdi_knn_workspace_bytes": (_I64, [_I64]),
    "mdi_knn": (_INT, [_P, _I64, _INT, _INT, _P, _P, _P, _P, _P]),
    "mdi_lof": (_INT, [_I64, _INT, _P, _P, _P, _P, _P, _P, _P]),
    "mdi_knn_build": (_INT, [_P, _I64, _P, _P]),
    "mdi_knn_split": (_INT, [_P, _I64, _INT, _P, _P]),
    "mdi_knn_query": (_INT, [_P, _I64, _INT, _INT, _I64, _I64, _P, _P, _P, _P, _P]),
    "mdi_lof_lrd": (_INT, [_I64, _INT, _P, _P, _P, _P, _P]),
    "mdi_lof_score": (_INT, [_I64, _INT, _P, _P, _P, _P, _P, _P]),
    "mdi_interp_workspace_bytes": (_I64, [_INT, _INT]),
    "mdi_interp_scale_map": (_INT, [_P, _P, _P, _P, _INT, C.c_double, _INT, _INT, _P, _P, _P, _P, _P]),
    "mdi_interp_scale_map_static": (_INT, [_P, _P, _P, _P, _INT, C.c_double, _INT, _INT, _P, _P, _P, _P]),
    "mdi_merge_workspace_bytes": (_I64, [_I64]),
    "mdi_merge_subsample": (_INT, [_P, _P, _P, _I64, C.c_float, C.c_float, _INT, _P, _P, _P,
                                   C.POINTER(C.c_int64), C.POINTER(C.c_int32), _P]),
}

_lib = None


def exported_symbols() -> list[str]:
    return list(_SIGNATURES)


def load() -> C.CDLL:
    """Load libmdi_b200.so; raises MdiError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise MdiError(
            f"{LIB_PATH} is missing: build it with `python __graft_entry__.py` (nvcc, sm_100a). "
            "There is no CPU fallback for this path."
        )
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.mdi_abi_version() != ABI_VERSION:
        raise MdiError(f"ABI mismatch: library {lib.mdi_abi_version()} vs binding {ABI_VERSION}")
    if lib.mdi_sizeof_view() != VIEW_DTYPE.itemsize:
        raise MdiError(f"MdiView layout mismatch: C {lib.mdi_sizeof_view()} vs numpy {VIEW_DTYPE.itemsize}")
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().mdi_last_error().decode(errors="replace")
        raise MdiError(f"{what or 'mdi call'} failed ({rc}): {msg}")


def ptr(t) -> int | None:
    """Device (or pinned host) pointer of a torch tensor, None passes NULL."""
    if t is None:
        return None
    return t.data_ptr()


def profile_collect() -> dict:
    """{kernel name: {"ms": total device ms, "calls": n}} since the last collect (profiling on)."""
    lib = load()
    cap = 64
    names = C.create_string_buffer(4096)
    ms = (C.c_float * cap)()
    calls = (C.c_int32 * cap)()
    n = lib.mdi_profile_collect(C.cast(names, C.c_void_p), 4096, C.cast(ms, C.c_void_p), C.cast(calls, C.c_void_p), cap)
    keys = names.value.decode().split("\n") if n else []
    return {k: {"ms": float(ms[i]), "calls": int(calls[i])} for i, k in enumerate(keys[:n])}
