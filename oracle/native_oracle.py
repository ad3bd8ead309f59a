"""TEST INFRASTRUCTURE -- ctypes loader for oracle/_build/libmdi_native_oracle.so (the C
restatement of native_modules/subsampling/src/impl.cpp; pinned against the reference's own compiled
module in oracle/_ref by tests/test_native_ref_pin.py, see native_oracle.c).
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this."""
import ctypes as C

import numpy as np

import build_oracle

_lib = None


def _load():
    global _lib
    if _lib is None:
        _lib = C.CDLL(str(build_oracle.build()))
        _lib.oracle_min_extents.restype = None
        _lib.oracle_min_extents.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                            C.c_void_p]
        _lib.oracle_merge_subsample.restype = C.c_int64
        _lib.oracle_merge_subsample.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float,
                                                C.c_void_p, C.c_void_p, C.c_int64]
    return _lib


def _f32(a, shape):
    a = np.ascontiguousarray(a, dtype=np.float32).reshape(shape)
    return a


def min_extents(pts, K, P, sizes):
    """impl.cpp:70-126.  pts [N,3], K [C,3,3], P [C,3,4], sizes [C,2] (width, height) -> ext [N]."""
    pts = _f32(pts, (-1, 3)); K = _f32(K, (-1, 3, 3)); P = _f32(P, (-1, 3, 4))
    sizes = np.ascontiguousarray(sizes, dtype=np.int32).reshape(-1, 2)
    ext = np.empty(pts.shape[0], np.float32)
    _load().oracle_min_extents(pts.ctypes.data, pts.shape[0], K.ctypes.data, P.ctypes.data, sizes.ctypes.data,
                               K.shape[0], ext.ctypes.data)
    return ext


def merge_subsample(pts, rgb, ext, max_bbox_aspect_ratio=1.1, min_extent_multiplier=1.0, max_frames=None):
    """impl.cpp:313-426.  Returns (pts', rgb') in the reference's emission order; raises when the
    reference would not terminate."""
    pts = _f32(pts, (-1, 3)); rgb = _f32(rgb, (-1, 3)); ext = _f32(ext, (-1,))
    n = pts.shape[0]
    out_p = np.empty((n, 3), np.float32); out_c = np.empty((n, 3), np.float32)
    if max_frames is None:
        max_frames = 4096 * max(n, 1)
    m = _load().oracle_merge_subsample(pts.ctypes.data, rgb.ctypes.data, ext.ctypes.data, n,
                                       max_bbox_aspect_ratio, min_extent_multiplier, out_p.ctypes.data,
                                       out_c.ctypes.data, max_frames)
    if m < 0:
        raise RuntimeError("merge_subsample: frame budget exhausted (the reference does not terminate here)")
    return out_p[:m].copy(), out_c[:m].copy()


def subsample_pointcloud(pts, rgb, K, P, sizes, max_bbox_aspect_ratio=1.1, min_extent_multiplier=1.0):
    """pointcloud_subsampling.cpp:22-67 minus the debug cloud."""
    ext = min_extents(pts, K, P, sizes)
    p, c = merge_subsample(pts, rgb, ext, max_bbox_aspect_ratio, min_extent_multiplier)
    return p, c, ext
