"""ctypes front-end of the C port in compute_overlap_port.c — TEST INFRASTRUCTURE ONLY.
Same signature and input strictness as the reference's Cython function
(/root/reference/PyraPose/utils/compute_overlap.pyx:13-16: float64, ndim 2, else ValueError)."""
import ctypes

import numpy as np

from . import build as _build

_lib = None


def _load():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(_build.build_port())
        lib.oracle_compute_overlap.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t,
                                               ctypes.c_size_t]
        lib.oracle_compute_overlap.restype = None
        lib.oracle_gt_assign.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t,
                                         ctypes.c_double, ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p,
                                         ctypes.c_void_p, ctypes.c_void_p]
        lib.oracle_gt_assign.restype = None
        _lib = lib
    return _lib


def _check(a, name):
    if not isinstance(a, np.ndarray) or a.dtype != np.float64:
        raise ValueError("Buffer dtype mismatch, expected 'double' for %s" % name)
    if a.ndim != 2:
        raise ValueError("Buffer has wrong number of dimensions (expected 2, got %d)" % a.ndim)


def compute_overlap(boxes, query_boxes):
    _check(boxes, "boxes")
    _check(query_boxes, "query_boxes")
    b = np.ascontiguousarray(boxes[:, :4])
    q = np.ascontiguousarray(query_boxes[:, :4])
    out = np.empty((b.shape[0], q.shape[0]), dtype=np.float64)
    _load().oracle_compute_overlap(b.ctypes.data, q.ctypes.data, out.ctypes.data, b.shape[0], q.shape[0])
    return out


def gt_assign(boxes, query_boxes, negative_overlap=0.4, positive_overlap=0.5):
    """Fused compute_gt_annotations: returns (positive bool[N], ignore bool[N], argmax int64[N], max f64[N])."""
    _check(boxes, "boxes")
    _check(query_boxes, "query_boxes")
    b = np.ascontiguousarray(boxes[:, :4])
    q = np.ascontiguousarray(query_boxes[:, :4])
    n = b.shape[0]
    argmax = np.empty(n, dtype=np.int64)
    pos = np.empty(n, dtype=np.uint8)
    ign = np.empty(n, dtype=np.uint8)
    mx = np.empty(n, dtype=np.float64)
    _load().oracle_gt_assign(b.ctypes.data, q.ctypes.data, n, q.shape[0], negative_overlap, positive_overlap,
                             argmax.ctypes.data, pos.ctypes.data, ign.ctypes.data, mx.ctypes.data)
    return pos.astype(bool), ign.astype(bool), argmax, mx
