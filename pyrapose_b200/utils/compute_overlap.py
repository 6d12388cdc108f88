"""Drop-in for the reference's Cython extension `PyraPose.utils.compute_overlap`
(PyraPose/utils/compute_overlap.pyx:13-53): same name, same argument strictness, numpy in / numpy out,
computed by the float64 CUDA kernel behind ppb_compute_overlap_f64 (bit-exact with the Cython loop).

CUDA is initialised on first call, not at import, so forked data-loader workers that never call it stay clean
(PyraPose/bin/train.py:372-390 forks workers; call this from the parent or from spawned workers)."""
import numpy as np


def _check(a, what):
    # Cython typed buffers `np.ndarray[double, ndim=2]` raise ValueError on a dtype or ndim mismatch.
    if not isinstance(a, np.ndarray):
        raise TypeError("Argument '%s' has incorrect type (expected numpy.ndarray, got %s)" % (what, type(a).__name__))
    if a.dtype != np.float64:
        raise ValueError("Buffer dtype mismatch, expected 'double' but got '%s'" % a.dtype.name)
    if a.ndim != 2:
        raise ValueError("Buffer has wrong number of dimensions (expected 2, got %d)" % a.ndim)


def compute_overlap(boxes, query_boxes):
    """
    Args
        boxes:       (N, 4) ndarray of float64
        query_boxes: (K, 4) ndarray of float64
    Returns
        overlaps: (N, K) ndarray of overlap between boxes and query_boxes
    """
    _check(boxes, "boxes")
    _check(query_boxes, "query_boxes")
    n, k = boxes.shape[0], query_boxes.shape[0]
    if n == 0 or k == 0:
        return np.zeros((n, k), dtype=np.float64)
    import torch
    from .. import _kernels
    dev = torch.device("cuda", torch.cuda.current_device())
    b = torch.from_numpy(np.ascontiguousarray(boxes[:, :4])).to(dev)
    q = torch.from_numpy(np.ascontiguousarray(query_boxes[:, :4])).to(dev)
    return _kernels.compute_overlap(b, q).cpu().numpy()


def compute_overlap_device(boxes, query_boxes):
    """Device-tensor overload: float64 CUDA tensors in, CUDA tensor out; batched query (B,K,4) supported."""
    from .. import _kernels
    return _kernels.compute_overlap(boxes, query_boxes)
