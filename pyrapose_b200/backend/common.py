"""Mirror of PyraPose/backend/common.py: same function names and argument meaning, CUDA tensors in/out
(numpy arrays are uploaded). The arithmetic runs in decode.cu / anchors.cu."""
import numpy as np


def _cuda(x, dtype=None):
    import torch
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(np.ascontiguousarray(x))
    if not x.is_cuda:
        x = x.cuda()
    if dtype is not None and x.dtype != dtype:
        x = x.to(dtype)
    return x


def bbox_transform_inv(boxes, deltas, mean=None, std=None):
    """common.py:6-22."""
    import torch
    from .. import _kernels
    return _kernels.bbox_transform_inv(_cuda(boxes, torch.float32), _cuda(deltas, torch.float32), mean, std)


def box3D_transform_inv(boxes, deltas, mean=None, std=None):
    """common.py:25-56."""
    import torch
    from .. import _kernels
    return _kernels.box3d_transform_inv(_cuda(boxes, torch.float32), _cuda(deltas, torch.float32), mean, std)


def box3D_denormalization(boxes, deltas, mean=None, std=None):
    """common.py:59-90."""
    import torch
    from .. import _kernels
    return _kernels.box3d_transform_inv(_cuda(boxes, torch.float32), _cuda(deltas, torch.float32), mean, std,
                                        denormalize=True)


def shift(shape, stride, anchors):
    """common.py:93-116: float32 anchors (A,4) shifted over a (H, W) feature grid -> (H*W*A, 4) float32."""
    import torch
    from .. import _kernels
    return _kernels.anchors_shift(_cuda(anchors, torch.float32), int(shape[0]), int(shape[1]), float(stride))
