"""numpy restatement of the reference's per-anchor decode ops — TEST INFRASTRUCTURE ONLY.

Follows /root/reference/PyraPose/backend/common.py (TensorFlow float32 elementwise graph ops). float32 throughout,
each multiply/add rounded separately exactly as the unfused TF graph evaluates them.
PINNED TO THE REFERENCE'S TEXT: tests/golden/reference_text.npz holds the outputs of the reference's own common.py
and layers/_misc.py, imported and run over numpy stand-ins for the tensorflow / keras primitives
(tests/golden/tf_shim.py, make_golden_shim.py); tests/test_oracle_reference_text.py compares bit for bit. What stays
unpinned is TensorFlow's own kernels (they cannot run here), i.e. that TF evaluates these float32 ops unfused.
"""
import numpy as np

F = np.float32


def bbox_transform_inv(boxes, deltas, mean=None, std=None):
    """common.py:6-22. boxes (B,N,4) f32, deltas (B,N,4) f32."""
    mean = np.zeros(4, F) if mean is None else np.asarray(mean, F)
    std = np.full(4, 0.2, F) if std is None else np.asarray(std, F)
    boxes = boxes.astype(F)
    deltas = deltas.astype(F)
    width = boxes[:, :, 2] - boxes[:, :, 0]
    height = boxes[:, :, 3] - boxes[:, :, 1]
    out = np.empty_like(deltas)
    for j in range(4):
        size = width if j % 2 == 0 else height
        out[:, :, j] = boxes[:, :, j] + (deltas[:, :, j] * std[j] + mean[j]) * size
    return out


def box3D_transform_inv(boxes, deltas, mean=None, std=None):
    """common.py:25-56: output j = boxes[j % 4] + (d[j]*std[j] + mean[j]) * (width if j even else height)."""
    mean = np.zeros(16, F) if mean is None else np.asarray(mean, F)
    std = np.full(16, 0.2, F) if std is None else np.asarray(std, F)
    boxes = boxes.astype(F)
    deltas = deltas.astype(F)
    width = boxes[:, :, 2] - boxes[:, :, 0]
    height = boxes[:, :, 3] - boxes[:, :, 1]
    out = np.empty_like(deltas)
    for j in range(16):
        size = width if j % 2 == 0 else height
        out[:, :, j] = boxes[:, :, j % 4] + (deltas[:, :, j] * std[j] + mean[j]) * size
    return out


def box3D_denormalization(boxes, deltas, mean=None, std=None):
    """common.py:59-90: as box3D_transform_inv, then x / 320.0 and y / 240.0."""
    out = box3D_transform_inv(boxes, deltas, mean, std)
    out[:, :, 0::2] = out[:, :, 0::2] / F(320.0)
    out[:, :, 1::2] = out[:, :, 1::2] / F(240.0)
    return out


def clip_boxes(image_shape_hw, boxes):
    """layers/_misc.py:259-280 (ClipBoxes): clip x to [0, width], y to [0, height]."""
    h, w = F(image_shape_hw[0]), F(image_shape_hw[1])
    out = boxes.astype(F).copy()
    out[:, :, 0] = np.clip(out[:, :, 0], 0, w)
    out[:, :, 1] = np.clip(out[:, :, 1], 0, h)
    out[:, :, 2] = np.clip(out[:, :, 2], 0, w)
    out[:, :, 3] = np.clip(out[:, :, 3], 0, h)
    return out


def hull_boxes(boxes3D):
    """Axis-aligned hull (x1,y1,x2,y2) of the 8 decoded corners — this repo's documented convenience
    input for FilterDetections (the PyraPose heads never produce 2-D boxes)."""
    xs = boxes3D[..., 0::2]
    ys = boxes3D[..., 1::2]
    return np.stack([xs.min(-1), ys.min(-1), xs.max(-1), ys.max(-1)], axis=-1).astype(F)
