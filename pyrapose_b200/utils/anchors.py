"""Anchor utilities — host-side mirror of PyraPose/utils/anchors.py with the heavy parts on the GPU.

  anchors_for_shape / shift      -> ppb_anchors_shift_f64 (bit-exact float64 grids; anchors.py:372-444)
  compute_gt_annotations         -> ppb_gt_assign_f64 (fused IoU + argmax + thresholds; anchors.py:290-318 and the
                                    Cython compute_overlap it calls)
  anchor_targets_bbox            -> same signature / outputs as anchors.py:72-287; assignment on the GPU for the
                                    whole batch in one launch, the per-object pose projection stays numpy
The nine base anchors per level are tiny and are computed on the host with the reference's own float64 recipe
(anchors.py:447-478). `Generator(compute_anchor_targets=anchor_targets_bbox, compute_shapes=guess_shapes)`
(PyraPose/preprocessing/generator.py:58-59) is the plugin seam these functions fit.
"""
import numpy as np


class AnchorParameters:
    """ The parameters that define how anchors are generated (anchors.py:26-42).

    Args
        sizes   : List of sizes to use. Each size corresponds to one feature level.
        strides : List of strides to use. Each stride correspond to one feature level.
        ratios  : List of ratios to use per location in a feature map.
        scales  : List of scales to use per location in a feature map.
    """

    def __init__(self, sizes, strides, ratios, scales):
        self.sizes = sizes
        self.strides = strides
        self.ratios = ratios
        self.scales = scales

    def num_anchors(self):
        return len(self.ratios) * len(self.scales)


# anchors.py:48-53 (ratios / scales are float32 arrays: keras.backend.floatx())
AnchorParameters.default = AnchorParameters(
    sizes=[32, 64, 128],
    strides=[8, 16, 32],
    ratios=np.array([0.5, 1, 2], np.float32),
    scales=np.array([2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0)], np.float32),
)


def _device():
    import torch
    return torch.device("cuda", torch.cuda.current_device())


def generate_anchors(base_size=16, ratios=None, scales=None):
    """Base (reference) windows: every ratio x every scale around the origin, float64, (A, 4) as (x1, y1, x2, y2)."""
    ratios = AnchorParameters.default.ratios if ratios is None else ratios
    scales = AnchorParameters.default.scales if scales is None else scales
    n_r, n_s = len(ratios), len(scales)
    # scale varies fastest; the product keeps the dtype of `scales` (float32 by default) before widening
    side = (base_size * np.tile(scales, n_r)).astype(np.float64)
    ratio = np.repeat(ratios, n_s)                                   # float32 values, promoted in the divisions
    width = np.sqrt((side * side) / ratio)
    height = width * ratio
    out = np.zeros((n_r * n_s, 4))
    out[:, 0] = 0.0 - width * 0.5
    out[:, 1] = 0.0 - height * 0.5
    out[:, 2] = width - width * 0.5
    out[:, 3] = height - height * 0.5
    return out


def guess_shapes(image_shape, pyramid_levels):
    """Feature-map shape per pyramid level: (dim + 2^l - 1) // 2^l."""
    hw = np.array(image_shape[:2])
    return [(hw + 2 ** lvl - 1) // (2 ** lvl) for lvl in pyramid_levels]


def shift(shape, stride, anchors):
    """All base anchors at every cell centre ((i + 0.5) * stride), x fastest; (H*W*A, 4) float64 numpy."""
    import torch
    from .. import _kernels
    base = torch.from_numpy(np.ascontiguousarray(anchors, dtype=np.float64)).to(_device())
    return _kernels.anchors_shift(base, int(shape[0]), int(shape[1]), float(stride)).cpu().numpy()


def anchors_for_shape_device(image_shape, pyramid_levels=None, anchor_params=None, shapes_callback=None):
    """anchors_for_shape, result left on the GPU (float64 CUDA tensor)."""
    import torch
    from .. import _kernels
    pyramid_levels = [3, 4, 5] if pyramid_levels is None else pyramid_levels
    anchor_params = AnchorParameters.default if anchor_params is None else anchor_params
    shapes = (guess_shapes if shapes_callback is None else shapes_callback)(image_shape, pyramid_levels)
    dev = _device()
    sizes = [int(s[0]) * int(s[1]) * anchor_params.num_anchors() for s in shapes]
    out = torch.empty((sum(sizes), 4), dtype=torch.float64, device=dev)
    row = 0
    for idx, _lvl in enumerate(pyramid_levels):
        base = generate_anchors(anchor_params.sizes[idx], anchor_params.ratios, anchor_params.scales)
        _kernels.anchors_shift(torch.from_numpy(base).to(dev), int(shapes[idx][0]), int(shapes[idx][1]),
                               float(anchor_params.strides[idx]), out=out[row:row + sizes[idx]])
        row += sizes[idx]
    return out


def anchors_for_shape(image_shape, pyramid_levels=None, anchor_params=None, shapes_callback=None):
    """ Generators anchors for a given shape: np.array (N, 4) float64 of (x1, y1, x2, y2)."""
    return anchors_for_shape_device(image_shape, pyramid_levels, anchor_params, shapes_callback).cpu().numpy()


def compute_gt_annotations(anchors, annotations, negative_overlap=0.4, positive_overlap=0.5):
    """ Obtain indices of gt annotations with the greatest overlap.

    Returns
        positive_indices: bool (N,)   max IoU >= positive_overlap
        ignore_indices:   bool (N,)   max IoU >  negative_overlap and not positive
        argmax_overlaps_inds: int64 (N,)
    """
    import torch
    from .. import _kernels
    dev = _device()
    a = torch.from_numpy(np.ascontiguousarray(anchors[:, :4].astype(np.float64))).to(dev)
    q = torch.from_numpy(np.ascontiguousarray(annotations[:, :4].astype(np.float64))).to(dev)
    pos, ign, arg = _kernels.gt_assign(a, q, None, negative_overlap, positive_overlap)
    return pos.cpu().numpy(), ign.cpu().numpy(), arg.cpu().numpy()


def _check_mean_std(v, default, what):
    if v is None:
        return np.asarray(default)
    if isinstance(v, (list, tuple)):
        return np.array(v)
    if not isinstance(v, np.ndarray):
        raise ValueError('Expected {} to be a np.ndarray, list or tuple. Received: {}'.format(what, type(v)))
    return v


def bbox_transform(anchors, gt_boxes, mean=None, std=None):
    """Box regression targets: (gt - anchor) / (w, h, w, h), then (t - mean) / std."""
    mean = _check_mean_std(mean, [0, 0, 0, 0], 'mean')
    std = _check_mean_std(std, [0.2, 0.2, 0.2, 0.2], 'std')
    size = np.stack([anchors[:, 2] - anchors[:, 0], anchors[:, 3] - anchors[:, 1]], axis=1)
    targets = (gt_boxes[:, :4] - anchors[:, :4]) / np.tile(size, 2)
    return (targets - mean) / std


def box3D_transform(anchors, gt_boxes, num_classes=None, mean=None, std=None):
    """3D-box-corner regression targets: corner j is measured from anchor coordinate j % 4 and divided by the anchor
    width (x values) or height (y values); then (t - mean) / std."""
    mean = _check_mean_std(mean, [0] * 16, 'mean')
    std = _check_mean_std(std, [0.2] * 16, 'std')
    size = np.stack([anchors[:, 2] - anchors[:, 0], anchors[:, 3] - anchors[:, 1]], axis=1)
    targets = (gt_boxes[:, :16] - np.tile(anchors[:, :4], 4)) / np.tile(size, 8)
    return (targets - mean) / std


def toPix_array(translation, fx=None, fy=None, cx=None, cy=None):
    """Pinhole projection of (n, 3) camera-frame points to pixels."""
    xpix = ((translation[:, 0] * fx) / translation[:, 2]) + cx
    ypix = ((translation[:, 1] * fy) / translation[:, 2]) + cy
    return np.stack((xpix, ypix), axis=1)


def _quat2mat(q):
    # transforms3d.quaternions.quat2mat, (w, x, y, z)
    w, x, y, z = q
    nq = w * w + x * x + y * y + z * z
    if nq < np.finfo(np.float64).eps:
        return np.eye(3)
    s = 2.0 / nq
    X, Y, Z = x * s, y * s, z * s
    return np.array([[1.0 - (y * Y + z * Z), x * Y - w * Z, x * Z + w * Y],
                     [x * Y + w * Z, 1.0 - (x * X + z * Z), y * Z - w * X],
                     [x * Z - w * Y, y * Z + w * X, 1.0 - (x * X + y * Y)]])


def _nearest_resize(mask, out_w, out_h):
    in_h, in_w = mask.shape[:2]
    ys = np.minimum(np.floor((np.arange(out_h) + 0.5) * in_h / out_h).astype(int), in_h - 1)
    xs = np.minimum(np.floor((np.arange(out_w) + 0.5) * in_w / out_w).astype(int), in_w - 1)
    return mask[ys][:, xs]


def _projected_corners(ann):
    """(n, 16) float64: image projections of the 8 corners of every object's 3D box (anchors.py:207-215)."""
    out = np.empty((len(ann['poses']), 16))
    for idx, pose in enumerate(ann['poses']):
        rot = np.asarray(_quat2mat(pose[3:]), dtype=np.float32)
        corners = rot[:3, :3].dot(ann['segmentations'][idx].T).T + np.repeat(pose[np.newaxis, 0:3], 8, axis=0)
        cam = ann['cam_params'][idx]
        out[idx] = np.reshape(toPix_array(corners, fx=cam[0], fy=cam[1], cx=cam[2], cy=cam[3]), (16))
    return out


def anchor_targets_bbox_device(anchors, image_group, annotations_group, num_classes, negative_overlap=0.4,
                               positive_overlap=0.5):
    """ anchor_targets_bbox with the targets left on the GPU: returns three CUDA tensors
    (regression_3D (B, N, 17), labels (B, N, C + 1), mask (B, H/8*W/8, C + 1)) that model.train_on_batch consumes
    without a host round trip. The host only packs the annotation lists (a few hundred numbers per image: boxes,
    labels, the projected 3D-box corners of anchors.py:207-215 in float64, mask ids); the N x K IoU assignment, the
    per-anchor states / one-hot labels / 16 regression targets and the mask targets run in CUDA kernels
    (ppb_gt_assign_f64, ppb_anchor_targets, ppb_mask_targets) with the reference's float64 arithmetic."""
    import torch
    from .. import _kernels
    assert (len(image_group) == len(annotations_group)), "The length of the images and annotations need to be equal."
    assert (len(annotations_group) > 0), "No data received to compute anchor targets for."
    for annotations in annotations_group:
        for key in ('bboxes', 'labels', 'poses', 'segmentations'):
            assert (key in annotations), "Annotations should contain %s." % key

    batch = len(image_group)
    dev = anchors.device if isinstance(anchors, torch.Tensor) else _device()
    kmax = max(1, max(a['bboxes'].shape[0] for a in annotations_group))
    q = np.zeros((batch, kmax, 4), dtype=np.float64)
    b16 = np.zeros((batch, kmax, 16), dtype=np.float64)
    lab = np.zeros((batch, kmax), dtype=np.int32)
    ids = np.full((batch, kmax), -1, dtype=np.int32)
    cnt = np.zeros(batch, dtype=np.int32)
    wh = np.full((batch, 2), -1, dtype=np.int32)
    mh, mw = [int(v) for v in guess_shapes(image_group[0].shape[:2], [3])[0]]
    masks = None
    for i, (image, ann) in enumerate(zip(image_group, annotations_group)):
        k = ann['bboxes'].shape[0]
        if np.shape(image):
            wh[i] = (image.shape[1], image.shape[0])
        if not k:
            continue
        cnt[i] = k
        q[i, :k] = ann['bboxes'][:, :4]
        lab[i, :k] = np.asarray(ann['labels']).astype(int)
        b16[i, :k] = _projected_corners(ann)
        ids[i, :k] = np.asarray(ann['mask_ids']).astype(int)
        m = np.asarray(ann['mask'][0])
        if masks is None:
            masks = np.zeros((batch,) + m.shape[:2], dtype=np.uint8)
        masks[i] = m
    a_dev = anchors if isinstance(anchors, torch.Tensor) else \
        torch.from_numpy(np.ascontiguousarray(anchors, dtype=np.float64)).to(dev)
    t = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    cnt_d, lab_d = t(cnt), t(lab)
    reg, cls = _kernels.anchor_targets(a_dev, t(q), cnt_d, lab_d, t(b16), t(wh), num_classes, negative_overlap,
                                       positive_overlap)
    if masks is None:
        mask = torch.zeros((batch, mh * mw, num_classes + 1), dtype=torch.float32, device=dev)
    else:
        mask = _kernels.mask_targets(t(masks), t(ids), lab_d, cnt_d, mh, mw, num_classes)
    return reg, cls, mask


def anchor_targets_bbox(anchors, image_group, annotations_group, num_classes, negative_overlap=0.4,
                        positive_overlap=0.5):
    """ Generate anchor targets for bbox detection (same contract as anchors.py:72-287).

    Returns (regression_3D (B, N, 17), labels_batch (B, N, num_classes + 1), mask_batch (B, H/8*W/8, num_classes + 1))
    float32 numpy; the last column is the anchor state (-1 ignore, 0 background, 1 foreground). Computed on the GPU
    (anchor_targets_bbox_device) and copied back.
    """
    return tuple(x.cpu().numpy() for x in anchor_targets_bbox_device(anchors, image_group, annotations_group,
                                                                     num_classes, negative_overlap, positive_overlap))
