"""numpy restatement of the reference's anchor utilities — TEST INFRASTRUCTURE ONLY.

Follows /root/reference/PyraPose/utils/anchors.py (line numbers cited per function). Pinned: every
function here is checked against the reference's own source executed in the build container
(tests/golden/make_golden.py -> tests/golden/anchors_*.npz, tests/test_oracle_anchors.py); anchor_targets_bbox as a
WHOLE (states, one-hot labels, projected-corner regression targets, mask targets, padded batches) and the Anchors
layer's float32 arithmetic are pinned by tests/golden/reference_text.npz (the reference's utils/anchors.py and
layers/_misc.py imported and run, make_golden_shim.py; transforms3d's quat2mat restated, PIL is installed) in
tests/test_oracle_reference_text.py, sha256 of the full arrays included.
"""
import numpy as np

from . import overlap as _overlap


class AnchorParameters:
    """anchors.py:26-42."""

    def __init__(self, sizes, strides, ratios, scales):
        self.sizes = sizes
        self.strides = strides
        self.ratios = ratios
        self.scales = scales

    def num_anchors(self):
        return len(self.ratios) * len(self.scales)


# anchors.py:48-53 — ratios and scales are float32 arrays (keras.backend.floatx()).
AnchorParameters.default = AnchorParameters(
    sizes=[32, 64, 128],
    strides=[8, 16, 32],
    ratios=np.array([0.5, 1, 2], np.float32),
    scales=np.array([2 ** 0, 2 ** (1.0 / 3.0), 2 ** (2.0 / 3.0)], np.float32),
)

# The 5-level variant the generic builder would use (sizes/strides of stock keras-retinanet; the level
# list [3,4,5,6,7] is the commented default at anchors.py:391).
AnchorParameters.five_level = AnchorParameters(
    sizes=[32, 64, 128, 256, 512],
    strides=[8, 16, 32, 64, 128],
    ratios=AnchorParameters.default.ratios,
    scales=AnchorParameters.default.scales,
)


def generate_anchors(base_size=16, ratios=None, scales=None):
    """anchors.py:447-478 (float64 arithmetic on float32-valued ratios/scales)."""
    if ratios is None:
        ratios = AnchorParameters.default.ratios
    if scales is None:
        scales = AnchorParameters.default.scales
    num_anchors = len(ratios) * len(scales)
    anchors = np.zeros((num_anchors, 4))
    anchors[:, 2:] = base_size * np.tile(scales, (2, len(ratios))).T
    areas = anchors[:, 2] * anchors[:, 3]
    anchors[:, 2] = np.sqrt(areas / np.repeat(ratios, len(scales)))
    anchors[:, 3] = anchors[:, 2] * np.repeat(ratios, len(scales))
    anchors[:, 0::2] -= np.tile(anchors[:, 2] * 0.5, (2, 1)).T
    anchors[:, 1::2] -= np.tile(anchors[:, 3] * 0.5, (2, 1)).T
    return anchors


def shift(shape, stride, anchors):
    """anchors.py:415-444."""
    shift_x = (np.arange(0, shape[1]) + 0.5) * stride
    shift_y = (np.arange(0, shape[0]) + 0.5) * stride
    shift_x, shift_y = np.meshgrid(shift_x, shift_y)
    shifts = np.vstack((shift_x.ravel(), shift_y.ravel(), shift_x.ravel(), shift_y.ravel())).transpose()
    A = anchors.shape[0]
    K = shifts.shape[0]
    all_anchors = anchors.reshape((1, A, 4)) + shifts.reshape((1, K, 4)).transpose((1, 0, 2))
    return all_anchors.reshape((K * A, 4))


def guess_shapes(image_shape, pyramid_levels):
    """anchors.py:357-369."""
    image_shape = np.array(image_shape[:2])
    return [(image_shape + 2 ** x - 1) // (2 ** x) for x in pyramid_levels]


def anchors_for_shape(image_shape, pyramid_levels=None, anchor_params=None, shapes_callback=None):
    """anchors.py:372-412."""
    if pyramid_levels is None:
        pyramid_levels = [3, 4, 5]
    if anchor_params is None:
        anchor_params = AnchorParameters.default
    if shapes_callback is None:
        shapes_callback = guess_shapes
    image_shapes = shapes_callback(image_shape, pyramid_levels)
    all_anchors = np.zeros((0, 4))
    for idx, _p in enumerate(pyramid_levels):
        anchors = generate_anchors(base_size=anchor_params.sizes[idx], ratios=anchor_params.ratios,
                                   scales=anchor_params.scales)
        shifted = shift(image_shapes[idx], anchor_params.strides[idx], anchors)
        all_anchors = np.append(all_anchors, shifted, axis=0)
    return all_anchors


def anchors_layer_f32(feature_shapes, anchor_params=None):
    """What the Keras `Anchors` layers + Concatenate produce for one image (float32):
    layers/_misc.py:52-56 casts the float64 base anchors to float32 (keras.backend.variable);
    backend/common.py:93-116 builds float32 shifts (arange + 0.5) * stride and adds them in float32;
    models/retinanet.py:232-243 concatenates the levels."""
    if anchor_params is None:
        anchor_params = AnchorParameters.default
    out = []
    for idx, (fh, fw) in enumerate(feature_shapes):
        base = generate_anchors(anchor_params.sizes[idx], anchor_params.ratios, anchor_params.scales).astype(np.float32)
        stride = np.float32(anchor_params.strides[idx])
        sx = (np.arange(0, fw, dtype=np.float32) + np.float32(0.5)) * stride
        sy = (np.arange(0, fh, dtype=np.float32) + np.float32(0.5)) * stride
        sx, sy = np.meshgrid(sx, sy)
        shifts = np.stack([sx.reshape(-1), sy.reshape(-1), sx.reshape(-1), sy.reshape(-1)], axis=0).T
        k = shifts.shape[0]
        a = base.shape[0]
        shifted = base.reshape(1, a, 4) + shifts.reshape(k, 1, 4).astype(np.float32)
        out.append(shifted.reshape(k * a, 4))
    return np.concatenate(out, axis=0).astype(np.float32)


def compute_gt_annotations(anchors, annotations, negative_overlap=0.4, positive_overlap=0.5, compute_overlap=None):
    """anchors.py:290-318. `compute_overlap` defaults to the C port (oracle.overlap)."""
    if compute_overlap is None:
        compute_overlap = _overlap.compute_overlap
    overlaps = compute_overlap(anchors.astype(np.float64), annotations.astype(np.float64))
    argmax_overlaps_inds = np.argmax(overlaps, axis=1)
    max_overlaps = overlaps[np.arange(overlaps.shape[0]), argmax_overlaps_inds]
    positive_indices = max_overlaps >= positive_overlap
    ignore_indices = (max_overlaps > negative_overlap) & ~positive_indices
    return positive_indices, ignore_indices, argmax_overlaps_inds


def bbox_transform(anchors, gt_boxes, mean=None, std=None):
    """anchors.py:481-512."""
    if mean is None:
        mean = np.array([0, 0, 0, 0])
    if std is None:
        std = np.array([0.2, 0.2, 0.2, 0.2])
    mean = np.asarray(mean)
    std = np.asarray(std)
    aw = anchors[:, 2] - anchors[:, 0]
    ah = anchors[:, 3] - anchors[:, 1]
    t = np.stack(((gt_boxes[:, 0] - anchors[:, 0]) / aw, (gt_boxes[:, 1] - anchors[:, 1]) / ah,
                  (gt_boxes[:, 2] - anchors[:, 2]) / aw, (gt_boxes[:, 3] - anchors[:, 3]) / ah)).T
    return (t - mean) / std


def box3D_transform(anchors, gt_boxes, num_classes=None, mean=None, std=None):
    """anchors.py:515-559: target j = (gt[j] - anchor[j % 4]) / (w if j even else h), then (t - mean) / std."""
    if mean is None:
        mean = np.zeros(16)
    if std is None:
        std = np.full(16, 0.2)
    mean = np.asarray(mean)
    std = np.asarray(std)
    aw = anchors[:, 2] - anchors[:, 0]
    ah = anchors[:, 3] - anchors[:, 1]
    cols = []
    for j in range(16):
        cols.append((gt_boxes[:, j] - anchors[:, j % 4]) / (aw if j % 2 == 0 else ah))
    t = np.stack(cols).T
    return (t - mean) / std


def toPix_array(translation, fx=None, fy=None, cx=None, cy=None):
    """anchors.py:562-567."""
    xpix = ((translation[:, 0] * fx) / translation[:, 2]) + cx
    ypix = ((translation[:, 1] * fy) / translation[:, 2]) + cy
    return np.stack((xpix, ypix), axis=1)


def quat2mat(q):
    """transforms3d.quaternions.quat2mat (w, x, y, z), as consumed at anchors.py:207 — restated from the
    published algorithm of transforms3d 0.3.1 (not vendored in the reference)."""
    w, x, y, z = q
    Nq = w * w + x * x + y * y + z * z
    if Nq < np.finfo(np.float64).eps:
        return np.eye(3)
    s = 2.0 / Nq
    X, Y, Z = x * s, y * s, z * s
    wX, wY, wZ = w * X, w * Y, w * Z
    xX, xY, xZ = x * X, x * Y, x * Z
    yY, yZ, zZ = y * Y, y * Z, z * Z
    return np.array([[1.0 - (yY + zZ), xY - wZ, xZ + wY],
                     [xY + wZ, 1.0 - (xX + zZ), yZ - wX],
                     [xZ - wY, yZ + wX, 1.0 - (xX + yY)]])


def anchor_targets_bbox(anchors, image_group, annotations_group, num_classes, negative_overlap=0.4,
                        positive_overlap=0.5, compute_overlap=None):
    """anchors.py:72-287 (bbox / 3D-box / mask-target logic; debug drawing omitted).

    annotations: dict with 'bboxes' (n,4), 'labels' (n,), 'poses' (n,7: t then quaternion wxyz),
    'segmentations' (n,8,3), 'cam_params' (n,4: fx,fy,cx,cy), 'mask' ([HxW] uint8 id image), 'mask_ids' (n,).
    """
    assert len(image_group) == len(annotations_group)
    assert len(annotations_group) > 0
    batch_size = len(image_group)
    n = anchors.shape[0]
    labels_batch = np.zeros((batch_size, n, num_classes + 1), dtype=np.float32)
    regression_3D = np.zeros((batch_size, n, 16 + 1), dtype=np.float32)
    image_shapes = guess_shapes(image_group[0].shape[:2], [3])
    mh, mw = int(image_shapes[0][0]), int(image_shapes[0][1])
    mask_batch = np.zeros((batch_size, mh * mw, num_classes + 1), dtype=np.float32)

    for index, (image, annotations) in enumerate(zip(image_group, annotations_group)):
        mask = annotations['mask'][0]
        if annotations['bboxes'].shape[0]:
            pos, ign, argmax = compute_gt_annotations(anchors, annotations['bboxes'], negative_overlap,
                                                      positive_overlap, compute_overlap)
            labels_batch[index, ign, -1] = -1
            labels_batch[index, pos, -1] = 1
            regression_3D[index, ign, -1] = -1
            regression_3D[index, pos, -1] = 1
            labels_batch[index, pos, annotations['labels'][argmax[pos]].astype(int)] = 1

            calculated_boxes = np.empty((0, 16))
            for idx, pose in enumerate(annotations['poses']):
                cls = int(annotations['labels'][idx])
                mask_id = annotations['mask_ids'][idx]
                # PIL Image.resize(NEAREST) of the id mask to the P3 grid (anchors.py:158)
                mask_flat = nearest_resize_pil(mask, mw, mh).flatten()
                anchors_spec = np.where(mask_flat == int(mask_id))[0]
                if len(anchors_spec) > 1:
                    mask_batch[index, anchors_spec, cls] = 1
                    mask_batch[index, anchors_spec, -1] = 1
                rot = np.asarray(quat2mat(pose[3:]), dtype=np.float32)
                tra = pose[:3]
                tDbox = rot[:3, :3].dot(annotations['segmentations'][idx].T).T
                tDbox = tDbox + np.repeat(tra[np.newaxis, 0:3], 8, axis=0)
                cam = annotations['cam_params'][idx]
                box3D = toPix_array(tDbox, fx=cam[0], fy=cam[1], cx=cam[2], cy=cam[3])
                box3D = np.reshape(box3D, (16))
                calculated_boxes = np.concatenate([calculated_boxes, [box3D]], axis=0)
            regression_3D[index, :, :-1] = box3D_transform(anchors, calculated_boxes[argmax, :], num_classes)

        if image.shape:
            centers = np.vstack([(anchors[:, 0] + anchors[:, 2]) / 2, (anchors[:, 1] + anchors[:, 3]) / 2]).T
            outside = np.logical_or(centers[:, 0] >= image.shape[1], centers[:, 1] >= image.shape[0])
            labels_batch[index, outside, -1] = -1
            regression_3D[index, outside, -1] = -1
    return regression_3D, labels_batch, mask_batch


def nearest_resize_pil(mask, out_w, out_h):
    """PIL's Image.resize(..., NEAREST) index rule: src = floor((dst + 0.5) * in / out)."""
    in_h, in_w = mask.shape[:2]
    ys = np.minimum((np.floor((np.arange(out_h) + 0.5) * in_h / out_h)).astype(int), in_h - 1)
    xs = np.minimum((np.floor((np.arange(out_w) + 0.5) * in_w / out_w)).astype(int), in_w - 1)
    return mask[ys][:, xs]
