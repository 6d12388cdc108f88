"""Training-side targets through the public utils API (the Generator's injection points,
PyraPose/preprocessing/generator.py:58-59): anchors_for_shape, compute_gt_annotations, compute_overlap and
anchor_targets_bbox against the oracle — bit-exact for anchors, IoU, states and one-hot labels; the float32
regression targets are produced from identical float64 arithmetic, so they are compared exactly too."""
import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu


def test_public_anchor_and_overlap_api_bit_exact(cuda_device):
    from oracle import anchors as oa, overlap as oo
    from pyrapose_b200.utils import anchors as ua
    from pyrapose_b200.utils.compute_overlap import compute_overlap
    for shape in ((480, 640, 3), (427, 561, 3)):
        assert np.array_equal(ua.anchors_for_shape(shape), oa.anchors_for_shape(shape))
    a = oa.anchors_for_shape((480, 640, 3))
    q = synth.gt_boxes(1000, 100)
    ov = compute_overlap(a, q)
    assert ov.dtype == np.float64 and np.array_equal(ov, oo.compute_overlap(a, q))
    ann = np.concatenate([q, np.zeros((100, 1))], axis=1)      # callers pass (K, 5): box + label (anchors.py:310)
    pos, ign, arg = ua.compute_gt_annotations(a, ann)
    p2, i2, a2 = oa.compute_gt_annotations(a, ann)
    assert np.array_equal(pos, p2) and np.array_equal(ign, i2) and np.array_equal(arg, a2)
    assert arg.dtype == np.int64 and pos.dtype == np.bool_


@pytest.mark.parametrize("num_classes,hw", [(21, (480, 640)), (15, (240, 320))], ids=["ycbv21_480x640", "linemod15_240x320"])
def test_anchor_targets_bbox_matches_oracle(cuda_device, num_classes, hw):
    from oracle import anchors as oa
    from pyrapose_b200.utils import anchors as ua
    H, W = hw
    anchors = oa.anchors_for_shape((H, W, 3))
    B = 4
    images = [np.zeros((H, W, 3), np.float32) for _ in range(B)]
    anns = [synth.pose_annotations(40 + i, num_classes, width=W, height=H) for i in range(B)]
    anns[2] = {k: (v[:0] if isinstance(v, np.ndarray) else v) for k, v in anns[2].items()}   # an image without objects
    got = ua.anchor_targets_bbox(anchors, images, anns, num_classes)
    ref = oa.anchor_targets_bbox(anchors, images, anns, num_classes)
    names = ("regression_3D", "labels", "mask")
    for g, r, nm in zip(got, ref, names):
        assert g.shape == r.shape and g.dtype == np.float32, nm
        assert np.array_equal(g, r), "%s differs (max abs %g)" % (nm, np.abs(g - r).max())
    assert got[0].shape == (B, anchors.shape[0], 17) and got[1].shape == (B, anchors.shape[0], num_classes + 1)
    assert (got[1][..., -1] == 1).sum() > 0 and (got[1][2, :, -1] == 1).sum() == 0
    with pytest.raises(AssertionError):
        ua.anchor_targets_bbox(anchors, images, anns[:2], num_classes)
