"""Independent cross-checks of the restated NMS / top-k of oracle/filter.py (TensorFlow is not installable here, so
tf.image.non_max_suppression itself cannot be run):

  * oracle.filter.non_max_suppression (the literal per-pair restatement of non_max_suppression_op.cc) against
    oracle.filter.non_max_suppression_fast (the vectorised form every parity test uses);
  * both against torchvision.ops.nms (torchvision 0.26, an independent implementation of the same rule: greedy by
    descending score, suppress iff IoU > threshold, IoU without the "+1" pixel convention);
  * filter_detections' concat -> top_k step against torch.topk.

torchvision breaks score ties by its own sort, so the seeded scores are tie-free by construction."""
import numpy as np
import pytest
import torch

from oracle import filter as of

torchvision = pytest.importorskip("torchvision")


def _case(seed, n, clustered):
    rng = np.random.default_rng(seed)
    if clustered:   # a few objects, many near-duplicate boxes each: long suppression chains
        centres = rng.uniform(60, 400, (6, 2))
        c = centres[rng.integers(0, 6, n)] + rng.normal(0, 6, (n, 2))
        wh = rng.uniform(30, 90, (n, 2))
    else:
        c = rng.uniform(0, 480, (n, 2))
        wh = rng.uniform(8, 200, (n, 2))
    boxes = np.concatenate([c - wh / 2, c + wh / 2], axis=1).astype(np.float32)
    scores = rng.permutation(n).astype(np.float32) / n * 0.9 + 0.05   # distinct
    return boxes, scores


@pytest.mark.parametrize("seed,n,clustered", [(0, 200, False), (1, 400, True), (2, 1200, True), (3, 1500, False)])
def test_fast_nms_equals_literal_nms_equals_torchvision(seed, n, clustered):
    boxes, scores = _case(seed, n, clustered)
    for thr, cap in ((0.5, 300), (0.3, 50), (0.7, 10 ** 6)):
        fast = of.non_max_suppression_fast(boxes, scores, cap, thr)
        if n <= 400:
            slow = of.non_max_suppression(boxes, scores, cap, thr)
            assert np.array_equal(slow, fast)
        tv = torchvision.ops.nms(torch.from_numpy(boxes), torch.from_numpy(scores), thr).numpy()[:cap]
        assert np.array_equal(tv, fast), (seed, thr, cap, len(tv), len(fast))


def test_iou_exactly_at_threshold_is_kept_by_all_three():
    # [0,0,10,10] vs [0,0,10,5]... IoU = 50 / 100 = 0.5 exactly: "> 0.5" does not suppress
    boxes = np.array([[0, 0, 10, 10], [0, 0, 10, 5], [100, 100, 110, 110]], np.float32)
    scores = np.array([0.9, 0.8, 0.7], np.float32)
    want = np.array([0, 1, 2])
    assert np.array_equal(of.non_max_suppression(boxes, scores, 300, 0.5), want)
    assert np.array_equal(of.non_max_suppression_fast(boxes, scores, 300, 0.5), want)
    assert np.array_equal(torchvision.ops.nms(torch.from_numpy(boxes), torch.from_numpy(scores), 0.5).numpy(), want)


def test_inverted_and_degenerate_boxes_follow_tf_not_torchvision():
    """TF normalises corners with min/max and gives area <= 0 an IoU of 0 (non_max_suppression_op.cc IOU());
    torchvision does neither, so here only the two restatements are compared with each other and with hand values."""
    boxes = np.array([[10, 10, 0, 0], [0, 0, 10, 10], [5, 5, 5, 20], [5, 5, 5, 20]], np.float32)
    scores = np.array([0.9, 0.8, 0.7, 0.6], np.float32)
    slow = of.non_max_suppression(boxes, scores, 300, 0.5)
    fast = of.non_max_suppression_fast(boxes, scores, 300, 0.5)
    # box 1 is box 0 with its corners in the other order: IoU 1 -> suppressed; the zero-area boxes never suppress
    # each other (IoU 0) although they coincide
    assert np.array_equal(slow, [0, 2, 3]) and np.array_equal(fast, slow)


def test_top_k_of_filter_detections_equals_torch_topk():
    rng = np.random.default_rng(5)
    n, c = 3000, 6
    boxes, _ = _case(6, n, True)
    cls = (rng.permutation(n * c).astype(np.float32) / (n * c)).reshape(n, c) * 0.3   # distinct, ~83 % pass 0.05
    b3 = rng.normal(size=(n, 16)).astype(np.float32)
    out = of.filter_detections(boxes, b3, cls, max_detections=100)
    # independent recomputation: per-class torchvision NMS, concatenate in class order, torch.topk
    idx, lab = [], []
    for k in range(c):
        cand = np.where(cls[:, k] > np.float32(0.05))[0]
        keep = torchvision.ops.nms(torch.from_numpy(boxes[cand]), torch.from_numpy(cls[cand, k]), 0.5).numpy()[:100]
        idx.append(cand[keep])
        lab.append(np.full(len(keep), k))
    idx, lab = np.concatenate(idx), np.concatenate(lab)
    sc = torch.from_numpy(cls[idx, lab])
    top = torch.topk(sc, 100).indices.numpy()
    assert np.array_equal(out[2], sc.numpy()[top])
    assert np.array_equal(out[3], lab[top].astype(np.int32))
    assert np.array_equal(out[0], boxes[idx[top]])
