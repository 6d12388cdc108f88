"""numpy restatement of filter_detections and the evaluators' vote filter — TEST INFRASTRUCTURE ONLY.

filter_detections follows /root/reference/PyraPose/layers/filter_detections.py:21-118. The
TensorFlow ops it calls are restated from their published semantics (TF 2.1.1, not in the tree):
  tf.where(mask)                        row-major ascending int64 indices
  tf.image.non_max_suppression (V3)     greedy in descending score; a candidate is dropped iff its IoU with an
                                        already selected box is > iou_threshold; IoU on min/max-normalised corners,
                                        no "+1", area<=0 -> 0, float32; stops at max_output_size; ties in score are
                                        broken towards the lower index (TF>=2.2 behaviour; TF 2.1 heap order is
                                        unspecified, synthetic scores are tie-free)
  tf.nn.top_k                           descending, ties -> lower index
  tf.pad(constant_values=-1)
PINNED TO THE REFERENCE'S TEXT for everything filter_detections.py itself decides (threshold, class loop,
concatenation order, the gather / top_k / pad sequence, dtypes, `other`): tests/golden/reference_text.npz holds the
outputs of the reference's own filter_detections / FilterDetections.call run over numpy stand-ins for the TF
primitives (tests/golden/tf_shim.py), compared exactly in tests/test_oracle_reference_text.py. The primitives above
remain restatements (TensorFlow cannot run here); NMS is cross-checked against torchvision.ops.nms.

vote_filter follows /root/reference/PyraPose/utils/linemod_eval.py:307-334, 427-430.
"""
import numpy as np

F = np.float32


def _iou_tf(boxes, i, j):
    """tensorflow/core/kernels/non_max_suppression_op.cc IOU(), float32; coordinates used as (y1,x1,y2,x2)
    positions 0..3 — symmetric in the two axes, so feeding (x1,y1,x2,y2) gives the same value."""
    bi = boxes[i]
    bj = boxes[j]
    ymin_i, xmin_i = min(bi[0], bi[2]), min(bi[1], bi[3])
    ymax_i, xmax_i = max(bi[0], bi[2]), max(bi[1], bi[3])
    ymin_j, xmin_j = min(bj[0], bj[2]), min(bj[1], bj[3])
    ymax_j, xmax_j = max(bj[0], bj[2]), max(bj[1], bj[3])
    area_i = F(F(ymax_i - ymin_i) * F(xmax_i - xmin_i))
    area_j = F(F(ymax_j - ymin_j) * F(xmax_j - xmin_j))
    if area_i <= 0 or area_j <= 0:
        return F(0.0)
    iy0, ix0 = max(ymin_i, ymin_j), max(xmin_i, xmin_j)
    iy1, ix1 = min(ymax_i, ymax_j), min(xmax_i, xmax_j)
    inter = F(max(F(iy1 - iy0), F(0.0)) * max(F(ix1 - ix0), F(0.0)))
    return F(inter / F(F(area_i + area_j) - inter))


def non_max_suppression(boxes, scores, max_output_size, iou_threshold):
    """Returns selected indices (into boxes) in selection order."""
    boxes = np.asarray(boxes, F)
    scores = np.asarray(scores, F)
    order = np.lexsort((np.arange(len(scores)), -scores.astype(np.float64)))  # score desc, index asc
    selected = []
    thr = F(iou_threshold)
    for cand in order:
        if len(selected) >= max_output_size:
            break
        keep = True
        for s in reversed(selected):
            if _iou_tf(boxes, cand, s) > thr:
                keep = False
                break
        if keep:
            selected.append(int(cand))
    return np.asarray(selected, dtype=np.int64)


def non_max_suppression_fast(boxes, scores, max_output_size, iou_threshold):
    """Same result as non_max_suppression, vectorised per candidate (for larger cases)."""
    boxes = np.asarray(boxes, F)
    scores = np.asarray(scores, F)
    n = len(scores)
    order = np.lexsort((np.arange(n), -scores.astype(np.float64)))
    a0 = np.minimum(boxes[:, 0], boxes[:, 2])
    a1 = np.minimum(boxes[:, 1], boxes[:, 3])
    a2 = np.maximum(boxes[:, 0], boxes[:, 2])
    a3 = np.maximum(boxes[:, 1], boxes[:, 3])
    area = ((a2 - a0).astype(F) * (a3 - a1).astype(F)).astype(F)
    thr = F(iou_threshold)
    sel = []
    for cand in order:
        if len(sel) >= max_output_size:
            break
        if sel:
            s = np.asarray(sel)
            iy0 = np.maximum(a0[cand], a0[s])
            ix0 = np.maximum(a1[cand], a1[s])
            iy1 = np.minimum(a2[cand], a2[s])
            ix1 = np.minimum(a3[cand], a3[s])
            inter = (np.maximum((iy1 - iy0).astype(F), F(0)) * np.maximum((ix1 - ix0).astype(F), F(0))).astype(F)
            denom = ((area[cand] + area[s]).astype(F) - inter).astype(F)
            with np.errstate(divide="ignore", invalid="ignore"):
                iou = (inter / denom).astype(F)
            iou = np.where((area[cand] <= 0) | (area[s] <= 0), F(0), iou)
            if np.any(iou > thr):
                continue
        sel.append(int(cand))
    return np.asarray(sel, dtype=np.int64)


def filter_detections(boxes, boxes3D, classification, other=(), class_specific_filter=True, nms=True,
                      score_threshold=0.05, max_detections=300, nms_threshold=0.5):
    """filter_detections.py:21-118 for ONE image. Returns [boxes (D,4), boxes3D (D,16), scores (D,), labels (D,) int32]
    + filtered `other`, padded with -1."""
    boxes = np.asarray(boxes, F)
    boxes3D = np.asarray(boxes3D, F)
    classification = np.asarray(classification, F)
    thr = F(score_threshold)

    def _filter(scores, labels):
        idx = np.where(scores > thr)[0].astype(np.int64)                       # :54
        if nms:
            keep = non_max_suppression_fast(boxes[idx], scores[idx], max_detections, nms_threshold)  # :57-61
            idx = idx[keep]                                                     # :64
        return np.stack([idx, labels[idx]], axis=1)                             # :67-68

    if class_specific_filter:
        all_idx = []
        for c in range(classification.shape[1]):                                # :75
            scores = classification[:, c]
            labels = np.full(scores.shape[0], c, dtype=np.int64)
            all_idx.append(_filter(scores, labels))
        indices = np.concatenate(all_idx, axis=0) if all_idx else np.zeros((0, 2), np.int64)  # :81
    else:
        scores = classification.max(axis=1)                                     # :83
        labels = classification.argmax(axis=1).astype(np.int64)                 # :84
        indices = _filter(scores, labels)

    scores = classification[indices[:, 0], indices[:, 1]]                       # :88
    labels = indices[:, 1]
    k = min(max_detections, scores.shape[0])                                    # :90
    top = np.lexsort((np.arange(len(scores)), -scores.astype(np.float64)))[:k]  # top_k: desc, ties -> lower index
    scores = scores[top]
    sel = indices[top, 0]                                                       # :93
    out_boxes = boxes[sel]
    out_boxes3D = boxes3D[sel]
    labels = labels[top]
    other_ = [np.asarray(o)[sel] for o in other]
    pad = max(0, max_detections - k)                                            # :100

    def padm(a):
        width = [(0, pad)] + [(0, 0)] * (a.ndim - 1)
        return np.pad(a, width, constant_values=-1)

    return [padm(out_boxes), padm(out_boxes3D), padm(scores), padm(labels).astype(np.int32)] + [padm(o) for o in other_]


def filter_detections_batch(boxes, boxes3D, classification, **kw):
    """FilterDetections.call (filter_detections.py:153-191): map over the batch."""
    outs = [filter_detections(boxes[b], boxes3D[b], classification[b], **kw) for b in range(boxes.shape[0])]
    return [np.stack([o[i] for o in outs], axis=0) for i in range(4)]


def vote_filter(scores, boxes3D, threshold=0.5, min_votes=10):
    """linemod_eval.py:307-334, 427-430 for ONE image: per class c, ascending anchor indices with
    score > threshold; classes with fewer than `min_votes` hits are skipped (returns empty arrays for them);
    votes = boxes3D[idx] (k,16); confidence = sum of the hit scores (pyrapose_node.py:638)."""
    scores = np.asarray(scores, F)
    boxes3D = np.asarray(boxes3D, F)
    out = []
    for c in range(scores.shape[1]):
        idx = np.where(scores[:, c] > F(threshold))[0].astype(np.int64)
        if len(idx) < min_votes:
            out.append((np.zeros(0, np.int64), np.zeros((0, 16), F), len(idx)))
            continue
        out.append((idx, boxes3D[idx, :], len(idx)))
    return out
