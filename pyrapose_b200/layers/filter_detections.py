"""FilterDetections — mirror of PyraPose/layers/filter_detections.py (function :21-118, layer :121-234).
Score threshold -> per-class NMS -> top-k -> -1 padding, for the whole batch, on the device (filter.cu)."""
from ..backend.common import _cuda


def filter_detections(boxes, boxes3D, classification, other=[], class_specific_filter=True, nms=True,
                      score_threshold=0.05, max_detections=300, nms_threshold=0.5):
    """ Filter detections of ONE image. boxes (N,4), boxes3D (N,16), classification (N,C), other: list of (N,...).
    Returns [boxes (D,4), boxes3D (D,16), scores (D,), labels (D,) int32] + filtered other, padded with -1."""
    outs = _filter_batch(_cuda(boxes)[None], _cuda(boxes3D)[None], _cuda(classification)[None],
                         [_cuda(o)[None] for o in other], class_specific_filter, nms, score_threshold, max_detections,
                         nms_threshold)
    return [o[0] for o in outs]


def _filter_batch(boxes, boxes3D, classification, other, class_specific_filter, nms, score_threshold, max_detections,
                  nms_threshold):
    import torch
    from .. import _kernels
    res = _kernels.filter_detections(boxes.float(), boxes3D.float(), classification.float(),
                                     class_specific_filter=class_specific_filter, nms=nms,
                                     score_threshold=score_threshold, max_detections=max_detections,
                                     nms_threshold=nms_threshold, want_indices=bool(other))
    outs = list(res[:4])
    for o in other:
        g = _kernels.gather_rows(o.float(), res[4])
        outs.append(g.to(o.dtype) if o.dtype != torch.float32 else g)
    return outs


class FilterDetections(object):
    """ Layer for filtering detections using score threshold, NMS and selecting the top-k detections."""

    def __init__(self, nms=True, class_specific_filter=True, nms_threshold=0.5, score_threshold=0.05,
                 max_detections=300, parallel_iterations=32, name='filtered_detections', **kwargs):
        self.nms = nms
        self.class_specific_filter = class_specific_filter
        self.nms_threshold = nms_threshold
        self.score_threshold = score_threshold
        self.max_detections = max_detections
        self.parallel_iterations = parallel_iterations  # kept for config round-trips; the batch runs in one launch set
        self.name = name

    def __call__(self, inputs, **kwargs):
        return self.call(inputs, **kwargs)

    def call(self, inputs, **kwargs):
        """ inputs: [boxes (B,N,4), boxes3D (B,N,16), classification (B,N,C), other[0], ...]."""
        boxes, boxes3D, classification = _cuda(inputs[0]), _cuda(inputs[1]), _cuda(inputs[2])
        other = [_cuda(o) for o in inputs[3:]]
        return _filter_batch(boxes, boxes3D, classification, other, self.class_specific_filter, self.nms,
                             self.score_threshold, self.max_detections, self.nms_threshold)

    def compute_output_shape(self, input_shape):
        return [
            (input_shape[0][0], self.max_detections, 4),
            (input_shape[1][0], self.max_detections, 16),
            (input_shape[2][0], self.max_detections),
            (input_shape[2][0], self.max_detections),
        ] + [tuple([input_shape[i][0], self.max_detections] + list(input_shape[i][3:])) for i in range(3, len(input_shape))]

    def compute_mask(self, inputs, mask=None):
        return (len(inputs) + 1) * [None]

    def get_config(self):
        return {
            'name': self.name,
            'nms': self.nms,
            'class_specific_filter': self.class_specific_filter,
            'nms_threshold': self.nms_threshold,
            'score_threshold': self.score_threshold,
            'max_detections': self.max_detections,
            'parallel_iterations': self.parallel_iterations,
        }
