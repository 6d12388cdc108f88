"""Layer objects — mirror of PyraPose/layers/_misc.py. Each is a callable with the reference's constructor
arguments, get_config() and compute_output_shape(); __call__ takes CUDA tensors (numpy is uploaded) and runs
the CUDA kernels behind the C-ABI."""
import numpy as np

from .. import backend
from ..backend.common import _cuda
from ..utils import anchors as utils_anchors


class _Layer(object):
    def __init__(self, name=None, **kwargs):
        self.name = name or self.__class__.__name__.lower()

    def get_config(self):
        return {'name': self.name}

    def __call__(self, inputs, **kwargs):
        return self.call(inputs, **kwargs)


class Anchors(_Layer):
    """ Generates anchors for a given feature map (_misc.py:24-93): float64 base anchors cast to float32, shifted
    in float32, tiled over the batch."""

    def __init__(self, size, stride, ratios=None, scales=None, *args, **kwargs):
        self.size = size
        self.stride = stride
        self.ratios = ratios
        self.scales = scales
        if ratios is None:
            self.ratios = utils_anchors.AnchorParameters.default.ratios
        elif isinstance(ratios, list):
            self.ratios = np.array(ratios)
        if scales is None:
            self.scales = utils_anchors.AnchorParameters.default.scales
        elif isinstance(scales, list):
            self.scales = np.array(scales)
        self.num_anchors = len(self.ratios) * len(self.scales)
        self.anchors = utils_anchors.generate_anchors(base_size=size, ratios=self.ratios,
                                                      scales=self.scales).astype(np.float32)
        super(Anchors, self).__init__(*args, **kwargs)

    def call(self, inputs, **kwargs):
        shape = tuple(inputs.shape)  # (B, H, W, C), channels_last
        anchors = backend.shift(shape[1:3], self.stride, self.anchors)
        return anchors.unsqueeze(0).expand(shape[0], -1, -1).contiguous()

    def compute_output_shape(self, input_shape):
        if None not in input_shape[1:]:
            return (input_shape[0], int(np.prod(input_shape[1:3])) * self.num_anchors, 4)
        return (input_shape[0], None, 4)

    def get_config(self):
        config = super(Anchors, self).get_config()
        config.update({'size': self.size, 'stride': self.stride, 'ratios': self.ratios.tolist(),
                       'scales': self.scales.tolist()})
        return config


class UpsampleLike(_Layer):
    """ Nearest-neighbour resize of `source` to the spatial size of `target` (_misc.py:96-115). On the hot path
    this is fused into the PFPN add kernel; the layer form is its own kernel (ppb_upsample_like_f32)."""

    def call(self, inputs, **kwargs):
        from .. import _kernels
        source, target = inputs
        return _kernels.upsample_like(_cuda(source), int(target.shape[1]), int(target.shape[2]))

    def compute_output_shape(self, input_shape):
        return (input_shape[0][0],) + tuple(input_shape[1][1:3]) + (input_shape[0][-1],)


def _mean_std(mean, std, n):
    if mean is None:
        mean = np.zeros(n)
    if std is None:
        std = np.full(n, 0.2)
    if isinstance(mean, (list, tuple)):
        mean = np.array(mean)
    elif not isinstance(mean, np.ndarray):
        raise ValueError('Expected mean to be a np.ndarray, list or tuple. Received: {}'.format(type(mean)))
    if isinstance(std, (list, tuple)):
        std = np.array(std)
    elif not isinstance(std, np.ndarray):
        raise ValueError('Expected std to be a np.ndarray, list or tuple. Received: {}'.format(type(std)))
    return mean, std


class _Regress(_Layer):
    _values = 4

    def __init__(self, mean=None, std=None, *args, **kwargs):
        self.mean, self.std = _mean_std(mean, std, self._values)
        super(_Regress, self).__init__(*args, **kwargs)

    def get_config(self):
        config = super(_Regress, self).get_config()
        config.update({'mean': self.mean.tolist(), 'std': self.std.tolist()})
        return config


class RegressBoxes(_Regress):
    """ Applies regression values to boxes (_misc.py:118-162)."""
    _values = 4

    def call(self, inputs, **kwargs):
        anchors, regression = inputs
        return backend.bbox_transform_inv(anchors, regression, mean=self.mean, std=self.std)

    def compute_output_shape(self, input_shape):
        return input_shape[0]


class RegressBoxes3D(_Regress):
    """ Applies the 16 corner regression values to anchors (_misc.py:165-209)."""
    _values = 16

    def call(self, inputs, **kwargs):
        anchors, regression = inputs
        return backend.box3D_transform_inv(anchors, regression, mean=self.mean, std=self.std)

    def compute_output_shape(self, input_shape):
        return input_shape[1]


class DenormBoxes3D(_Regress):
    """ RegressBoxes3D followed by x / 320, y / 240 (_misc.py:212-256)."""
    _values = 16

    def call(self, inputs, **kwargs):
        anchors, regression = inputs
        return backend.box3D_denormalization(anchors, regression, mean=self.mean, std=self.std)

    def compute_output_shape(self, input_shape):
        return input_shape[1]


class ClipBoxes(_Layer):
    """ Clips boxes to the image: x to [0, width], y to [0, height] (_misc.py:259-280)."""

    def call(self, inputs, **kwargs):
        import torch
        from .. import _kernels
        image, boxes = inputs
        shape = tuple(image.shape)  # channels_last (B, H, W, C)
        return _kernels.clip_boxes(_cuda(boxes, torch.float32), float(shape[2]), float(shape[1]))

    def compute_output_shape(self, input_shape):
        return input_shape[1]
