"""preprocess_image — the step right before the hot path (PyraPose/utils/image.py:35-62)."""
import numpy as np

_CAFFE_MEAN_BGR = np.array([103.939, 116.779, 123.68], dtype=np.float32)


def preprocess_image(x, mode='caffe'):
    """BGR image -> float32, ImageNet mean removed ('caffe') or scaled to [-1, 1] ('tf')."""
    x = x.astype(np.float32)
    if mode == 'tf':
        x /= 127.5
        x -= 1.
    elif mode == 'caffe':
        x -= _CAFFE_MEAN_BGR
    return x
