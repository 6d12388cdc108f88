"""Input staging — the step right before the hot path:

  preprocess_image       PyraPose/utils/image.py:35-62
  compute_resize_scale   PyraPose/utils/image.py:281-305
  resize_image           PyraPose/utils/image.py:308-323   (cv2.resize(img, None, fx=scale, fy=scale), bilinear)
  compute_inputs         PyraPose/preprocessing/generator.py:320-336 (zero-padded batch)

`stage_batch` does the three per-image steps in ONE kernel launch per image on the device (ppb_resize_preprocess_f32):
raw uint8 BGR pixels in, the float32 mean-subtracted, resized, zero-padded batch tensor out — what the reference
builds on the host before predict_on_batch. `resize_image` alone runs the same kernel on a float32 image (the form the
reference resizes: preprocess_image comes first, PyraPose/utils/linemod_eval.py:283-288)."""
import ctypes

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


def compute_resize_scale(image_shape, min_side=480, max_side=640):
    """ Scale such that the smallest side becomes min_side, unless the largest side would then exceed max_side."""
    rows, cols = image_shape[0], image_shape[1]
    scale = min_side / min(rows, cols)
    if max(rows, cols) * scale > max_side:
        scale = max_side / max(rows, cols)
    return scale


def _resized_shape(shape, scale):
    # cv::resize with dsize = None: Size(saturate_cast<int>(cols * fx), saturate_cast<int>(rows * fy)), i.e. cvRound
    return int(np.rint(shape[0] * scale)), int(np.rint(shape[1] * scale))


def _launch(src_dev, is_u8, sh, sw, dst_dev, dh, dw, ph, pw, scale, mean):
    import torch
    from .. import _native as nat
    m = None if mean is None else (ctypes.c_float * 3)(*[float(v) for v in mean])
    with torch.cuda.device(dst_dev.device):
        nat.check(nat.load().ppb_resize_preprocess_f32(nat.ptr(src_dev), int(is_u8), sh, sw, nat.ptr(dst_dev), dh, dw,
                                                       ph, pw, ctypes.c_double(scale), ctypes.c_double(scale), m,
                                                       nat.current_stream_ptr()))


def resize_image(img, min_side=480, max_side=640):
    """ Resize a float32 image such that its size is constrained to min_side / max_side. Returns (image, scale)."""
    import torch
    img = np.asarray(img)
    if img.ndim != 3 or img.shape[2] != 3:
        raise ValueError("expected an image of shape (H, W, 3), got %s" % (img.shape,))
    if img.dtype == np.uint8:
        raise ValueError("resize_image resizes the float32 image preprocess_image returns (as the reference's callers "
                         "do); for raw uint8 images use stage_batch, which fuses preprocess + resize + padding")
    scale = compute_resize_scale(img.shape, min_side=min_side, max_side=max_side)
    dh, dw = _resized_shape(img.shape, scale)
    dev = torch.device("cuda", torch.cuda.current_device())
    src = torch.from_numpy(np.ascontiguousarray(img, dtype=np.float32)).to(dev)
    dst = torch.empty((dh, dw, 3), dtype=torch.float32, device=dev)
    _launch(src, False, img.shape[0], img.shape[1], dst, dh, dw, dh, dw, scale, None)
    return dst.cpu().numpy(), scale


def compute_inputs(image_group, batch_size=None):
    """ Zero batch of the largest shape in the group with every image in its upper-left corner (host, numpy)."""
    max_shape = tuple(max(image.shape[x] for image in image_group) for x in range(3))
    n = len(image_group) if batch_size is None else batch_size
    batch = np.zeros((n,) + max_shape, dtype=np.float32)
    for i, image in enumerate(image_group):
        batch[i, :image.shape[0], :image.shape[1], :image.shape[2]] = image
    return batch


def stage_batch(raw_images, min_side=480, max_side=640, mode='caffe', out=None):
    """ raw_images: list of uint8 (or float32) BGR arrays (H_i, W_i, 3), sizes may differ.
    Returns (batch, scales): batch = CUDA float32 tensor (B, Hmax, Wmax, 3) = compute_inputs([resize_image(
    preprocess_image(img))[0] for img in raw_images]), built on the device from the raw pixels; `out` may name a
    preallocated tensor of that shape (e.g. the engine's bound input)."""
    import torch
    if mode not in ('caffe', None):
        raise ValueError("stage_batch implements mode='caffe' (the reference's resnet backbone) or None")
    scales = [compute_resize_scale(im.shape, min_side=min_side, max_side=max_side) for im in raw_images]
    shapes = [_resized_shape(im.shape, s) for im, s in zip(raw_images, scales)]
    ph, pw = max(s[0] for s in shapes), max(s[1] for s in shapes)
    dev = out.device if out is not None else torch.device("cuda", torch.cuda.current_device())
    if out is None:
        out = torch.empty((len(raw_images), ph, pw, 3), dtype=torch.float32, device=dev)
    elif tuple(out.shape) != (len(raw_images), ph, pw, 3) or out.dtype != torch.float32:
        raise ValueError("out must be a float32 tensor of shape %s" % ((len(raw_images), ph, pw, 3),))
    mean = _CAFFE_MEAN_BGR if mode == 'caffe' else None
    for i, (im, s, (dh, dw)) in enumerate(zip(raw_images, scales, shapes)):
        im = np.asarray(im)
        u8 = im.dtype == np.uint8
        src = torch.from_numpy(np.ascontiguousarray(im if u8 else im.astype(np.float32))).to(dev, non_blocking=True)
        _launch(src, u8, im.shape[0], im.shape[1], out[i], dh, dw, ph, pw, s, mean)
    return out, scales
