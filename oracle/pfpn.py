"""fp32 CPU restatement (torch-CPU convolutions) of the reference's PFPN graph — TEST INFRASTRUCTURE ONLY.

WIRING PINNED TO THE REFERENCE'S TEXT, NUMERICS OF THE LIBRARIES UNPINNED. TensorFlow 2.1.1 / Keras 2.3.1 /
keras-resnet 0.1.0 are not installable here and the reference has no tests. What IS done: the reference's own
model-building code (models/__init__.py, models/resnet.py, models/retinanet.py, initializers.py, layers/_misc.py) is
imported and run in the build container over a miniature Keras (tests/golden/keras_shim.py), loaded with the same
weights through the Keras layer-name map, and its outputs are committed (tests/golden/reference_model.npz,
make_golden_model.py); this file reproduces them — C3-C5, P3-P5, 3Dbox, cls, mask, boxes3D — to <= 2.5e-6 relative L2
(tests/test_oracle_reference_model.py). That fixes the graph (who feeds whom, strides, paddings, activations, shared
heads, Reshape / Concatenate order). The arithmetic of Conv2D / 'same' padding / MaxPooling2D / BatchNormalization and
keras-resnet's ResNet50 are restatements on both sides. This follows the model-building code line by line:
  backbone : keras_resnet.models.ResNet50(include_top=False, freeze_bn=True) as called at
             /root/reference/PyraPose/models/resnet.py:87,110 (architecture restated from keras-resnet 0.1.0
             models/_2d.py::ResNet50 + blocks/_2d.py::bottleneck_2d: ZeroPad(3)+7x7/2 VALID, BN(eps=1e-5), ReLU,
             MaxPool 3x3/2 'same'; bottlenecks [3,4,6,3], stride on the first 1x1, ZeroPad(1)+3x3 VALID, no bias)
  PFPN     : /root/reference/PyraPose/models/retinanet.py:180-214  (__create_sparceFPN)
  heads    : retinanet.py:9-54 (cls), 57-98 (mask), 101-131 (3D regression); pyramid concat :224-225
  inference: retinanet.py:302-335 (retinanet_bbox): Anchors + RegressBoxes3D
The backbone restatement is cross-checked against an unrelated public implementation of the same Caffe ResNet-50
(HuggingFace transformers ResNet, downsample_in_bottleneck=True, same weights): tests/test_oracle_resnet_crosscheck.py.
Weights come in as a dict keyed "<keras layer name>/<param>" with HWIO kernels (pyrapose_b200.models.spec
names); this file never creates weights.
"""
import numpy as np
import torch
import torch.nn.functional as F

from . import anchors as oanchors
from . import decode as odecode

BN_EPS = 1e-5  # keras-resnet BatchNormalization(epsilon=1e-5)


# Where the oracle's torch ops run. None = CPU (the default, and the only mode of the CPU test suite). A GPU test may
# set it to a CUDA device to execute THE SAME fp32 graph through torch's library convolutions (TF32 off) at sizes the
# CPU cannot afford — BASELINE config 2 at its full batch of 32 (tests/test_engine_gpu.py); still test infrastructure.
DEVICE = [None]


def _t(a):
    if isinstance(a, torch.Tensor):  # oracle/train.py passes leaf tensors so that autograd supplies gradients
        return a
    t = torch.from_numpy(np.ascontiguousarray(a))
    return t.to(DEVICE[0]) if DEVICE[0] is not None else t


# Rounding hook applied wherever the native engine stores an activation as bf16. Identity for the fp32 oracle;
# oracle/train.py swaps in a bf16 round (straight-through, gradient rounded too) for its "bf16-emulating" mode, which
# checks the backward arithmetic independently of ReLU-mask flips between an fp32 and a bf16 forward pass.
_Q = [lambda x, name: x, lambda z, name: torch.relu(z)]


def _q(x, name=None):
    """A stored linear activation (convolution or Add output); `name` is the engine's tensor name."""
    return _Q[0](x, name)


def _qrelu(z, name=None):
    """A stored ReLU activation."""
    return _Q[0](_Q[1](z, name), name) if _Q[1] is not None else _Q[0](z, name)


def _kernel(w, name):
    return _t(w[name + "/kernel"]).permute(3, 2, 0, 1).contiguous()  # HWIO -> OIHW


def tf_same_pad(size, k, stride):
    """TensorFlow 'SAME': out = ceil(in/stride); total = max((out-1)*stride + k - in, 0); the extra pixel goes
    to the bottom/right."""
    out = -(-size // stride)
    total = max((out - 1) * stride + k - size, 0)
    return total // 2, total - total // 2


def conv_same(x, w, name, stride=1):
    """keras.layers.Conv2D(padding='same') with bias, NCHW tensors."""
    k = _kernel(w, name)
    kh, kw = k.shape[2], k.shape[3]
    pt, pb = tf_same_pad(x.shape[2], kh, stride)
    pl, pr = tf_same_pad(x.shape[3], kw, stride)
    x = F.pad(x, (pl, pr, pt, pb))
    b = _t(w[name + "/bias"]) if (name + "/bias") in w else None
    return F.conv2d(x, k, b, stride=stride)


def conv_valid_nobias(x, w, name, stride=1, pad=0):
    """keras-resnet: explicit ZeroPadding2D(pad) then Conv2D(padding='valid', use_bias=False)."""
    if pad:
        x = F.pad(x, (pad, pad, pad, pad))
    return F.conv2d(x, _kernel(w, name), None, stride=stride)


def frozen_bn(x, w, name):
    g = _t(w[name + "/gamma"]).view(1, -1, 1, 1)
    b = _t(w[name + "/beta"]).view(1, -1, 1, 1)
    m = _t(w[name + "/moving_mean"]).view(1, -1, 1, 1)
    v = _t(w[name + "/moving_variance"]).view(1, -1, 1, 1)
    return (x - m) / torch.sqrt(v + BN_EPS) * g + b


def maxpool_same_3x3_s2(x):
    pt, pb = tf_same_pad(x.shape[2], 3, 2)
    pl, pr = tf_same_pad(x.shape[3], 3, 2)
    x = F.pad(x, (pl, pr, pt, pb), value=float("-inf"))
    return F.max_pool2d(x, 3, 2)


def upsample_like(src, target):
    """UpsampleLike (layers/_misc.py:96-115) -> tf.image.resize(..., NEAREST) in TF 2.1 (half_pixel_centers):
    src index = min(floor((dst + 0.5) * in / out), in - 1)."""
    ih, iw = src.shape[2], src.shape[3]
    oh, ow = target.shape[2], target.shape[3]
    ys = torch.clamp(torch.floor((torch.arange(oh, dtype=torch.float32) + 0.5) * (ih / oh)).long(), max=ih - 1).to(src.device)
    xs = torch.clamp(torch.floor((torch.arange(ow, dtype=torch.float32) + 0.5) * (iw / ow)).long(), max=iw - 1).to(src.device)
    return src[:, :, ys][:, :, :, xs]


def resnet50(x, w, taps=None):
    """Returns [C2, C3, C4, C5] (NCHW)."""
    x = conv_valid_nobias(x, w, "conv1", stride=2, pad=3)
    x = _qrelu(frozen_bn(x, w, "bn_conv1"), "conv1_relu")
    if taps is not None:
        taps["conv1_relu"] = x
    x = maxpool_same_3x3_s2(x)
    if taps is not None:
        taps["pool1"] = x
    outs = []
    for stage, blocks in enumerate([3, 4, 6, 3]):
        for block in range(blocks):
            tag = "%d%s" % (stage + 2, chr(ord("a") + block))
            stride = 2 if (block == 0 and stage > 0) else 1
            y = conv_valid_nobias(x, w, "res%s_branch2a" % tag, stride=stride)
            y = _qrelu(frozen_bn(y, w, "bn%s_branch2a" % tag), "res%s_branch2a_relu" % tag)
            y = conv_valid_nobias(y, w, "res%s_branch2b" % tag, pad=1)
            y = _qrelu(frozen_bn(y, w, "bn%s_branch2b" % tag), "res%s_branch2b_relu" % tag)
            y = conv_valid_nobias(y, w, "res%s_branch2c" % tag)
            y = frozen_bn(y, w, "bn%s_branch2c" % tag)
            if block == 0:
                sc = conv_valid_nobias(x, w, "res%s_branch1" % tag, stride=stride)
                sc = _q(frozen_bn(sc, w, "bn%s_branch1" % tag), "res%s_branch1_bn" % tag)
            else:
                sc = x
            x = _qrelu(y + sc, "res%s_relu" % tag)
            if taps is not None:
                taps["res%s_relu" % tag] = x
        outs.append(x)
    return outs


def sparce_fpn(C3, C4, C5, w, taps=None):
    """retinanet.py:180-214."""
    P3 = _q(conv_same(C3, w, "P3_lateral"), "P3_lateral")                        # :192
    P4 = _q(conv_same(C4, w, "P4_lateral"), "P4_lateral")                        # :193
    P5 = _q(conv_same(C5, w, "P5_lateral"), "P5_lateral")                        # :194
    P5_up = upsample_like(P5, P4)                                                # :196
    P4_up = upsample_like(P4, P3)                                                # :197
    P4_mid = _q(conv_same(_q(P5_up + P4, "P4_mid_in"), w, "P4_mid"), "P4_mid")   # :198-199
    P3_mid = _q(conv_same(_q(P4_up + P3, "P3_mid_in"), w, "P3_mid"), "P3_mid")   # :200-201
    P3_down = _q(conv_same(P3_mid, w, "P3_down", stride=2), "P3_down")           # :202
    P3_out = _q(conv_same(_q(P3_mid + P3, "P3_fin"), w, "P3"), "P3")             # :203-204
    P4_fin = P3_down + P4_mid                                                    # :206
    P4_down = _q(conv_same(P4_mid, w, "P4_down", stride=2), "P4_down")           # :207
    P4_out = _q(conv_same(_q(P4_fin + P4, "P4_fin"), w, "P4"), "P4")             # :208-209
    P5_out = _q(conv_same(_q(P4_down + P5, "P5_fin"), w, "P5"), "P5")            # :211-212
    if taps is not None:
        taps.update({"P3_lateral": P3, "P4_lateral": P4, "P5_lateral": P5, "P4_mid": P4_mid, "P3_mid": P3_mid,
                     "P3_down": P3_down, "P4_down": P4_down, "P3": P3_out, "P4": P4_out, "P5": P5_out})
    return [P3_out, P4_out, P5_out]


def _tower(x, w, prefix, out_name, tname=None):
    for i in range(4):
        x = _qrelu(conv_same(x, w, "%s_%d" % (prefix, i)), None if tname is None else tname % i)
    return conv_same(x, w, out_name)


def regression_head(feat, w, level=None):
    """retinanet.py:101-131: NHWC Reshape((-1, 16)). `level` only names the engine tensors for the rounding hooks."""
    y = _tower(feat, w, "pyramid_regression3D", "pyramid_regression3D",
               None if level is None else "reg_tower_t%%d@%d" % level)
    return y.permute(0, 2, 3, 1).reshape(y.shape[0], -1, 16)


def classification_head(feat, w, num_classes, level=None):
    """retinanet.py:9-54: Reshape((-1, C)) then sigmoid."""
    y = _tower(feat, w, "pyramid_classification", "pyramid_classification",
               None if level is None else "cls_tower_t%%d@%d" % level)
    return torch.sigmoid(y.permute(0, 2, 3, 1).reshape(y.shape[0], -1, num_classes))


def mask_head(feat, w, num_classes, level=None):
    """retinanet.py:57-98."""
    y = _tower(feat, w, "mask", "mask_out", None if level is None else "mask_tower_t%%d@%d" % level)
    return torch.sigmoid(y.permute(0, 2, 3, 1).reshape(y.shape[0], -1, num_classes))


def forward_training(images_nhwc, w, num_classes, taps=None, num_threads=None):
    """retinanet() (retinanet.py:260-299): returns numpy [3Dbox (B,N,16), cls (B,N,C), mask (B,HW/64,C)]."""
    if num_threads:
        torch.set_num_threads(num_threads)
    with torch.no_grad():
        x = _t(np.asarray(images_nhwc, np.float32)).permute(0, 3, 1, 2).contiguous()
        _, C3, C4, C5 = resnet50(x, w, taps)
        feats = sparce_fpn(C3, C4, C5, w, taps)
        reg = torch.cat([regression_head(f, w) for f in feats], dim=1)                     # :224-225
        cls = torch.cat([classification_head(f, w, num_classes) for f in feats], dim=1)
        msk = mask_head(feats[0], w, num_classes)                                           # :296-297
        shapes = [(int(f.shape[2]), int(f.shape[3])) for f in feats]
    return reg.cpu().numpy(), cls.cpu().numpy(), msk.cpu().numpy(), shapes


def forward_inference(images_nhwc, w, num_classes, anchor_params=None, taps=None, num_threads=None):
    """retinanet_bbox() (retinanet.py:302-335): [boxes3D (B,N,16), classification (B,N,C), mask]."""
    reg, cls, msk, shapes = forward_training(images_nhwc, w, num_classes, taps, num_threads)
    anchors = oanchors.anchors_layer_f32(shapes, anchor_params)
    anchors = np.broadcast_to(anchors[None], (reg.shape[0],) + anchors.shape)
    boxes3D = odecode.box3D_transform_inv(anchors, reg)
    return boxes3D, cls, msk


def taps_to_nhwc(taps):
    return {k: v.permute(0, 2, 3, 1).contiguous().cpu().numpy() for k, v in taps.items()}
