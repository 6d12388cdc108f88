"""Architecture of the PFPN network as a flat list of convolution layers (pure Python, no CUDA).

Names follow the Keras layer names a reference checkpoint would carry: keras-resnet 0.1.0 for the
backbone (`conv1`, `bn_conv1`, `res{2-5}{a-f}_branch{1,2a,2b,2c}`, `bn...`; PyraPose/models/resnet.py:87),
'P3'/'P4'/'P5' for the named pyramid outputs (PyraPose/models/retinanet.py:204,209,212). Layers the
reference leaves auto-numbered (`conv2d_N`) get stable descriptive names here; the head names are the
ones commented out in retinanet.py:33,44,119,126.
"""
from collections import namedtuple

import numpy as np

from .. import initializers

ConvSpec = namedtuple("ConvSpec", "name ksize cin cout stride use_bias bn init")

RESNET50_BLOCKS = [3, 4, 6, 3]


def resnet50_layers():
    """keras_resnet.models.ResNet50(include_top=False, freeze_bn=True): Caffe-style bottlenecks with the
    stride on the first 1x1 conv, no conv bias, BatchNorm(eps=1e-5) after every conv."""
    layers = [ConvSpec("conv1", 7, 3, 64, 2, False, "bn_conv1", "he")]
    cin = 64
    features = 64
    for stage, blocks in enumerate(RESNET50_BLOCKS):
        for block in range(blocks):
            tag = "%d%s" % (stage + 2, chr(ord("a") + block))
            stride = 2 if (block == 0 and stage > 0) else 1
            layers.append(ConvSpec("res%s_branch2a" % tag, 1, cin, features, stride, False, "bn%s_branch2a" % tag, "he"))
            layers.append(ConvSpec("res%s_branch2b" % tag, 3, features, features, 1, False, "bn%s_branch2b" % tag, "he"))
            layers.append(ConvSpec("res%s_branch2c" % tag, 1, features, features * 4, 1, False, "bn%s_branch2c" % tag, "he"))
            if block == 0:
                layers.append(ConvSpec("res%s_branch1" % tag, 1, cin, features * 4, stride, False, "bn%s_branch1" % tag, "he"))
            cin = features * 4
        features *= 2
    return layers


def pfpn_layers(feature_size=256):
    """__create_sparceFPN (PyraPose/models/retinanet.py:180-214). All convs: bias, 'same', linear."""
    f = feature_size
    return [
        ConvSpec("P3_lateral", 1, 512, f, 1, True, None, "glorot"),    # :192
        ConvSpec("P4_lateral", 1, 1024, f, 1, True, None, "glorot"),   # :193
        ConvSpec("P5_lateral", 1, 2048, f, 1, True, None, "glorot"),   # :194
        ConvSpec("P4_mid", 3, f, f, 1, True, None, "glorot"),          # :199
        ConvSpec("P3_mid", 3, f, f, 1, True, None, "glorot"),          # :201
        ConvSpec("P3_down", 3, f, f, 2, True, None, "glorot"),         # :202
        ConvSpec("P3", 3, f, f, 1, True, None, "glorot"),              # :204
        ConvSpec("P4_down", 3, f, f, 2, True, None, "glorot"),         # :207
        ConvSpec("P4", 3, f, f, 1, True, None, "glorot"),              # :209
        ConvSpec("P5", 3, f, f, 1, True, None, "glorot"),              # :212
    ]


def head_layers(num_classes, num_anchors, feature_size=256, regression_feature_size=512,
                classification_feature_size=256):
    """default_3Dregression_model (:101-131), default_classification_model (:9-54), default_mask_model (:57-98)."""
    layers = []
    cin = feature_size
    for i in range(4):
        layers.append(ConvSpec("pyramid_regression3D_%d" % i, 3, cin, regression_feature_size, 1, True, None, "normal"))
        cin = regression_feature_size
    layers.append(ConvSpec("pyramid_regression3D", 3, cin, num_anchors * 16, 1, True, None, "normal"))
    cin = feature_size
    for i in range(4):
        layers.append(ConvSpec("pyramid_classification_%d" % i, 3, cin, classification_feature_size, 1, True, None, "normal"))
        cin = classification_feature_size
    layers.append(ConvSpec("pyramid_classification", 3, cin, num_classes * num_anchors, 1, True, None, "prior"))
    cin = feature_size
    for i in range(4):
        layers.append(ConvSpec("mask_%d" % i, 3, cin, classification_feature_size, 1, True, None, "normal"))
        cin = classification_feature_size
    layers.append(ConvSpec("mask_out", 3, cin, num_classes, 1, True, None, "prior"))
    return layers


def all_layers(num_classes, num_anchors=9):
    return resnet50_layers() + pfpn_layers() + head_layers(num_classes, num_anchors)


def init_weights(num_classes, num_anchors=9, seed=1, head_init="keras", randomize_bn=False):
    """Random-init weights shaped and distributed like a freshly built reference model.

    Returns {"<layer>/kernel": HWIO f32, "<layer>/bias": f32, "<bn>/gamma|beta|moving_mean|moving_variance": f32}.
    head_init="keras": N(0, 0.01) kernels and PriorProbability(0.01) output bias, exactly the reference's
    initialisers; "spread": he-normal head kernels so that random-init scores and deltas are not all ~0.01 / ~0
    (used by parity tests to exercise the decode and filter paths with non-degenerate values).
    randomize_bn=True draws non-trivial frozen BatchNorm statistics so that BN folding is exercised.
    """
    rng = np.random.default_rng(seed)
    w = {}
    for l in all_layers(num_classes, num_anchors):
        shape = (l.ksize, l.ksize, l.cin, l.cout)
        if l.init == "he":
            k = initializers.he_normal(rng, shape)
            if l.name == "conv1":
                k = k * np.float32(1.0 / 64.0)  # inputs are mean-subtracted 8-bit pixels (std ~70)
        elif l.init == "glorot":
            k = initializers.glorot_uniform(rng, shape)
        elif head_init == "spread":
            k = initializers.he_normal(rng, shape)
            if l.cout not in (256, 512):  # the three output convs: keep logits / deltas O(1)
                k = k * np.float32(0.1)
        else:
            k = initializers.normal(rng, shape, 0.0, 0.01)
        w[l.name + "/kernel"] = k
        if l.use_bias:
            if l.init == "prior":
                w[l.name + "/bias"] = initializers.PriorProbability(0.01)((l.cout,), np.float32).astype(np.float32)
            else:
                w[l.name + "/bias"] = np.zeros((l.cout,), np.float32)
        if l.bn is not None:
            # The reference never runs a random-init backbone (it loads the ImageNet checkpoint,
            # resnet.py:89-98). A small gain on each block's last BatchNorm keeps the synthetic
            # backbone's activations O(1) through the 16 residual additions, like a trained one.
            gain = np.float32(0.25 if l.name.endswith("branch2c") else 1.0)
            if randomize_bn:
                w[l.bn + "/gamma"] = rng.uniform(0.8, 1.2, l.cout).astype(np.float32) * gain
                w[l.bn + "/beta"] = rng.normal(0, 0.1, l.cout).astype(np.float32)
                w[l.bn + "/moving_mean"] = rng.normal(0, 0.1, l.cout).astype(np.float32)
                w[l.bn + "/moving_variance"] = rng.uniform(0.8, 1.2, l.cout).astype(np.float32)
            else:
                w[l.bn + "/gamma"] = np.ones(l.cout, np.float32) * gain
                w[l.bn + "/beta"] = np.zeros(l.cout, np.float32)
                w[l.bn + "/moving_mean"] = np.zeros(l.cout, np.float32)
                w[l.bn + "/moving_variance"] = np.ones(l.cout, np.float32)
    return w
