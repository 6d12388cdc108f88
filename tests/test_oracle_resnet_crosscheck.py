"""Independent cross-check of the oracle's ResNet-50 restatement (oracle/pfpn.py::resnet50).

keras-resnet 0.1.0 — where the reference's backbone arithmetic lives (PyraPose/models/resnet.py:87) — is not
installable here, so the oracle restates its block structure from knowledge of the package. The HuggingFace
`transformers` ResNet (installed, offline, random-initialised) is a second, unrelated implementation of the same
original Caffe ResNet-50 when configured with `downsample_in_bottleneck=True` (stride on the first 1x1 of a
down-sampling block), bottleneck depths [3, 4, 6, 3], projection shortcuts with BatchNorm, eps = 1e-5, no conv bias.
Loaded with the SAME weights, its stem and its four stages must reproduce the oracle's to fp32 rounding. This pins the
oracle's block wiring (stride placement, shortcut projection, where the ReLUs sit, BN epsilon) to a public
implementation; it does not pin TensorFlow's numerics.

The one place the two differ by design is the stem's max-pool: Keras `MaxPooling2D(3, strides=2, padding='same')`
pads bottom/right only for even inputs, PyTorch's `MaxPool2d(3, 2, padding=1)` pads all four sides — the stages are
therefore fed the ORACLE's pooled tensor, and the difference itself is asserted."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
transformers = pytest.importorskip("transformers")


def _load_conv_bn(hf_layer, w, conv_name, bn_name):
    with torch.no_grad():
        hf_layer.convolution.weight.copy_(torch.from_numpy(w[conv_name + "/kernel"]).permute(3, 2, 0, 1))
        bn = hf_layer.normalization
        bn.weight.copy_(torch.from_numpy(w[bn_name + "/gamma"]))
        bn.bias.copy_(torch.from_numpy(w[bn_name + "/beta"]))
        bn.running_mean.copy_(torch.from_numpy(w[bn_name + "/moving_mean"]))
        bn.running_var.copy_(torch.from_numpy(w[bn_name + "/moving_variance"]))


def _hf_resnet50(w):
    from transformers import ResNetConfig, ResNetModel
    cfg = ResNetConfig(num_channels=3, embedding_size=64, hidden_sizes=[256, 512, 1024, 2048], depths=[3, 4, 6, 3],
                       layer_type="bottleneck", hidden_act="relu", downsample_in_first_stage=False,
                       downsample_in_bottleneck=True)
    m = ResNetModel(cfg).eval()
    for mod in m.modules():
        if isinstance(mod, torch.nn.BatchNorm2d):
            assert mod.eps == 1e-5
    _load_conv_bn(m.embedder.embedder, w, "conv1", "bn_conv1")
    for s, blocks in enumerate([3, 4, 6, 3]):
        for b in range(blocks):
            tag = "%d%s" % (s + 2, chr(ord("a") + b))
            layer = m.encoder.stages[s].layers[b]
            for i, br in enumerate(("2a", "2b", "2c")):
                _load_conv_bn(layer.layer[i], w, "res%s_branch%s" % (tag, br), "bn%s_branch%s" % (tag, br))
            if b == 0:
                _load_conv_bn(layer.shortcut, w, "res%s_branch1" % tag, "bn%s_branch1" % tag)
            else:
                assert isinstance(layer.shortcut, torch.nn.Identity)
    return m


def _rel(a, b):
    return float(torch.linalg.norm((a - b).double()) / torch.linalg.norm(b.double()))


def test_oracle_resnet50_matches_an_independent_implementation():
    from oracle import pfpn
    from pyrapose_b200.models import spec
    w = spec.init_weights(3, seed=5, head_init="keras", randomize_bn=True)
    rng = np.random.default_rng(0)
    img = (rng.integers(0, 256, (2, 96, 136, 3)).astype(np.float32) - np.array([103.939, 116.779, 123.68], np.float32))
    x = torch.from_numpy(img).permute(0, 3, 1, 2).contiguous()
    taps = {}
    with torch.no_grad():
        c2, c3, c4, c5 = pfpn.resnet50(x, w, taps=taps)
        hf = _hf_resnet50(w)
        # stem: ZeroPadding2D(3) + 7x7/2 VALID == Conv2d(7, stride 2, padding 3); BN; ReLU
        stem = hf.embedder.embedder(x)
        assert stem.shape == taps["conv1_relu"].shape and _rel(stem, taps["conv1_relu"]) < 1e-5
        # the max-pool is where Keras 'same' and PyTorch padding=1 differ (even input: Keras pads bottom/right only)
        pooled_torch = hf.embedder.pooler(stem)
        assert pooled_torch.shape == taps["pool1"].shape
        assert not torch.allclose(pooled_torch, taps["pool1"])
        # the four stages on the oracle's pooled tensor
        h = taps["pool1"]
        for s, (want, last) in enumerate(zip((c2, c3, c4, c5), ("res2c_relu", "res3d_relu", "res4f_relu", "res5c_relu"))):
            h = hf.encoder.stages[s](h)
            assert h.shape == want.shape, (s, h.shape, want.shape)
            err = _rel(h, want)
            print("stage %d: relative L2 between the two implementations %.2e" % (s + 2, err))
            assert err < 2e-5, (s, err)
            assert torch.equal(want, taps[last])
    assert tuple(c3.shape[1:]) == (512, 12, 17) and tuple(c5.shape[1:]) == (2048, 3, 5)
