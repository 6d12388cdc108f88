"""The reference's LAYER objects called as a PyraPose maintainer would call them (PyraPose/layers/_misc.py,
layers/filter_detections.py): same constructor arguments, tensors in, tensors out, against the CPU oracle. The kernels
underneath are also covered one by one in test_kernels_gpu.py; these go through the layer classes."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _np(t):
    return t.detach().cpu().numpy()


def test_anchors_layer_matches_reference_arithmetic(cuda_device):
    """Anchors(size, stride)(features): float64 base anchors cast to float32, float32 shifts (_misc.py:24-93)."""
    from oracle import anchors as oa
    from pyrapose_b200 import layers
    import torch
    for (fh, fw), size, stride in (((60, 80), 32, 8), ((30, 40), 64, 16), ((15, 20), 128, 32), ((17, 25), 32, 8)):
        feat = torch.zeros((2, fh, fw, 256), device=cuda_device)
        layer = layers.Anchors(size=size, stride=stride)
        got = _np(layer(feat))
        # the layer's own arithmetic is float32 throughout (backend/common.py:93-116); the oracle restates it
        want = oa.anchors_layer_f32([(fh, fw)], oa.AnchorParameters([size], [stride], oa.AnchorParameters.default.ratios,
                                                                      oa.AnchorParameters.default.scales))
        assert got.shape == (2, fh * fw * 9, 4) and got.dtype == np.float32
        assert np.array_equal(got[0], want) and np.array_equal(got[1], want)
        assert layer.compute_output_shape((2, fh, fw, 256)) == (2, fh * fw * 9, 4)
        cfg = layer.get_config()
        assert layers.Anchors(**{k: v for k, v in cfg.items() if k != 'name'}).num_anchors == 9


def test_upsample_like_layer(cuda_device):
    """UpsampleLike()([source, target]) (_misc.py:96-115): nearest resize to the target's H x W; exact copies."""
    import torch
    from oracle import pfpn
    from pyrapose_b200 import layers
    rng = np.random.default_rng(0)
    for (ih, iw, oh, ow, c) in ((15, 20, 30, 40, 256), (30, 40, 60, 80, 256), (9, 13, 17, 25, 6), (4, 4, 4, 4, 3)):
        src = rng.standard_normal((2, ih, iw, c)).astype(np.float32)
        tgt = np.zeros((2, oh, ow, 8), np.float32)
        got = _np(layers.UpsampleLike()([torch.from_numpy(src).to(cuda_device), torch.from_numpy(tgt).to(cuda_device)]))
        want = pfpn.upsample_like(torch.from_numpy(src).permute(0, 3, 1, 2), torch.from_numpy(tgt).permute(0, 3, 1, 2))
        want = want.permute(0, 2, 3, 1).numpy()
        assert got.shape == (2, oh, ow, c) and np.array_equal(got, want)
    assert layers.UpsampleLike().compute_output_shape([(2, 15, 20, 256), (2, 30, 40, 8)]) == (2, 30, 40, 256)


def test_regress_layers_match_oracle(cuda_device):
    """RegressBoxes / RegressBoxes3D / DenormBoxes3D / ClipBoxes through the layer objects, bit for bit."""
    import torch
    from oracle import anchors as oa, decode as od
    from pyrapose_b200 import layers
    rng = np.random.default_rng(1)
    anchors = oa.anchors_layer_f32([(15, 20), (8, 10)])
    B, N = 3, anchors.shape[0]
    a = np.broadcast_to(anchors, (B, N, 4)).copy()
    d4 = rng.standard_normal((B, N, 4)).astype(np.float32)
    d16 = rng.standard_normal((B, N, 16)).astype(np.float32)
    a_t, d4_t, d16_t = (torch.from_numpy(x).to(cuda_device) for x in (a, d4, d16))
    assert np.array_equal(_np(layers.RegressBoxes()([a_t, d4_t])), od.bbox_transform_inv(a, d4))
    assert np.array_equal(_np(layers.RegressBoxes3D()([a_t, d16_t])), od.box3D_transform_inv(a, d16))
    mean, std = np.linspace(-0.1, 0.1, 16), np.linspace(0.1, 0.3, 16)
    assert np.array_equal(_np(layers.RegressBoxes3D(mean=mean, std=list(std))([a_t, d16_t])),
                          od.box3D_transform_inv(a, d16, mean=mean, std=std))
    assert np.array_equal(_np(layers.DenormBoxes3D()([a_t, d16_t])), od.box3D_denormalization(a, d16))
    boxes = od.bbox_transform_inv(a, d4) * 4.0
    img = torch.zeros((B, 120, 160, 3), device=cuda_device)
    assert np.array_equal(_np(layers.ClipBoxes()([img, torch.from_numpy(boxes).to(cuda_device)])),
                          od.clip_boxes((120, 160), boxes))
    with pytest.raises(ValueError):
        layers.RegressBoxes3D(mean=0.0)


def test_filter_detections_layer_with_other(cuda_device):
    """FilterDetections(...)([boxes, boxes3D, classification, other]) (filter_detections.py:121-234)."""
    import torch
    from oracle import decode as od, filter as of
    from pyrapose_b200 import layers
    rng = np.random.default_rng(2)
    B, N, C = 2, 4000, 5
    centers = rng.uniform(20, 300, (B, N, 2)).astype(np.float32)
    wh = rng.uniform(8, 60, (B, N, 2)).astype(np.float32)
    boxes = np.concatenate([centers - wh / 2, centers + wh / 2], axis=2)
    boxes3D = rng.standard_normal((B, N, 16)).astype(np.float32)
    scores = (1.0 / (1.0 + np.exp(-(rng.standard_normal((B, N, C)) * 1.5 - 3.0)))).astype(np.float32)
    other = rng.standard_normal((B, N, 3)).astype(np.float32)
    for kw in (dict(), dict(class_specific_filter=False), dict(nms=False, max_detections=50), dict(score_threshold=0.3, nms_threshold=0.3)):
        layer = layers.FilterDetections(**kw)
        got = [_np(t) for t in layer([torch.from_numpy(x).to(cuda_device) for x in (boxes, boxes3D, scores, other)])]
        want = of.filter_detections_batch(boxes, boxes3D, scores, **kw)
        for g, w in zip(got[:4], want):
            assert g.dtype == w.dtype and np.array_equal(g, w)
        # `other` is gathered with the same indices: rows of kept detections equal the source rows, padding is -1
        D = layer.max_detections
        assert got[4].shape == (B, D, 3)
        for b in range(B):
            k = int((want[3][b] >= 0).sum())
            assert np.all(got[4][b, k:] == -1)
        shapes = layer.compute_output_shape([(B, N, 4), (B, N, 16), (B, N, C), (B, N, 3)])
        # filter_detections.py:209 slices the other inputs' shapes with [3:], so a (B, N, 3) input is announced as
        # (B, D): the mirror keeps the reference's answer
        assert shapes[0] == (B, D, 4) and shapes[4] == (B, D)
        assert layers.FilterDetections(**layer.get_config()).get_config() == layer.get_config()
