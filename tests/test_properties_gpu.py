"""Size-independent properties of the bandwidth-class kernels at the FULL sizes of BASELINE.json (configs 2 and 3),
where the CPU oracle would take minutes: symmetry and self-overlap of the IoU, agreement of the fused assignment with the
materialised matrix, the encode -> decode round trip of the 3D-box transform, consistency of the score pass with a
library reduction, and the fixed-point property of FilterDetections. Reference formulas:
PyraPose/utils/compute_overlap.pyx:13-53, PyraPose/utils/anchors.py:290-318, 515-559, PyraPose/backend/common.py:25-56,
PyraPose/layers/filter_detections.py:21-118."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _gt(seed, k, w=640, h=480):
    rng = np.random.default_rng(seed)
    x1 = rng.uniform(0, w - 16, k)
    y1 = rng.uniform(0, h - 16, k)
    bw = rng.uniform(16, 256, k)
    bh = rng.uniform(16, 256, k)
    return np.stack([x1, y1, np.minimum(x1 + bw, w), np.minimum(y1 + bh, h)], axis=1)


def test_overlap_full_size_symmetry_self_and_fused_assignment(cuda_device):
    """Config 3: 64 images x 56 700 anchors x 100 boxes (2.9 GB matrix)."""
    import torch
    from pyrapose_b200 import _kernels as K
    from pyrapose_b200.utils import anchors as ua
    B, Kq = 64, 100
    anchors = ua.anchors_for_shape_device((480, 640, 3))           # f64 [N, 4] on the device
    N = anchors.shape[0]
    assert N == 56700
    q = torch.from_numpy(np.stack([_gt(1000 + i, Kq) for i in range(B)])).to(cuda_device)
    m = K.compute_overlap(anchors, q)                               # [B, N, K]
    assert m.shape == (B, N, Kq) and m.dtype == torch.float64
    assert float(m.min()) >= 0.0 and float(m.max()) <= 1.0
    # symmetry, bit for bit: IoU(a, q) == IoU(q, a) (the two areas commute in the union) on one image
    mt = K.compute_overlap(q[3].contiguous(), anchors)              # [K, N]
    assert torch.equal(m[3], mt.t())
    # a box overlaps itself exactly 1.0
    self_iou = K.compute_overlap(anchors[:4096].contiguous(), anchors[:4096].contiguous())
    assert torch.equal(torch.diagonal(self_iou), torch.ones(4096, dtype=torch.float64, device=cuda_device))
    # fused assignment == numpy lines of compute_gt_annotations on the materialised matrix (first maximum wins)
    positive, ignore, argmax, mx = K.gt_assign(anchors, q, want_max=True)
    ref_max, ref_arg = m.max(dim=2)
    assert torch.equal(mx, ref_max)
    # torch.max does not promise the FIRST maximal index: check that ours is the first one attaining the maximum
    picked = torch.gather(m, 2, argmax.unsqueeze(2)).squeeze(2)
    assert torch.equal(picked, ref_max)
    first = (m == ref_max.unsqueeze(2)).to(torch.uint8).argmax(dim=2)
    assert torch.equal(argmax, first)
    assert torch.equal(positive.bool(), ref_max >= 0.5)
    assert torch.equal(ignore.bool(), (ref_max > 0.4) & ~(ref_max >= 0.5))
    del m


def test_box3d_encode_decode_round_trip_full_size(cuda_device):
    """box3D_transform (anchors.py:515-559, numpy, pinned by the goldens) followed by the device
    box3D_transform_inv (common.py:25-56) returns the projected corners: (gt - a) / w / std * std * w + a."""
    import torch
    from oracle import anchors as oa
    from pyrapose_b200 import _kernels as K
    B = 32
    anchors = oa.anchors_for_shape((480, 640, 3))                    # f64 [56700, 4]
    N = anchors.shape[0]
    rng = np.random.default_rng(11)
    gt16 = rng.uniform(-50, 700, (N, 16))
    targets = oa.box3D_transform(anchors, gt16)                      # f64, mean 0 / std 0.2
    a32 = torch.from_numpy(anchors.astype(np.float32)).to(cuda_device)
    d = torch.from_numpy(targets.astype(np.float32)).to(cuda_device).unsqueeze(0).repeat(B, 1, 1).contiguous()
    out = K.box3d_transform_inv(a32, d)
    ref = torch.from_numpy(gt16.astype(np.float32)).to(cuda_device)
    err = (out - ref.unsqueeze(0)).abs().max()
    assert float(err) < 2e-3, float(err)                             # fp32 round-off on ~700 px coordinates
    assert torch.equal(out[0], out[B - 1])                           # the same anchors for every image of the batch


def test_score_pass_full_size_counts(cuda_device):
    """Config 2 score tensor (32 x 56 700 x 15): popcount of the mask == fused per-class counts == library reduction."""
    import torch
    from pyrapose_b200 import _kernels as K
    g = torch.Generator(device=cuda_device).manual_seed(5)
    sc = torch.sigmoid(torch.randn(32, 56700, 15, generator=g, device=cuda_device) * 1.5 - 4.6)
    for thr in (0.05, 0.5):
        mask, counts = K.score_mask(sc, thr)
        ref = (sc > thr).sum(dim=1).to(torch.int32)
        assert torch.equal(counts, ref)
        words = mask.cpu().numpy().view(np.uint32)
        pop = np.unpackbits(words.view(np.uint8), axis=-1).reshape(32, 15, -1).sum(-1)
        assert np.array_equal(pop, ref.cpu().numpy())


def test_filter_detections_is_a_fixed_point(cuda_device):
    """Feeding the kept detections back through the same filter keeps every one of them, in the same order
    (greedy NMS survivors do not suppress each other; top-k of a sorted list is the list)."""
    import torch
    from pyrapose_b200 import _kernels as K
    rng = np.random.default_rng(21)
    B, N, C = 4, 56700, 15
    cls = (1.0 / (1.0 + np.exp(-rng.normal(-6.0, 1.5, (B, N, C))))).astype(np.float32)
    ctr = rng.uniform(0, 640, (B, N, 2))
    wh = rng.uniform(8, 160, (B, N, 2))
    boxes = np.concatenate([ctr - wh / 2, ctr + wh / 2], axis=2).astype(np.float32)
    b3 = rng.normal(0, 100, (B, N, 16)).astype(np.float32)
    t = lambda a: torch.from_numpy(a).to(cuda_device)
    ob, o3, osc, ol = K.filter_detections(t(boxes), t(b3), t(cls))
    kept = (ol >= 0)
    assert int(kept.sum()) > 0
    # rebuild a dense (B, 300, C) score tensor holding each kept score in its class column
    sc2 = torch.zeros((B, 300, C), dtype=torch.float32, device=cuda_device)
    idx = ol.clamp_min(0).long().unsqueeze(2)
    sc2.scatter_(2, idx, torch.where(kept, osc, torch.zeros_like(osc)).unsqueeze(2))
    ob2, o32, osc2, ol2 = K.filter_detections(torch.where(kept.unsqueeze(2), ob, torch.zeros_like(ob)),
                                              torch.where(kept.unsqueeze(2), o3, torch.zeros_like(o3)), sc2)
    assert torch.equal(ol2, ol)
    assert torch.equal(osc2, osc)
    assert torch.equal(ob2, ob)
    assert torch.equal(o32, o3)


def test_engine_full_batch_equals_single_image_runs(cuda_device):
    """Config 2 at its full size (batch 32, 480x640, 15 classes): every image's outputs are bit-identical to the
    outputs of the same image run alone — the batch-1 case is the one checked against the fp32 oracle in
    test_engine_gpu.py, so the full-size run inherits that parity. (Holds because an output element's K loop — taps x
    channel blocks, in order — does not depend on which tile, CTA or CTA pair computes it.)"""
    from pyrapose_b200 import models
    from pyrapose_b200.models import spec
    C, B, H, W = 15, 32, 480, 640
    weights = spec.init_weights(C, seed=1, head_init="spread", randomize_bn=True)
    train = models.backbone("resnet50").retinanet(C)
    train.set_weights_dict(weights)
    infer = models.convert_model(train)
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (B, H, W, 3)).astype(np.float32) - np.array([103.939, 116.779, 123.68], np.float32)
    # image 0 IS the image (and these are the weights) of tests/test_engine_gpu.py::test_engine_matches_oracle
    # [480x640_C15]: the batch-32 run is chained directly to the case compared with the fp32 oracle
    from test_engine_gpu import make_images
    img[0] = make_images(1, 1, H, W)[0]
    full = [np.array(o, copy=True) for o in infer.predict_on_batch(img)]
    assert full[0].shape == (B, 56700, 16) and full[1].shape == (B, 56700, C) and full[2].shape == (B, 4800, C)
    det_full = [np.array(o, copy=True) for o in infer.detect_on_batch(img)]
    for i in (0, 17, 31):
        one = infer.predict_on_batch(img[i:i + 1])
        for a, b in zip(one, full):
            assert np.array_equal(a[0], b[i])
        det_one = infer.detect_on_batch(img[i:i + 1])
        for a, b in zip(det_one, det_full):
            assert np.array_equal(a[0], b[i])
