"""Parity cases the round-1 review found untested:

  * BASELINE.json config 3, second run: 5 pyramid levels (N = 57 600 anchors), K = 100 boxes, batch 64 — anchors,
    the materialised IoU matrix and the fused assignment, bit-exact against the oracle (which is pinned to the
    reference's own Cython / numpy code, tests/test_oracle_anchors_overlap.py)
  * end-to-end FilterDetections agreement: the WHOLE fp32 oracle pipeline (network + decode + filter) against the
    whole bf16 device pipeline on the same image — the fraction of identical (anchor index, label) detections and the
    largest score difference, as a measured metric with a stated floor
  * vote filter (evaluators' np.where(scores > thr)) on the reference's own initialisers with a +-1 % band
"""
import hashlib
import os

import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_config3_five_levels_batch64_bit_exact(cuda_device):
    import torch
    from oracle import anchors as oa, overlap as oo
    from pyrapose_b200 import _kernels as K
    from pyrapose_b200.utils import anchors as ua
    five = ua.AnchorParameters(sizes=[32, 64, 128, 256, 512], strides=[8, 16, 32, 64, 128],
                               ratios=ua.AnchorParameters.default.ratios, scales=ua.AnchorParameters.default.scales)
    a_dev = ua.anchors_for_shape_device((480, 640, 3), pyramid_levels=[3, 4, 5, 6, 7], anchor_params=five)
    a = a_dev.cpu().numpy()
    assert a.shape == (57600, 4) and a.dtype == np.float64
    sha = dict(l.split() for l in open(os.path.join(GOLD, "sha256.txt")))
    assert hashlib.sha256(a.tobytes()).hexdigest() == sha["anchors_5level_480x640"]   # the reference's own output
    B, Kq = 64, 100
    qs = np.stack([synth.gt_boxes(1000 + i, Kq) for i in range(B)])
    q_dev = torch.from_numpy(qs).to(cuda_device)
    iou = K.compute_overlap(a_dev, q_dev)                                # (64, 57600, 100) float64 = 2.95 GB
    assert tuple(iou.shape) == (B, 57600, Kq)
    pos, ign, arg, mx = K.gt_assign(a_dev, q_dev, want_max=True)
    # fused assignment == max / first-argmax / thresholds of the materialised matrix, all 64 images, on the device
    m, am = iou.max(dim=2)
    assert torch.equal(mx, m)
    assert torch.equal(pos, m >= 0.5) and torch.equal(ign, (m > 0.4) & ~(m >= 0.5))
    first = (iou == m.unsqueeze(2)).to(torch.uint8).argmax(dim=2)       # np.argmax: the FIRST maximum wins
    assert torch.equal(arg, first)
    del m, am, first
    # bit-exact against the oracle for a sample of the images (the C port runs one image in ~0.1 s)
    for i in (0, 31, 63):
        ref = oo.compute_overlap(a, qs[i])
        assert np.array_equal(iou[i].cpu().numpy(), ref)
        p, g, amx, mxx = oo.gt_assign(a, qs[i])
        assert np.array_equal(pos[i].cpu().numpy(), p) and np.array_equal(ign[i].cpu().numpy(), g)
        assert np.array_equal(arg[i].cpu().numpy(), amx) and np.array_equal(mx[i].cpu().numpy(), mxx)
    assert int(pos.sum()) > 0


def test_end_to_end_filter_detections_agreement(cuda_device):
    """fp32 oracle network -> decode -> filter_detections against bf16 device network -> fused decode -> device
    FilterDetections, config 1 / 2 shape (480x640, 15 classes), the bench's weights (he-normal heads, cls bias -6:
    ~2 % of the 850 500 scores pass 0.05, so per-class NMS and the top-300 cut both do real work).

    A bf16 network cannot reproduce an fp32 one bit for bit, so this is a MEASURED agreement with a floor:
      matched = |{(anchor index, label)} of the oracle's 300 detections that the device also returns| / 300
    Floor 0.80 (a detection is lost when a 1e-2 score perturbation reorders a suppression chain or moves it across
    the rank-300 cut); scores of matched detections within 2e-2; decoded corners of matched detections 1e-2 rel."""
    import torch
    from oracle import decode as od, filter as of, pfpn
    from pyrapose_b200 import _kernels, models
    from pyrapose_b200.models import spec
    from test_engine_gpu import make_images
    C, H, W = 15, 480, 640
    weights = spec.init_weights(C, 9, seed=1, head_init="spread")
    weights["pyramid_classification/bias"] = np.full_like(weights["pyramid_classification/bias"], -6.0)
    train = models.backbone("resnet50").retinanet(num_classes=C)
    train.set_weights_dict(weights)
    model = models.convert_model(train)
    img = make_images(10, 1, H, W)
    ref_b3, ref_cls, _ = pfpn.forward_inference(img, weights, C)
    ref = of.filter_detections_batch(od.hull_boxes(ref_b3), ref_b3, ref_cls)
    buf = model.forward_device(img, want_hull=True)
    outs = _kernels.filter_detections(buf["hull"], buf["boxes3D"], buf["cls"], want_indices=True)
    got = [o.cpu().numpy() for o in outs]
    g_idx, g_lab, g_sc, g_b3 = got[4][0], got[3][0], got[2][0], got[1][0]
    # the oracle's kept anchor indices: recover them from the gathered boxes3D rows (rows of ref_b3 are distinct)
    r_lab, r_sc, r_b3 = ref[3][0], ref[2][0], ref[1][0]
    n_ref = int((r_lab >= 0).sum())
    assert n_ref == 300 and int((g_lab >= 0).sum()) == 300
    key = {}
    for j in range(300):
        key[(int(g_idx[j]), int(g_lab[j]))] = j
    lut = {ref_b3[0, i].tobytes(): i for i in np.where((ref_cls[0] > 0.05).any(axis=1))[0]}
    matched, dscore, dcorner = 0, 0.0, 0.0
    for j in range(300):
        i = lut[r_b3[j].tobytes()]
        k = key.get((i, int(r_lab[j])))
        if k is None:
            continue
        matched += 1
        dscore = max(dscore, abs(float(g_sc[k]) - float(r_sc[j])))
        dcorner = max(dcorner, float(np.linalg.norm(g_b3[k] - r_b3[j]) / np.linalg.norm(r_b3[j])))
    frac = matched / 300.0
    n_cand = int((ref_cls[0] > 0.05).sum())
    print("end-to-end FilterDetections agreement: %d / 300 identical (index, label) = %.3f; candidates %d; "
          "max |dscore| %.3e; max rel corner error %.3e" % (matched, frac, n_cand, dscore, dcorner))
    assert frac >= 0.80
    assert dscore <= 2e-2 and dcorner <= 1e-2


def test_vote_filter_band_one_percent_on_reference_initialisers(cuda_device):
    """kept indices must agree outside a +-1 % score band. With the reference's own initialisers
    every score sits near PriorProbability(0.01), so the evaluators' np.where(scores > thr) is exercised at
    thr = the median oracle score (half of the 850 500 scores on each side) and at its 90th percentile."""
    from oracle import pfpn
    from pyrapose_b200 import models
    from test_engine_gpu import make_images
    m = models.convert_model(models.backbone("resnet50").retinanet(num_classes=15, seed=1))
    img = make_images(0, 1, 480, 640)
    _, cls, _ = m.predict_on_batch(img)
    _, ref_cls, _ = pfpn.forward_inference(img, m.get_weights_dict(), 15)
    for q in (0.5, 0.9):
        thr = float(np.quantile(ref_cls, q))
        safe = np.abs(ref_cls - thr) > 0.01 * thr
        frac_safe = float(safe.mean())
        agree = np.array_equal((cls > thr)[safe], (ref_cls > thr)[safe])
        inside = float(((cls > thr) == (ref_cls > thr))[~safe].mean()) if (~safe).any() else 1.0
        print("vote filter thr %.6f (q%.0f): %.1f%% of scores outside the +-1%% band, all agree: %s; inside the band "
              "%.1f%% agree" % (thr, 100 * q, 100 * frac_safe, agree, 100 * inside))
        assert frac_safe > 0.2, "the band must not swallow the whole distribution"
        assert agree
