"""Pins the TensorFlow / Keras side of the oracle to the REFERENCE'S OWN TEXT.

tests/golden/reference_text.npz was produced by importing and running the reference's modules — backend/common.py,
layers/_misc.py, layers/filter_detections.py, losses.py, utils/anchors.py — over numpy stand-ins for the handful of
tensorflow / keras primitives they call (tests/golden/tf_shim.py, generator tests/golden/make_golden_shim.py; both
committed, run in the build container where /root/reference exists). Everything the reference's code decides (operand
order, the class loop and concatenation order of filter_detections, padding, loss normalisers, the 24 edge differences,
all of anchor_targets_bbox) is therefore fixed by these vectors; the library primitives (NMS-V3, top_k tie rule,
Keras' binary_crossentropy, nearest resize, quat2mat) are restated in the shim and cross-checked elsewhere
(tests/test_oracle_nms_crosscheck.py, tests/test_oracle_loss_crosscheck.py).

Bar: bit-exact for everything that is elementwise float32 or integer / index work; relative 2e-6 for the two loss
scalars (a float32 sum over ~1e5 terms: TF's, numpy's and torch's reduction orders differ)."""
import hashlib
import os

import numpy as np
import pytest

from synth import pose_annotations

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "reference_text.npz"))


@pytest.fixture(scope="module")
def sha():
    return dict(l.split() for l in open(os.path.join(GOLD, "reference_text_sha256.txt")) if l.strip())


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_decode_ops_bit_exact(gold):
    from oracle import decode as od
    anc, d = gold["dec_anchors"], gold["dec_deltas"]
    for name, got in (
        ("dec_box3D", od.box3D_transform_inv(anc, d)),
        ("dec_box3D_meanstd", od.box3D_transform_inv(anc, d, mean=gold["dec_mean"], std=gold["dec_std"])),
        ("dec_bbox", od.bbox_transform_inv(anc, d[:, :, :4])),
        ("dec_denorm", od.box3D_denormalization(anc, d)),
        ("layer_regress3D", od.box3D_transform_inv(anc, d)),
        ("layer_clip", od.clip_boxes((480, 640), gold["clip_in"])),
    ):
        assert got.dtype == np.float32 and got.shape == gold[name].shape, name
        assert np.array_equal(got, gold[name]), name


def test_anchors_layer_bit_exact(gold, sha):
    from oracle import anchors as oa
    full = oa.anchors_layer_f32([(60, 80), (30, 40), (15, 20)])
    assert full.dtype == np.float32 and _sha(full) == sha["anchors_layer_480x640_f32"]
    assert np.array_equal(full[:90], gold["anchors_layer_head"]) and np.array_equal(full[-90:], gold["anchors_layer_tail"])
    odd = oa.anchors_layer_f32([(7, 5)])
    assert np.array_equal(np.stack([odd, odd]), gold["anchors_layer_odd_7x5"])   # the layer tiles it over the batch


def test_upsample_like_bit_exact(gold):
    import torch
    from oracle import pfpn
    src = torch.from_numpy(gold["up_src"]).permute(0, 3, 1, 2)
    for key, (h, w) in (("up_30x40", (30, 40)), ("up_17x23", (17, 23))):
        got = pfpn.upsample_like(src, torch.zeros(2, 1, h, w)).permute(0, 2, 3, 1).numpy()
        assert np.array_equal(got, gold[key]), key


CASES = [("default", dict()), ("agnostic", dict(class_specific_filter=False)), ("no_nms", dict(nms=False)),
         ("no_nms_agnostic", dict(nms=False, class_specific_filter=False)),
         ("tight", dict(score_threshold=0.5, max_detections=100, nms_threshold=0.3))]


@pytest.mark.parametrize("name,kw", CASES, ids=[c[0] for c in CASES])
def test_filter_detections_exact(gold, name, kw):
    from oracle import filter as of
    got = of.filter_detections(gold["fd_boxes"], gold["fd_boxes3D"], gold["fd_cls"], other=[gold["fd_other"]], **kw)
    for key, g in zip(("boxes", "boxes3D", "scores", "labels", "other"), got):
        want = gold["fd_%s_%s" % (name, key)]
        assert g.shape == want.shape and g.dtype == want.dtype, (name, key, g.dtype, want.dtype)
        assert np.array_equal(g, want), (name, key)
    assert (gold["fd_%s_labels" % name] >= 0).sum() > 0


def test_filter_detections_dense_empty_and_layer(gold):
    from oracle import filter as of
    got = of.filter_detections(gold["fdd_in_boxes"], gold["fdd_in_boxes3D"], gold["fdd_in_cls"], max_detections=50)
    for key, g in zip(("boxes", "boxes3D", "scores", "labels"), got):
        assert np.array_equal(g, gold["fdd_%s" % key]), key
    assert (gold["fdd_labels"] >= 0).all()                       # more survivors than max_detections: no padding
    got = of.filter_detections(gold["fd_boxes"], gold["fd_boxes3D"], (gold["fd_cls"] * np.float32(0.01)).astype(np.float32))
    assert np.array_equal(got[3], gold["fd_empty_labels"]) and (got[3] == -1).all()
    assert np.array_equal(got[2], gold["fd_empty_scores"])
    # FilterDetections.call over a batch of two images
    b = np.stack([gold["fd_boxes"], gold["fdl_boxes2"]])
    b3 = np.stack([gold["fd_boxes3D"], gold["fdl_boxes3D2"]])
    c = np.stack([gold["fd_cls"], gold["fdl_cls2"]])
    got = of.filter_detections_batch(b, b3, c)
    for key, g in zip(("boxes", "boxes3D", "scores", "labels"), got):
        assert g.dtype == gold["fdl_%s" % key].dtype and np.array_equal(g, gold["fdl_%s" % key]), key


def _annotations(gold):
    C = 15
    anns = [pose_annotations(int(s), C) for s in gold["tgt_seeds"][:2]]
    empty = {k: (v[:0] if isinstance(v, np.ndarray) else v) for k, v in pose_annotations(int(gold["tgt_seeds"][2]), C).items()}
    return anns + [empty], C


def test_anchor_targets_bbox_bit_exact(gold, sha):
    from oracle import anchors as oa
    anns, C = _annotations(gold)
    anchors = oa.anchors_for_shape((480, 640, 3))
    reg, lab, msk = oa.anchor_targets_bbox(anchors, [np.zeros((480, 640, 3), np.float32)] * 3, anns, C)
    assert [list(a.shape) for a in (reg, lab, msk)] == gold["tgt_shapes"].tolist()
    assert np.array_equal(reg[:, :, -1].astype(np.int8), gold["tgt_states"])
    pos = gold["tgt_pos_index"]
    assert len(pos) > 100
    assert np.array_equal(reg[pos[:, 0], pos[:, 1], :16], gold["tgt_pos_regression"])
    assert np.array_equal(lab[pos[:, 0], pos[:, 1], :C].argmax(axis=1), gold["tgt_pos_label"])
    assert np.array_equal(np.argwhere(msk != 0).astype(np.int32), gold["tgt_mask_nonzero"])
    assert _sha(reg) == sha["targets_480x640_C15_regression_3D"]
    assert _sha(lab) == sha["targets_480x640_C15_labels"]
    assert _sha(msk) == sha["targets_480x640_C15_mask"]
    # image 2 has no objects: everything background
    assert not reg[2].any() and not lab[2].any() and not msk[2].any()
    # padded batch: anchors centred outside the smaller image are ignored
    images_p = [np.zeros((480, 640, 3), np.float32), np.zeros((360, 500, 3), np.float32)]
    reg_p, lab_p, _ = oa.anchor_targets_bbox(anchors, images_p, anns[:2], C)
    assert np.array_equal(reg_p[:, :, -1].astype(np.int8), gold["tgt_pad_states"])
    assert _sha(reg_p) == sha["targets_padded_regression_3D"] and _sha(lab_p) == sha["targets_padded_labels"]


def test_losses_match_reference_text(gold):
    import torch
    from oracle import losses as ol
    y_reg, y_cls = torch.from_numpy(gold["loss_y_reg"]), torch.from_numpy(gold["loss_y_cls"])
    p_reg, p_cls = torch.from_numpy(gold["loss_p_reg"]), torch.from_numpy(gold["loss_p_cls"])
    assert int((y_reg[:, :, -1] == 1).sum()) > 50

    def close(got, want, tol=2e-6):
        got, want = float(got), float(want[0])
        assert abs(got - want) <= tol * abs(want), (got, want)

    close(ol.focal(y_cls, p_cls), gold["loss_focal"])
    close(ol.focal(y_cls, p_cls, alpha=0.5, gamma=1.0), gold["loss_focal_a5_g1"])
    close(ol.orthogonal_l1(y_reg, p_reg), gold["loss_orth"])
    none = y_reg.clone()
    none[:, :, -1] = 0
    assert float(ol.orthogonal_l1(none, p_reg)) == float(gold["loss_orth_no_positives"][0]) == 0.0
    nonec = y_cls.clone()
    nonec[:, :, -1] = torch.where(nonec[:, :, -1] == 1, torch.zeros(()), nonec[:, :, -1])
    nonec[:, :, :15] = 0
    close(ol.focal(nonec, p_cls), gold["loss_focal_no_positives"])        # normaliser max(1, 0) = 1
    # float64 evaluation of the same restatement agrees too (the comparison is not a float32 accident)
    close(ol.focal(y_cls.double(), p_cls.double()), gold["loss_focal"], tol=2e-6)
    close(ol.orthogonal_l1(y_reg.double(), p_reg.double()), gold["loss_orth"], tol=2e-6)


def _raw_images(gold):
    rng = np.random.default_rng(int(gold["img_seed"][0]))
    return [rng.integers(0, 256, (int(h), int(w), 3)).astype(np.uint8) for h, w in gold["img_shapes"]]


def test_input_staging_host_mirror_equals_the_reference_functions(gold, sha):
    """utils.image.preprocess_image / compute_resize_scale / compute_inputs (the host-side mirror; no GPU) against the
    reference's own utils/image.py and Generator.compute_inputs. preprocess and the scale are exact; the resized
    pixels come from cv2.resize on both sides (the reference's call) and are compared at 2 ulp of the pixel range,
    since OpenCV dispatches SIMD code by CPU."""
    cv2 = pytest.importorskip("cv2")
    from pyrapose_b200.utils import image as ui
    staged = []
    for k, raw in enumerate(_raw_images(gold)):
        x = ui.preprocess_image(raw)
        assert x.dtype == np.float32 and _sha(x) == sha["preprocess_%dx%d" % raw.shape[:2]]
        s = ui.compute_resize_scale(x.shape)
        assert s == float(gold["img_scales"][k])
        y = cv2.resize(x, None, fx=s, fy=s)
        assert list(y.shape) == gold["img_resized_shape_%d" % k].tolist()
        assert np.abs(y[::7, ::11] - gold["img_resized_sample_%d" % k]).max() <= 3.1e-5
        staged.append(y)
    batch = ui.compute_inputs(staged)
    assert list(batch.shape) == gold["img_batch_shape"].tolist() and batch.dtype == np.float32
    assert np.abs(batch[:, ::9, ::13] - gold["img_batch_sample"]).max() <= 3.1e-5
    assert gold["img_batch_shape"].tolist() == [4, 640, 640, 3]
