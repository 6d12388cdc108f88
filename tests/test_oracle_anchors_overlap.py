"""Pins the CPU oracle (oracle/anchors.py, oracle/overlap.py) against golden vectors produced by the
REFERENCE's own code (tests/golden/make_golden.py) and, when oracle/_ref is present, against the
reference Cython extension directly. Bit-exact everywhere."""
import hashlib
import os

import numpy as np
import pytest

from oracle import anchors as oa
from oracle import build as obuild
from oracle import overlap as oo

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "anchors_overlap.npz"))


@pytest.fixture(scope="module")
def sha():
    d = {}
    for line in open(os.path.join(GOLD, "sha256.txt")):
        k, v = line.split()
        d[k] = v
    return d


def gt_boxes(seed, k=100, w=640, h=480):
    rng = np.random.default_rng(seed)
    x1 = rng.uniform(0, w - 16, k)
    y1 = rng.uniform(0, h - 16, k)
    bw = rng.uniform(16, 256, k)
    bh = rng.uniform(16, 256, k)
    return np.stack([x1, y1, np.minimum(x1 + bw, w), np.minimum(y1 + bh, h)], axis=1).astype(np.float64)


def test_generate_anchors_golden(gold):
    for size in (32, 64, 128):
        assert np.array_equal(oa.generate_anchors(base_size=size), gold["generate_anchors_%d" % size])
    # SURVEY.md §4 literal values
    a = oa.generate_anchors(base_size=32)
    assert a[0].tolist() == [-22.627416997969522, -11.313708498984761, 22.627416997969522, 11.313708498984761]
    assert a[3].tolist() == [-16.0, -16.0, 16.0, 16.0]
    assert a[4, 2] == 20.158737182617188


def test_anchors_for_shape_golden(gold, sha):
    a = oa.anchors_for_shape((480, 640, 3))
    assert a.shape == (56700, 4) and a.dtype == np.float64
    assert hashlib.sha256(a.tobytes()).hexdigest() == sha["anchors_480x640"]
    assert sha["anchors_480x640"] == "b02e1877e13002e595cc9f1db771db3b2812409c6f6bb4cdd3ef3e64fe1f5571"
    assert np.array_equal(a[:64], gold["anchors_480x640_head"])
    assert np.array_equal(a[-64:], gold["anchors_480x640_tail"])
    assert a.sum() == 63504000.0
    a2 = oa.anchors_for_shape((960, 1280, 3))
    assert a2.shape == (226800, 4)
    assert hashlib.sha256(a2.tobytes()).hexdigest() == sha["anchors_960x1280"]
    a5 = oa.anchors_for_shape((480, 640, 3), pyramid_levels=[3, 4, 5, 6, 7], anchor_params=oa.AnchorParameters.five_level)
    assert a5.shape == (57600, 4)
    assert hashlib.sha256(a5.tobytes()).hexdigest() == sha["anchors_5level_480x640"]
    ao = oa.anchors_for_shape((427, 561, 3))
    assert tuple(gold["anchors_odd_shape"]) == ao.shape
    assert hashlib.sha256(ao.tobytes()).hexdigest() == sha["anchors_odd_427x561"]


def test_overlap_port_matches_reference_golden(gold, sha):
    a = oa.anchors_for_shape((480, 640, 3))
    sub = a[::int(gold["iou_anchor_stride"][0])].copy()
    ov = oo.compute_overlap(sub, gold["iou_query"])
    assert np.array_equal(ov, gold["iou_matrix"])
    pos, ign, arg = oa.compute_gt_annotations(sub, gold["iou_query"])
    assert np.array_equal(pos, gold["assign_positive"])
    assert np.array_equal(ign, gold["assign_ignore"])
    assert np.array_equal(arg, gold["assign_argmax"])
    p2, i2, a2, _ = oo.gt_assign(sub, gold["iou_query"])
    assert np.array_equal(p2, pos) and np.array_equal(i2, ign) and np.array_equal(a2, arg)


@pytest.mark.parametrize("seed", [1000, 1001])
def test_overlap_full_size_hashes(gold, sha, seed):
    a = oa.anchors_for_shape((480, 640, 3))
    q = gt_boxes(seed)
    ov = oo.compute_overlap(a, q)
    assert hashlib.sha256(ov.tobytes()).hexdigest() == sha["iou_full_seed%d" % seed]
    p, i, arg, _ = oo.gt_assign(a, q)
    h = hashlib.sha256(p.astype(np.uint8).tobytes() + i.astype(np.uint8).tobytes() + arg.astype(np.int64).tobytes())
    assert h.hexdigest() == sha["assign_full_seed%d" % seed]
    assert [int(p.sum()), int(i.sum())] == gold["assign_full_counts_seed%d" % seed].tolist()


def test_overlap_known_answers(gold):
    ov = oo.compute_overlap(gold["kat_boxes"], gold["kat_query"])
    assert np.array_equal(ov, gold["kat_iou"])
    assert ov[0, 0] == 1.0 and ov[0, 1] == 0.0
    assert ov[0, 2] == 50.0 / 150.0      # iw=5, ih=10 with the +1 convention
    assert ov[0, 3] == 10.0 / 190.0      # shared edge pixel


def test_overlap_rejects_wrong_dtype_like_cython():
    with pytest.raises(ValueError):
        oo.compute_overlap(np.zeros((3, 4), np.float32), np.zeros((2, 4)))
    with pytest.raises(ValueError):
        oo.compute_overlap(np.zeros(4), np.zeros((2, 4)))


def test_overlap_port_vs_reference_extension_live():
    ref = obuild.load_ref_compute_overlap()
    if ref is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    rng = np.random.default_rng(3)
    b = rng.uniform(-50, 700, (4000, 4))
    b[:, 2:] = b[:, :2] + rng.uniform(0, 300, (4000, 2))
    q = gt_boxes(11, k=37)
    assert np.array_equal(ref(b, q), oo.compute_overlap(b, q))
    # degenerate / empty
    assert oo.compute_overlap(np.zeros((0, 4)), q).shape == (0, 37)
    assert np.array_equal(ref(b[:5], np.zeros((0, 4))), oo.compute_overlap(b[:5], np.zeros((0, 4))))


def test_regression_targets_golden(gold):
    assert np.array_equal(oa.bbox_transform(gold["xform_anchors"], gold["xform_gt4"]), gold["bbox_transform"])
    assert np.array_equal(oa.box3D_transform(gold["xform_anchors"], gold["xform_gt16"], 15), gold["box3D_transform"])
    assert np.array_equal(oa.toPix_array(gold["toPix_in"], fx=572.4114, fy=573.57043, cx=325.2611, cy=242.04899),
                          gold["toPix_out"])


def test_anchors_layer_f32_close_to_f64_and_shape():
    a32 = oa.anchors_layer_f32([(60, 80), (30, 40), (15, 20)])
    a64 = oa.anchors_for_shape((480, 640, 3))
    assert a32.shape == (56700, 4) and a32.dtype == np.float32
    assert np.max(np.abs(a32.astype(np.float64) - a64)) < 1e-4
