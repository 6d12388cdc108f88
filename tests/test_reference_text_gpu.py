"""The CUDA path against vectors produced by the REFERENCE'S OWN TEXT (tests/golden/reference_text.npz: the reference's
backend/common.py, layers/_misc.py, layers/filter_detections.py and utils/anchors.py imported and run over numpy
stand-ins for the tensorflow / keras primitives — tests/golden/tf_shim.py, make_golden_shim.py). Same layer objects and
call signatures as the reference; every comparison is bit-exact (elementwise float32, integer and index work).

tests/test_oracle_reference_text.py pins the CPU oracle to the same vectors on the CPU box; this file removes the
oracle from the chain for the decode ops, the Anchors / UpsampleLike / RegressBoxes3D / ClipBoxes / FilterDetections
layers and anchor_targets_bbox."""
import hashlib
import os

import numpy as np
import pytest

from synth import pose_annotations

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "reference_text.npz"))


@pytest.fixture(scope="module")
def sha():
    return dict(l.split() for l in open(os.path.join(GOLD, "reference_text_sha256.txt")) if l.strip())


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _np(t):
    return t.detach().cpu().numpy()


def test_decode_layers_equal_the_reference_text(cuda_device, gold):
    import torch
    from pyrapose_b200 import layers
    dev = cuda_device
    anc, d = torch.from_numpy(gold["dec_anchors"]).to(dev), torch.from_numpy(gold["dec_deltas"]).to(dev)
    assert np.array_equal(_np(layers.RegressBoxes3D()([anc, d])), gold["dec_box3D"])
    assert np.array_equal(_np(layers.RegressBoxes3D()([anc, d])), gold["layer_regress3D"])
    got = layers.RegressBoxes3D(mean=gold["dec_mean"], std=gold["dec_std"])([anc, d])
    assert np.array_equal(_np(got), gold["dec_box3D_meanstd"])
    assert np.array_equal(_np(layers.RegressBoxes()([anc, d[:, :, :4].contiguous()])), gold["dec_bbox"])
    assert np.array_equal(_np(layers.DenormBoxes3D()([anc, d])), gold["dec_denorm"])
    img = torch.zeros((2, 480, 640, 3), device=dev)
    assert np.array_equal(_np(layers.ClipBoxes()([img, torch.from_numpy(gold["clip_in"]).to(dev)])), gold["layer_clip"])


def test_anchors_and_upsample_layers_equal_the_reference_text(cuda_device, gold, sha):
    import torch
    from pyrapose_b200 import layers
    dev = cuda_device
    levels = []
    for size, stride, (fh, fw) in ((32, 8, (60, 80)), (64, 16, (30, 40)), (128, 32, (15, 20))):
        levels.append(_np(layers.Anchors(size=size, stride=stride)(torch.zeros((1, fh, fw, 4), device=dev)))[0])
    full = np.concatenate(levels, axis=0)
    assert full.dtype == np.float32 and _sha(full) == sha["anchors_layer_480x640_f32"]
    odd = _np(layers.Anchors(size=32, stride=8)(torch.zeros((2, 7, 5, 3), device=dev)))
    assert np.array_equal(odd, gold["anchors_layer_odd_7x5"])
    src = torch.from_numpy(gold["up_src"]).to(dev)
    for key, (h, w) in (("up_30x40", (30, 40)), ("up_17x23", (17, 23))):
        got = _np(layers.UpsampleLike()([src, torch.zeros((2, h, w, 1), device=dev)]))
        assert np.array_equal(got, gold[key]), key


CASES = [("default", dict()), ("agnostic", dict(class_specific_filter=False)), ("no_nms", dict(nms=False)),
         ("no_nms_agnostic", dict(nms=False, class_specific_filter=False)),
         ("tight", dict(score_threshold=0.5, max_detections=100, nms_threshold=0.3))]


@pytest.mark.parametrize("name,kw", CASES, ids=[c[0] for c in CASES])
def test_filter_detections_layer_equals_the_reference_text(cuda_device, gold, name, kw):
    import torch
    from pyrapose_b200 import layers
    dev = cuda_device
    ins = [torch.from_numpy(gold[k][None]).to(dev) for k in ("fd_boxes", "fd_boxes3D", "fd_cls", "fd_other")]
    got = [_np(t)[0] for t in layers.FilterDetections(**kw)(ins)]
    for key, g in zip(("boxes", "boxes3D", "scores", "labels", "other"), got):
        want = gold["fd_%s_%s" % (name, key)]
        assert g.shape == want.shape and g.dtype == want.dtype, (name, key, g.dtype, want.dtype)
        assert np.array_equal(g, want), (name, key)


def test_filter_detections_dense_empty_and_batch(cuda_device, gold):
    import torch
    from pyrapose_b200 import layers
    dev = cuda_device
    ins = [torch.from_numpy(gold[k][None]).to(dev) for k in ("fdd_in_boxes", "fdd_in_boxes3D", "fdd_in_cls")]
    got = [_np(t)[0] for t in layers.FilterDetections(max_detections=50)(ins)]
    for key, g in zip(("boxes", "boxes3D", "scores", "labels"), got):
        assert np.array_equal(g, gold["fdd_%s" % key]), key
    low = (gold["fd_cls"] * np.float32(0.01)).astype(np.float32)
    ins = [torch.from_numpy(a[None]).to(dev) for a in (gold["fd_boxes"], gold["fd_boxes3D"], low)]
    got = [_np(t)[0] for t in layers.FilterDetections()(ins)]
    assert np.array_equal(got[3], gold["fd_empty_labels"]) and np.array_equal(got[2], gold["fd_empty_scores"])
    # the layer over a batch of two images (FilterDetections.call maps filter_detections over the batch)
    b = np.stack([gold["fd_boxes"], gold["fdl_boxes2"]])
    b3 = np.stack([gold["fd_boxes3D"], gold["fdl_boxes3D2"]])
    c = np.stack([gold["fd_cls"], gold["fdl_cls2"]])
    got = [_np(t) for t in layers.FilterDetections()([torch.from_numpy(a).to(dev) for a in (b, b3, c)])]
    for key, g in zip(("boxes", "boxes3D", "scores", "labels"), got):
        assert g.dtype == gold["fdl_%s" % key].dtype and np.array_equal(g, gold["fdl_%s" % key]), key


def test_anchor_targets_bbox_equals_the_reference_text(cuda_device, gold, sha):
    from pyrapose_b200.utils import anchors as ua
    C = 15
    anns = [pose_annotations(int(s), C) for s in gold["tgt_seeds"][:2]]
    empty = {k: (v[:0] if isinstance(v, np.ndarray) else v) for k, v in pose_annotations(int(gold["tgt_seeds"][2]), C).items()}
    anns.append(empty)
    anchors = ua.anchors_for_shape((480, 640, 3))
    reg, lab, msk = ua.anchor_targets_bbox(anchors, [np.zeros((480, 640, 3), np.float32)] * 3, anns, C)
    assert np.array_equal(reg[:, :, -1].astype(np.int8), gold["tgt_states"])
    pos = gold["tgt_pos_index"]
    assert np.array_equal(reg[pos[:, 0], pos[:, 1], :16], gold["tgt_pos_regression"])
    assert np.array_equal(np.argwhere(msk != 0).astype(np.int32), gold["tgt_mask_nonzero"])
    assert _sha(reg) == sha["targets_480x640_C15_regression_3D"]
    assert _sha(lab) == sha["targets_480x640_C15_labels"]
    assert _sha(msk) == sha["targets_480x640_C15_mask"]
    images_p = [np.zeros((480, 640, 3), np.float32), np.zeros((360, 500, 3), np.float32)]
    reg_p, lab_p, _ = ua.anchor_targets_bbox(anchors, images_p, anns[:2], C)
    assert np.array_equal(reg_p[:, :, -1].astype(np.int8), gold["tgt_pad_states"])
    assert _sha(reg_p) == sha["targets_padded_regression_3D"] and _sha(lab_p) == sha["targets_padded_labels"]


def test_stage_batch_equals_the_reference_functions(cuda_device, gold):
    """Raw uint8 images -> ui.stage_batch (preprocess + bilinear resize + zero-padded batch in CUDA kernels) against
    the reference's own preprocess_image + resize_image + Generator.compute_inputs (utils/image.py:35-62, 281-323;
    preprocessing/generator.py:320-336). Scales, shapes and the zero padding exact; pixels within 2 ulp of the pixel
    range (OpenCV's SIMD bilinear kernel may fuse multiply-adds; tests/test_image_gpu.py)."""
    from pyrapose_b200.utils import image as ui
    rng = np.random.default_rng(int(gold["img_seed"][0]))
    raws = [rng.integers(0, 256, (int(h), int(w), 3)).astype(np.uint8) for h, w in gold["img_shapes"]]
    got, scales = ui.stage_batch(raws)
    got = got.cpu().numpy()
    assert scales == [float(s) for s in gold["img_scales"]]
    assert list(got.shape) == gold["img_batch_shape"].tolist()
    d = np.abs(got[:, ::9, ::13] - gold["img_batch_sample"])
    print("stage_batch vs the reference's staging functions: max |diff| %.2e, identical %.1f%%" % (d.max(), 100 * (d == 0).mean()))
    assert d.max() <= 3.1e-5
    for b, (rows, cols) in enumerate(gold["img_batch_nonzero_rows"]):
        assert not got[b, rows:].any() and not got[b, :, cols:].any()
        assert got[b, rows - 1].any() and got[b, :, cols - 1].any()
