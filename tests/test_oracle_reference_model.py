"""The PFPN forward oracle (oracle/pfpn.py) against the REFERENCE'S OWN model-building code.

tests/golden/reference_model.npz holds the outputs of the model that PyraPose/models/__init__.py, models/resnet.py and
models/retinanet.py BUILD (backbone('resnet50').retinanet(num_classes) and convert_model), run in the build container
over the miniature Keras of tests/golden/keras_shim.py (generator: tests/golden/make_golden_model.py). That pins the
wiring of the sparse pyramid and the three heads, every layer argument the reference passes, the Reshape /
Concatenate output order, Anchors + RegressBoxes3D of the prediction model, Keras' auto-naming of the unnamed layers
and the frozen set of resnet.py:100-103. Convolution / 'same' padding / max-pool / BatchNormalization arithmetic and
keras-resnet's ResNet50 are restated in the stand-in (cross-checked in tests/test_oracle_resnet_crosscheck.py).

Tolerance: both sides are float32 torch-CPU convolutions but fold BatchNormalization differently and sum in different
orders through ~60 layers: relative L2 <= 1e-5 on every tensor (measured <= 2.5e-6), scores max |diff| <= 5e-6."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "reference_model.npz"))


def _weights(gold):
    from pyrapose_b200.models import spec
    C, B, H, W, seed = [int(v) for v in gold["config"]]
    return spec.init_weights(C, seed=seed, head_init="spread", randomize_bn=True), C


def _rel(a, b):
    return float(np.linalg.norm((a - b).astype(np.float64)) / np.linalg.norm(b.astype(np.float64)))


def test_oracle_forward_matches_the_reference_built_model(gold):
    from oracle import pfpn
    w, C = _weights(gold)
    taps = {}
    reg, cls, msk, shapes = pfpn.forward_training(gold["image"], w, C, taps=taps)
    assert shapes == [(12, 17), (6, 9), (3, 5)]
    nhwc = pfpn.taps_to_nhwc(taps)
    rows = [("3Dbox", reg, gold["train_3Dbox"]), ("cls", cls, gold["train_cls"]), ("mask", msk, gold["train_mask"])]
    for ours, theirs in (("res3d_relu", "tap_C3"), ("res4f_relu", "tap_C4"), ("res5c_relu", "tap_C5"),
                         ("P3", "tap_P3"), ("P4", "tap_P4"), ("P5", "tap_P5")):
        rows.append((ours, nhwc[ours], gold[theirs]))
    for name, got, want in rows:
        assert got.shape == want.shape and got.dtype == np.float32, (name, got.shape, want.shape)
        err = _rel(got, want)
        print("%-12s relative L2 vs the reference-built model %.2e" % (name, err))
        assert err <= 1e-5, (name, err)
    assert np.abs(cls - gold["train_cls"]).max() <= 5e-6 and np.abs(msk - gold["train_mask"]).max() <= 5e-6
    boxes3D, cls2, msk2 = pfpn.forward_inference(gold["image"], w, C)
    assert _rel(boxes3D, gold["pred_boxes3D"]) <= 1e-5
    # the scores are spread (not all at the 0.01 prior), so the comparison is not vacuous
    assert gold["train_cls"].max() > 0.2 and gold["train_cls"].min() < 0.01


def test_keras_name_map_is_the_one_the_reference_code_produces():
    """pyrapose_b200/models/keras_io.py's engine-name -> Keras-name map against the layer table of the model the
    reference's code built (auto-numbered conv2d_<k> in model_1 / model_2 / mask / top level)."""
    from pyrapose_b200.models import keras_io, spec
    table = {}
    for line in open(os.path.join(GOLD, "reference_model_layers.txt")):
        if line.startswith("#"):
            assert "['3Dbox', 'cls', 'mask']" in line and "['boxes3D', 'cls', 'mask']" in line
            continue
        path, cls, trainable, *shapes = line.split()
        table[path] = (cls, int(trainable.split("=")[1]), dict(s.split("=") for s in shapes))
    name_map = keras_io.keras_layer_names()
    w = spec.init_weights(4, seed=5)
    seen = set()
    for key, arr in w.items():
        layer, param = key.split("/")
        group, kname = name_map.get(layer, (layer, layer))
        path = kname if group == kname else group + "/" + kname
        assert path in table, (key, path)
        assert table[path][2][param] == "x".join(str(d) for d in arr.shape), (key, path)
        seen.add(path)
    assert seen == set(table), sorted(set(table) - seen)
    # what the map claims, spelled out (PyraPose/models/retinanet.py creation order in a fresh process)
    assert name_map["pyramid_regression3D_0"] == ("model_1", "conv2d_1") and name_map["pyramid_regression3D"] == ("model_1", "conv2d_5")
    assert name_map["pyramid_classification"] == ("model_2", "conv2d_10") and name_map["mask_out"] == ("mask", "conv2d_15")
    assert table["model_1/conv2d_5"][2]["kernel"] == "3x3x512x144" and table["mask/conv2d_15"][2]["kernel"] == "3x3x256x4"
    assert [name_map[n][0] for n in ("P3_lateral", "P4_lateral", "P5_lateral", "P4_mid", "P3_mid", "P3_down", "P4_down")] == \
        ["conv2d_%d" % k for k in range(16, 23)]
    assert table["conv2d_16"][2]["kernel"] == "1x1x512x256" and table["conv2d_18"][2]["kernel"] == "1x1x2048x256"
    # frozen set after the reference's loop (resnet.py:100-103): conv1 and every convolution of stage 2, nothing else
    frozen = sorted(p for p, (cls, tr, _) in table.items() if cls == "Conv2D" and not tr)
    assert frozen == sorted(["conv1"] + ["res2%s_branch%s" % (b, br) for b in "abc" for br in ("2a", "2b", "2c")] +
                            ["res2a_branch1"])
    assert all(tr for p, (cls, tr, _) in table.items() if cls == "BatchNormalization")


def test_a_wrong_name_map_would_be_caught(gold):
    """Swapping two same-shaped pyramid convolutions (P4_mid <-> P3_down, both 3x3 256 -> 256) changes the oracle's
    output far beyond the tolerance: the vectors do discriminate between assignments that shapes cannot."""
    from oracle import pfpn
    w, C = _weights(gold)
    w = dict(w)
    for p in ("kernel", "bias"):
        w["P4_mid/" + p], w["P3_down/" + p] = w["P3_down/" + p], w["P4_mid/" + p]
    reg, cls, msk, _ = pfpn.forward_training(gold["image"], w, C)
    assert _rel(reg, gold["train_3Dbox"]) > 1e-2
