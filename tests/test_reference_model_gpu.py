"""The CUDA forward pass against the outputs of the model the REFERENCE'S OWN code builds
(tests/golden/reference_model.npz: PyraPose/models/{__init__,resnet,retinanet}.py run over the miniature Keras of
tests/golden/keras_shim.py; generator tests/golden/make_golden_model.py). No oracle in the chain.

Stated tolerances (BASELINE.json north_star): relative L2 <= 1e-4 on the fp32 path, <= 1e-2 on the bf16 path, for
feature maps, scores and decoded corners; the trainable / frozen split of the training engine equals the one the
reference's loop leaves behind (resnet.py:100-103)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "reference_model.npz"))


def _rel(a, b):
    return float(np.linalg.norm((a - b).astype(np.float64)) / np.linalg.norm(b.astype(np.float64)))


def _models(gold, precision):
    from pyrapose_b200 import models
    from pyrapose_b200.models import spec
    C, B, H, W, seed = [int(v) for v in gold["config"]]
    train = models.backbone("resnet50").retinanet(num_classes=C)
    train.set_weights_dict(spec.init_weights(C, seed=seed, head_init="spread", randomize_bn=True))
    return train, models.convert_model(train, precision=precision), (C, B, H, W)


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-4), ("bf16", 1e-2)])
def test_forward_matches_the_reference_built_model(cuda_device, gold, precision, tol):
    train, pred, (C, B, H, W) = _models(gold, precision)
    boxes3D, cls, msk = pred.predict_on_batch(gold["image"])
    rows = [("boxes3D", boxes3D, gold["pred_boxes3D"]), ("cls", cls, gold["train_cls"]), ("mask", msk, gold["train_mask"])]
    for ours, theirs in (("res3d_relu", "tap_C3"), ("res4f_relu", "tap_C4"), ("res5c_relu", "tap_C5"),
                         ("P3", "tap_P3"), ("P4", "tap_P4"), ("P5", "tap_P5")):
        rows.append((ours, pred.activation(ours, H, W), gold[theirs]))
    for name, got, want in rows:
        assert got.shape == want.shape and got.dtype == np.float32, (name, got.shape, want.shape)
        err = _rel(got, want)
        print("%s path %-12s relative L2 vs the reference-built model %.2e" % (precision, name, err))
        assert err <= tol, (precision, name, err)
    assert np.abs(cls - gold["train_cls"]).max() <= (1e-4 if precision == "fp32" else 2e-2)
    if precision == "bf16":     # the training model's raw outputs ([3Dbox, cls, mask], retinanet.py:260-299)
        reg, cls_t, msk_t = train.predict_on_batch(gold["image"])
        # raw regression deltas are not in the north_star list (decoded corners are, checked above at 1e-2); the
        # stated bar for them is 2e-2, as in tests/test_engine_gpu.py
        print("bf16 path 3Dbox (raw) relative L2 vs the reference-built model %.2e" % _rel(reg, gold["train_3Dbox"]))
        assert _rel(reg, gold["train_3Dbox"]) <= 2e-2 and _rel(cls_t, gold["train_cls"]) <= tol


def test_frozen_set_equals_the_reference_loop(cuda_device, gold):
    """resnet.py:100-103 over keras-resnet's layer list freezes conv1 and stage 2; BatchNormalization is frozen by
    freeze_bn=True. The training engine's per-layer `trainable` flag against the table the reference code left."""
    from pyrapose_b200 import optimizers
    from pyrapose_b200.models import keras_io
    train, _, (C, B, H, W) = _models(gold, "bf16")
    train.compile(optimizer=optimizers.Adam(lr=1e-5, clipnorm=0.001))
    y = [np.zeros((1, 2457, 17), np.float32), np.zeros((1, 2457, C + 1), np.float32),
         np.zeros((1, (H // 8) * (W // 8), C + 1), np.float32)]
    train.train_on_batch(gold["image"][:1], y)
    eng = train._train_engine
    name_map = keras_io.keras_layer_names()
    inverse = {(v[1] if v[0] == v[1] else v[0] + "/" + v[1]): k for k, v in name_map.items()}
    checked = 0
    for line in open(os.path.join(GOLD, "reference_model_layers.txt")):
        if line.startswith("#"):
            continue
        path, cls, trainable = line.split()[:3]
        if cls != "Conv2D":
            continue
        _, _, tr = eng.train_param_info(inverse.get(path, path))
        assert tr == bool(int(trainable.split("=")[1])), path
        checked += 1
    assert checked == 53 + 25
