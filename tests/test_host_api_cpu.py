"""Host-side mirror of the reference's Python surface: names, constructor arguments, config
round trips and error behaviour — everything that does not need a device."""
import math

import numpy as np
import pytest

from oracle import anchors as oa
from pyrapose_b200 import initializers, layers, models
from pyrapose_b200.models import spec
from pyrapose_b200.utils import anchors as ua
from pyrapose_b200.utils import compute_overlap as uco
from pyrapose_b200.utils import shard
from pyrapose_b200.utils.image import preprocess_image


def test_backbone_factory_and_custom_objects():
    b = models.backbone("resnet50")
    for k in ("UpsampleLike", "PriorProbability", "RegressBoxes", "FilterDetections", "Anchors", "ClipBoxes",
              "RegressBoxes3D", "DenormBoxes3D"):
        assert k in b.custom_objects                              # PyraPose/models/__init__.py:10-21
    with pytest.raises(ValueError):
        models.backbone("resnet18")                               # resnet.py:63-69 allowed list
    with pytest.raises(NotImplementedError):
        models.backbone("vgg16")                                  # models/__init__.py:55-65


def test_training_and_inference_model_surface():
    train = models.backbone("resnet50").retinanet(num_classes=3)
    assert train.output_names == ["3Dbox", "cls", "mask"]         # models/__init__.py:79-81
    models.assert_training_model(train)
    infer = models.convert_model(train)
    assert infer.output_names == ["boxes3D", "cls", "mask"]       # retinanet.py:335
    with pytest.raises(AssertionError):
        models.assert_training_model(infer)
    with pytest.raises(ValueError):
        models.check_training_model(infer)                        # reference sys.exit(1) -> exception
    with pytest.raises(ValueError):
        infer.get_layer("nope")
    k, b = infer.get_layer("pyramid_classification").get_weights()
    assert k.shape == (3, 3, 256, 27) and np.allclose(b, -math.log(99.0))   # PriorProbability(0.01)
    assert infer.get_layer("pyramid_regression3D").get_weights()[0].shape == (3, 3, 512, 144)


def test_layer_table_matches_reference_graph():
    ls = spec.all_layers(15, 9)
    names = [l.name for l in ls]
    assert names[0] == "conv1" and "res5c_branch2c" in names and names.count("res3a_branch1") == 1
    assert sum(1 for n in names if n.startswith("res")) == 52     # 16 bottlenecks x 3 + 4 projection shortcuts
    for n in ("P3_lateral", "P4_mid", "P3_down", "P4_down", "P3", "P4", "P5", "mask_out"):
        assert n in names
    params = sum(l.ksize ** 2 * l.cin * l.cout + (l.cout if l.use_bias else 0) for l in ls)
    assert 40e6 < params < 45e6                                   # ~42.5 M parameters


def test_save_load_roundtrip(tmp_path):
    m = models.backbone("resnet50").retinanet(num_classes=2)
    p = str(tmp_path / "m.npz")
    m.save(p)
    m2 = models.load_model(p, backbone_name="resnet50")
    assert m2.num_classes == 2 and m2.output_names == m.output_names
    for k in ("conv1/kernel", "P3/bias", "bn2a_branch2a/gamma"):
        assert np.array_equal(m.weights[k], m2.weights[k])
    w = {"P3/bias": np.ones(256, np.float32), "unknown_layer/kernel": np.zeros(3)}
    m2.set_weights_dict(w)                                         # by_name: unknown names ignored
    assert m2.weights["P3/bias"][0] == 1.0
    with pytest.raises(ValueError):
        m2.set_weights_dict({"P3/bias": np.ones(7, np.float32)})
    m2.set_weights_dict({"P3/bias": np.ones(7, np.float32)}, skip_mismatch=True)


def test_regress_layers_validate_mean_std_like_the_reference():
    for cls, n in ((layers.RegressBoxes, 4), (layers.RegressBoxes3D, 16), (layers.DenormBoxes3D, 16)):
        l = cls()
        cfg = l.get_config()
        assert np.allclose(cfg["mean"], 0) and np.allclose(cfg["std"], 0.2) and len(cfg["std"]) == n
        with pytest.raises(ValueError):
            cls(mean=3.0)                                          # _misc.py:181-189
        with pytest.raises(ValueError):
            cls(std="x")
        l2 = cls(mean=cfg["mean"], std=cfg["std"])
        assert l2.get_config()["std"] == cfg["std"]


def test_anchors_layer_config_and_shapes():
    a = layers.Anchors(size=32, stride=8)
    cfg = a.get_config()
    assert cfg["size"] == 32 and cfg["stride"] == 8 and len(cfg["ratios"]) == 3 and len(cfg["scales"]) == 3
    assert a.num_anchors == 9 and a.anchors.dtype == np.float32    # _misc.py:52-56 cast
    assert a.compute_output_shape((2, 60, 80, 256)) == (2, 60 * 80 * 9, 4)
    assert a.compute_output_shape((None, None, None, 256)) == (None, None, 4)
    assert np.array_equal(a.anchors, oa.generate_anchors(32).astype(np.float32))


def test_filter_detections_layer_config_and_shapes():
    f = layers.FilterDetections(nms=False, max_detections=100, score_threshold=0.3)
    cfg = f.get_config()
    assert cfg["nms"] is False and cfg["max_detections"] == 100 and cfg["parallel_iterations"] == 32
    f2 = layers.FilterDetections(**cfg)
    assert f2.get_config() == cfg
    shapes = f.compute_output_shape([(4, 500, 4), (4, 500, 16), (4, 500, 7), (4, 500, 3, 2)])
    # `other` shapes keep input_shape[i][3:] exactly as filter_detections.py:208-210 does (a quirk of the reference:
    # the trailing dims start at index 3, not 2)
    assert shapes == [(4, 100, 4), (4, 100, 16), (4, 100), (4, 100), (4, 100, 2)]
    assert f.compute_mask([1, 2, 3]) == [None] * 4


def test_host_anchor_helpers_match_oracle():
    assert np.array_equal(ua.generate_anchors(64), oa.generate_anchors(64))
    assert [tuple(s) for s in ua.guess_shapes((480, 640, 3), [3, 4, 5])] == [(60, 80), (30, 40), (15, 20)]
    assert [tuple(s) for s in ua.guess_shapes((427, 561), [3, 4, 5])] == [(54, 71), (27, 36), (14, 18)]
    rng = np.random.default_rng(0)
    anc = oa.anchors_for_shape((64, 96, 3))
    gt16 = rng.uniform(0, 90, (anc.shape[0], 16))
    assert np.array_equal(ua.box3D_transform(anc, gt16), oa.box3D_transform(anc, gt16))
    gt4 = rng.uniform(0, 90, (anc.shape[0], 4))
    assert np.array_equal(ua.bbox_transform(anc, gt4), oa.bbox_transform(anc, gt4))
    with pytest.raises(ValueError):
        ua.bbox_transform(anc, gt4, mean=1.0)
    assert ua.AnchorParameters.default.num_anchors() == 9


def test_compute_overlap_argument_strictness_matches_cython():
    with pytest.raises(ValueError):
        uco.compute_overlap(np.zeros((3, 4), np.float32), np.zeros((2, 4)))
    with pytest.raises(ValueError):
        uco.compute_overlap(np.zeros((3, 4)), np.zeros(4))
    with pytest.raises(TypeError):
        uco.compute_overlap([[0, 0, 1, 1]], np.zeros((2, 4)))
    assert uco.compute_overlap(np.zeros((0, 4)), np.zeros((2, 4))).shape == (0, 2)   # no device needed


def test_prior_probability_and_preprocess():
    b = initializers.PriorProbability(0.01)((27,))
    assert b.dtype == np.float32 and np.allclose(b, -4.59511985)
    img = np.full((2, 2, 3), 128, np.uint8)
    out = preprocess_image(img)
    assert out.dtype == np.float32 and np.allclose(out[0, 0], 128 - np.array([103.939, 116.779, 123.68]))
    assert np.allclose(preprocess_image(img, mode="tf"), 128 / 127.5 - 1, atol=1e-6)


def test_shard_range_partitions_exactly():
    for total in (0, 1, 7, 32, 64, 65):
        for world in (1, 2, 3, 8):
            parts = [shard.shard_range(total, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == total
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in parts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.shard_range(4, 2, 2)


def test_compile_mirrors_the_reference_training_configuration():
    """models.PFPNModel.compile accepts exactly what PyraPose/bin/train.py:95-102 compiles and nothing else; no CUDA
    is touched until train_on_batch."""
    from pyrapose_b200 import losses, models, optimizers
    from pyrapose_b200.models.retinanet import PFPNModel
    m = PFPNModel(3, weights={})
    with pytest.raises(RuntimeError):
        m.train_on_batch(np.zeros((1, 64, 64, 3), np.float32), [None, None, None])
    m.compile(loss={'3Dbox': losses.orthogonal_l1(), 'cls': losses.focal(), 'mask': losses.focal()},
              optimizer=optimizers.Adam(lr=1e-5, clipnorm=0.001))
    assert m.metrics_names == ['loss', '3Dbox_loss', 'cls_loss', 'mask_loss']
    assert m.optimizer.get_config() == {'lr': 1e-5, 'beta_1': 0.9, 'beta_2': 0.999, 'epsilon': 1e-7, 'clipnorm': 0.001}
    with pytest.raises(ValueError):
        m.compile(loss={'3Dbox': losses.focal(), 'cls': losses.focal(), 'mask': losses.focal()})
    with pytest.raises(ValueError):
        m.compile(loss={'3Dbox': losses.orthogonal_l1(weight=1.0), 'cls': losses.focal(), 'mask': losses.focal()})
    with pytest.raises(ValueError):
        m.compile(loss={'cls': losses.focal()})
    with pytest.raises(ValueError):
        m.compile(optimizer="sgd")
    with pytest.raises(ValueError):
        models.convert_model(m).compile()
    with pytest.raises(TypeError):
        optimizers.Adam(nesterov=True)
    with pytest.raises(ValueError):
        m.train_on_batch(np.zeros((1, 64, 64, 3), np.float32), [None, None])


def test_training_oracle_modes_agree_on_a_tiny_case():
    """oracle/train.py: the bf16-emulating pass stays within 2e-2 of the fp32 pass on the losses (it only rounds), and
    frozen tensors get no gradient."""
    from oracle import train as otrain
    from pyrapose_b200.models import spec
    C, B, H, W = 2, 1, 64, 64
    w = spec.init_weights(C, seed=2, head_init="spread", randomize_bn=True)
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, (B, H, W, 3)).astype(np.float32) - 110.0
    N = 9 * (8 * 8 + 4 * 4 + 2 * 2)
    y_reg = np.zeros((B, N, 17), np.float32)
    y_cls = np.zeros((B, N, C + 1), np.float32)
    y_mask = np.zeros((B, 64, C + 1), np.float32)
    state = rng.choice([-1, 0, 1], size=(B, N), p=[.1, .7, .2]).astype(np.float32)
    y_reg[..., :16] = rng.normal(0, 1, (B, N, 16))
    y_reg[..., 16] = state
    y_cls[..., C] = state
    y_cls[..., 0] = state == 1
    y_mask[..., C] = 1
    y_mask[..., 1] = 1
    a = otrain.step(img, w, (y_reg, y_cls, y_mask), C, num_threads=4)
    b = otrain.step(img, w, (y_reg, y_cls, y_mask), C, emulate_bf16=True, num_threads=4)
    for k in ("3Dbox", "cls", "mask"):
        assert abs(a["losses"][k] - b["losses"][k]) <= 2e-2 * abs(a["losses"][k])
    assert a["losses"]["l2"] == b["losses"]["l2"]
    assert not any(k.startswith("conv1") or k.startswith("res2") or "bn" in k.split("/")[0] for k in a["grads"])
    assert "res3a_branch2a/kernel" in a["grads"] and "mask_out/bias" in a["grads"]
    assert a["grad_norm"] > 0 and np.isfinite(a["grad_norm"])
