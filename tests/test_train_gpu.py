"""Parity of the native training step (models.PFPNModel.compile / train_on_batch -> ppb_engine_train_*) against the
fp32 CPU oracle (oracle/train.py: torch autograd over the restated graph + restated Keras-2.3.1 Adam with global-norm
clipping) on the same random-init weights, synthetic images and anchor targets.

Stated tolerances (bf16 activations, weights AND gradient maps against an fp32 oracle):
  * loss values and the global gradient norm: relative 1e-2 (north_star's bf16 tolerance for forward quantities)
  * per-tensor raw gradients (before clipping, BatchNorm fold undone, l2 term included), for every trainable tensor
    whose gradient is not negligible (norm >= 1e-4 of the global norm): cosine >= 0.985 and relative L2 <= 0.17.
    Why not 1e-2: a bf16 forward pass differs from an fp32 one by ~1e-2 per activation, which flips the ReLU
    mask of the units nearest zero; every flipped unit contributes its whole gradient as error, so the deviation
    grows by ~1.5e-2 per ReLU layer on the way down (measured on this case: 4e-3 at the output convolutions, 3e-2
    after one tower layer, 7e-2 at the tower inputs, 0.13 worst at res5a_branch2a; profiles/r01_train_grad_parity.txt).
    It is noise, not bias: test_gradients_match_autograd_on_pinned_activations repeats the comparison with the
    oracle's stored activations and ReLU masks pinned to the engine's own (same piecewise-linear function on both
    sides) and every tensor then agrees to rel. L2 <= 1e-2 (measured worst 6e-3, cosine 0.99998); the kernels alone
    are pinned at ~1.7e-3 (bf16 output rounding) by tests/test_dgrad_gpu.py and ~1e-6 by tests/test_wgrad_gpu.py.
  * one Adam step: per-tensor weight delta: same bound as the gradients (the step is linear in the clipped gradient
    here: |g| << epsilon after the global-norm clip); frozen tensors (conv1, stage 2) bit-identical
"""
import numpy as np
import pytest

from synth import pose_annotations

pytestmark = pytest.mark.gpu

C, B, H, W = 3, 2, 96, 128
LR = 1e-3  # larger than the reference's 1e-5 so that one step moves fp32 weights by more than their rounding error


def _targets(anchors_fn, targets_fn):
    anchors = anchors_fn((H, W, 3))
    imgs = [np.zeros((H, W, 3), np.float32) for _ in range(B)]
    anns = []
    for i in range(B):
        a = pose_annotations(40 + i, C, n_obj=4, width=W, height=H)
        # objects sized for a 96x128 image: widen the 2-D boxes so that several anchors reach IoU 0.5
        a['bboxes'][:, 0] = np.maximum(a['bboxes'][:, 0] - 12, 0)
        a['bboxes'][:, 1] = np.maximum(a['bboxes'][:, 1] - 12, 0)
        a['bboxes'][:, 2] = np.minimum(a['bboxes'][:, 2] + 12, W - 1)
        a['bboxes'][:, 3] = np.minimum(a['bboxes'][:, 3] + 12, H - 1)
        anns.append(a)
    return targets_fn(anchors, imgs, anns, C)


@pytest.fixture(scope="module")
def step(cuda_device):
    from oracle import train as otrain
    from pyrapose_b200 import losses, models, optimizers
    from pyrapose_b200.models import spec
    from pyrapose_b200.utils.anchors import anchor_targets_bbox, anchors_for_shape
    weights = spec.init_weights(C, seed=5, head_init="spread", randomize_bn=True)
    rng = np.random.default_rng(7)
    img = rng.integers(0, 256, (B, H, W, 3)).astype(np.float32) - np.array([103.939, 116.779, 123.68], np.float32)
    y = _targets(anchors_for_shape, anchor_targets_bbox)
    assert (y[0][..., -1] == 1).sum() > 20, "synthetic targets need positives"
    ref = otrain.step(img, weights, y, C, lr=LR, clipnorm=0.001)

    model = models.backbone("resnet50").retinanet(num_classes=C)
    model.set_weights_dict({k: v.copy() for k, v in weights.items()})
    model.compile(loss={'3Dbox': losses.orthogonal_l1(), 'cls': losses.focal(), 'mask': losses.focal()},
                  optimizer=optimizers.Adam(lr=LR, clipnorm=0.001))
    out = model.train_on_batch(img, y)
    eng = model._train_engine
    # second oracle pass with every stored activation pinned to the engine's own copy (see the module docstring)
    pinned = {n: eng.activation(n) for n in pinned_names()}
    ref_pinned = otrain.step(img, weights, y, C, lr=LR, clipnorm=0.001, pinned=pinned)
    return dict(ref=ref, ref_pinned=ref_pinned, model=model, eng=eng, out=out, weights=weights)


def pinned_names():
    names = ["conv1_relu"]
    for stage, blocks in enumerate([3, 4, 6, 3]):
        for b in range(blocks):
            tag = "res%d%s" % (stage + 2, chr(ord("a") + b))
            names += [tag + "_branch2a_relu", tag + "_branch2b_relu", tag + "_relu"]
            if b == 0:
                names.append(tag + "_branch1_bn")
    names += ["P3_lateral", "P4_lateral", "P5_lateral", "P4_mid_in", "P4_mid", "P3_mid_in", "P3_mid", "P3_down",
              "P3_fin", "P3", "P4_down", "P4_fin", "P4", "P5_fin", "P5"]
    for i in range(4):
        names += ["reg_tower_t%d@%d" % (i, l) for l in range(3)] + ["cls_tower_t%d@%d" % (i, l) for l in range(3)]
        names.append("mask_tower_t%d@0" % i)
    return names


def _rel(a, b):
    return float(np.linalg.norm((a.astype(np.float64) - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-30))


def test_losses_match_oracle(step):
    ref, out, last = step["ref"]["losses"], step["out"], step["model"].last_step
    print("losses", out, ref, last)
    assert abs(out[1] - ref["3Dbox"]) <= 1e-2 * abs(ref["3Dbox"])
    assert abs(out[2] - ref["cls"]) <= 1e-2 * abs(ref["cls"])
    assert abs(out[3] - ref["mask"]) <= 1e-2 * abs(ref["mask"])
    assert abs(last["l2"] - ref["l2"]) <= 1e-4 * abs(ref["l2"])
    assert abs(out[0] - ref["total"]) <= 1e-2 * abs(ref["total"])
    assert abs(last["grad_norm"] - step["ref"]["grad_norm"]) <= 1e-2 * step["ref"]["grad_norm"]


def test_gradients_match_autograd(step):
    eng, ref, model = step["eng"], step["ref"], step["model"]
    gnorm = ref["grad_norm"]
    worst = []
    for l in model.layers:
        for bias in (False, True):
            key = l.name + ("/bias" if bias else "/kernel")
            got = eng.train_layer_tensor("grads", l.name, l.ksize, l.cin, l.cout, bias=bias)
            if key not in ref["grads"]:
                assert got is None, "frozen tensor %s has a gradient slot" % key
                continue
            assert got is not None, key
            want = ref["grads"][key]
            n = np.linalg.norm(want)
            if n < 1e-4 * gnorm:
                continue
            rel = _rel(got, want)
            cos = float((got.astype(np.float64) * want).sum() / (np.linalg.norm(got) * n))
            worst.append((rel, cos, key))
    worst.sort(reverse=True)
    print("worst gradient tensors:", worst[:8])
    assert len(worst) > 60
    for rel, cos, key in worst:
        assert rel <= 0.17 and cos >= 0.985, (key, rel, cos)


def test_gradients_match_autograd_on_pinned_activations(step):
    """The tight check of the backward pass: the oracle differentiates the same piecewise-linear function as the
    engine (its stored activations and ReLU masks are the engine's own, weights rounded like the engine's), so what
    is left is fp32 accumulation order and the bf16 rounding of the gradient maps. Every non-negligible tensor:
    relative L2 <= 1e-2, cosine >= 0.9999 (measured worst: 6.0e-3 / 0.99998)."""
    eng, ref, model = step["eng"], step["ref_pinned"], step["model"]
    gnorm = ref["grad_norm"]
    worst = []
    for l in model.layers:
        for bias in (False, True):
            key = l.name + ("/bias" if bias else "/kernel")
            if key not in ref["grads"]:
                continue
            got = eng.train_layer_tensor("grads", l.name, l.ksize, l.cin, l.cout, bias=bias)
            want = ref["grads"][key]
            n = np.linalg.norm(want)
            if n < 1e-4 * gnorm:
                continue
            rel = _rel(got, want)
            cos = float((got.astype(np.float64) * want).sum() / (np.linalg.norm(got) * n))
            worst.append((rel, cos, key))
    worst.sort(reverse=True)
    print("pinned: worst gradient tensors:", worst[:8])
    assert len(worst) > 60
    for rel, cos, key in worst:
        assert rel <= 1e-2 and cos >= 0.9999, (key, rel, cos)


def test_adam_step_matches_keras_semantics(step):
    eng, ref, model, w0 = step["eng"], step["ref"], step["model"], step["weights"]
    new = model.get_weights_dict()
    checked = 0
    for key, want in ref["weights"].items():
        d_want = want - w0[key]
        d_got = new[key] - w0[key]
        if np.linalg.norm(d_want) < 1e-12:
            continue
        rel = _rel(d_got, d_want)
        assert rel <= 0.17, (key, rel)
        checked += 1
    assert checked > 60
    for key in w0:
        if key not in ref["weights"]:
            assert np.array_equal(new[key], w0[key]), "frozen tensor %s changed" % key


def test_forward_after_step_uses_updated_weights(step):
    """The bf16 forward matrices are rewritten by the Adam kernel: a forward pass after the step must match the
    oracle forward on the UPDATED weights (rel. L2 2e-2 on the raw regression, as in test_engine_gpu)."""
    from oracle import pfpn
    model = step["model"]
    new = model.get_weights_dict()
    rng = np.random.default_rng(8)
    img = rng.integers(0, 256, (B, H, W, 3)).astype(np.float32) - np.array([103.939, 116.779, 123.68], np.float32)
    reg, cls, mask = model.predict_on_batch(img)
    ref_reg, ref_cls, ref_mask, _ = pfpn.forward_training(img, new, C)
    assert _rel(reg, ref_reg) <= 2e-2
    assert np.abs(cls - ref_cls).max() <= 2e-2


def test_second_step_decreases_nothing_weird(step):
    """Two more steps on the same batch run through the captured graph; losses stay finite and the step counter
    advances the bias correction (no NaN from the moments)."""
    model = step["model"]
    from pyrapose_b200.utils.anchors import anchor_targets_bbox, anchors_for_shape
    rng = np.random.default_rng(7)
    img = rng.integers(0, 256, (B, H, W, 3)).astype(np.float32) - np.array([103.939, 116.779, 123.68], np.float32)
    y = _targets(anchors_for_shape, anchor_targets_bbox)
    a = model.train_on_batch(img, y)
    b = model.train_on_batch(img, y)
    assert all(np.isfinite(a)) and all(np.isfinite(b))
    assert b[0] < a[0] * 1.05


def test_prediction_model_follows_training(cuda_device, tmp_path):
    """The reference builds prediction_model = retinanet_bbox(model) BEFORE training and evaluates through it after
    every epoch (PyraPose/bin/train.py:86-92, 119-129): the two share their layers. Here they share the engine: a
    predict_on_batch after train_on_batch must run on the TRAINED weights, convert_model() after training too, and
    save -> load_model must reproduce them."""
    from pyrapose_b200 import models, optimizers
    from pyrapose_b200.models import spec
    from pyrapose_b200.utils.anchors import anchor_targets_bbox, anchors_for_shape
    weights = spec.init_weights(C, seed=9, head_init="spread", randomize_bn=True)
    rng = np.random.default_rng(17)
    img = rng.integers(0, 256, (B, H, W, 3)).astype(np.float32) - np.array([103.939, 116.779, 123.68], np.float32)
    y = _targets(anchors_for_shape, anchor_targets_bbox)
    train = models.backbone("resnet50").retinanet(num_classes=C)
    train.set_weights_dict({k: v.copy() for k, v in weights.items()})
    pred = models.convert_model(train)                      # built before training, like train.py:92
    before = pred.predict_on_batch(img)
    train.compile(optimizer=optimizers.Adam(lr=1e-2, clipnorm=0.001))
    for _ in range(3):
        train.train_on_batch(img, y)
    after = pred.predict_on_batch(img)
    assert np.abs(after[1] - before[1]).max() > 1e-3, "the prediction model still runs the initial weights"
    t_out = train.predict_on_batch(img)                     # training model: [3Dbox raw, cls, mask]
    assert np.array_equal(after[1], t_out[1]) and np.array_equal(after[2], t_out[2])
    late = models.convert_model(train)                      # built after training
    assert np.array_equal(late.predict_on_batch(img)[1], after[1])
    # another training step after the inference binds: the engine goes back to the training plan with its state
    l1 = train.train_on_batch(img, y)
    assert all(np.isfinite(l1))
    path = str(tmp_path / "snapshot.npz")
    pred.save(path)                                         # syncs the trained master weights into the shared dict
    again = models.load_model(path, backbone_name="resnet50")
    a, b = again.predict_on_batch(img), pred.predict_on_batch(img)
    # fp32 master -> bf16 on load vs the Adam kernel's own bf16 rewrite: the same rounding of the same fp32 values
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[0], b[0])
    # loading weights into a model that has trained refreshes the training master copy as well
    train.set_weights_dict({k: v.copy() for k, v in weights.items()})
    assert np.array_equal(pred.predict_on_batch(img)[1], before[1])
    new = train.get_weights_dict()
    assert all(np.array_equal(new[k], weights[k]) for k in weights)


def test_separate_engines_follow_training_too(cuda_device):
    """Prediction models that do NOT share the training engine — the production arena (alias_arena=True) and the fp32
    path — are refilled from the trained master weights when the model has stepped since they were loaded."""
    from pyrapose_b200 import models, optimizers
    from pyrapose_b200.models import spec
    from pyrapose_b200.utils.anchors import anchor_targets_bbox, anchors_for_shape
    weights = spec.init_weights(C, seed=9, head_init="spread", randomize_bn=True)
    rng = np.random.default_rng(18)
    img = rng.integers(0, 256, (B, H, W, 3)).astype(np.float32) - np.array([103.939, 116.779, 123.68], np.float32)
    y = _targets(anchors_for_shape, anchor_targets_bbox)
    train = models.backbone("resnet50").retinanet(num_classes=C)
    train.set_weights_dict({k: v.copy() for k, v in weights.items()})
    shared = models.convert_model(train)
    prod = models.convert_model(train, alias_arena=True)
    fp32 = models.convert_model(train, precision='fp32')
    b_prod, b_fp32 = prod.predict_on_batch(img)[1].copy(), fp32.predict_on_batch(img)[1].copy()
    train.compile(optimizer=optimizers.Adam(lr=1e-2, clipnorm=0.001))
    for _ in range(2):
        train.train_on_batch(img, y)
    ref = shared.predict_on_batch(img)[1].copy()          # the engine the step updates in place
    a_prod, a_fp32 = prod.predict_on_batch(img)[1], fp32.predict_on_batch(img)[1]
    assert np.abs(a_prod - b_prod).max() > 1e-3 and np.abs(a_fp32 - b_fp32).max() > 1e-3
    assert np.array_equal(a_prod, ref)                     # same bf16 weights, same kernels, other buffers
    assert np.abs(a_fp32 - ref).max() < 5e-2               # fp32 path on the fp32 master weights vs the bf16 path
    train.train_on_batch(img, y)                           # and again after a further step
    assert np.array_equal(prod.predict_on_batch(img)[1], shared.predict_on_batch(img)[1])
