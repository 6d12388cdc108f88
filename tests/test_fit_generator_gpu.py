"""model.fit_generator (models/_fit.py; the reference's training loop, PyraPose/bin/train.py:381-390) against a plain
loop of model.train_on_batch over the same batches in the same order.

The two run the same kernels; the only run-to-run variation is the order of the fp32 atomics that flush the weight
gradients, so the stated tolerance is the one tests/test_train_config4_gpu.py uses between two runs of one step:
per-batch losses relative 1e-3, final weights relative L2 1e-3 of the accumulated update."""
import numpy as np
import pytest

from synth import pose_annotations

pytestmark = pytest.mark.gpu

C, B, H, W = 3, 2, 96, 128
N_BATCHES = 5
LR = 1e-3


def _annotations(seed):
    a = pose_annotations(seed, C, n_obj=4, width=W, height=H)
    a['bboxes'][:, 0] = np.maximum(a['bboxes'][:, 0] - 12, 0)
    a['bboxes'][:, 1] = np.maximum(a['bboxes'][:, 1] - 12, 0)
    a['bboxes'][:, 2] = np.minimum(a['bboxes'][:, 2] + 12, W - 1)
    a['bboxes'][:, 3] = np.minimum(a['bboxes'][:, 3] + 12, H - 1)
    return a


class SyntheticSequence(object):
    """Shaped like the reference's Generator (preprocessing/generator.py:45-398): `__getitem__` returns
    (image batch, [regression_3D, labels, mask]) and the anchor-target function is injected
    (`compute_anchor_targets`, generator.py:58-59)."""

    def __init__(self, compute_anchor_targets, anchors, uint8=True):
        self.compute_anchor_targets = compute_anchor_targets
        self.anchors = anchors
        self.uint8 = uint8
        self.epoch_ends = 0

    def __len__(self):
        return N_BATCHES

    def __getitem__(self, index):
        rng = np.random.default_rng(100 + index)
        img = rng.integers(0, 256, (B, H, W, 3)).astype(np.uint8)
        if not self.uint8:
            img = img.astype(np.float32) - np.array([103.939, 116.779, 123.68], np.float32)
        anns = [_annotations(1000 + 10 * index + i) for i in range(B)]
        y = self.compute_anchor_targets(self.anchors, [np.zeros((H, W, 3), np.uint8)] * B, anns, C)
        return img, list(y)

    def on_epoch_end(self):
        self.epoch_ends += 1


class Recorder(object):
    def __init__(self):
        self.batches, self.epochs, self.events = [], [], []

    def on_train_begin(self, logs):
        self.events.append("train_begin")

    def on_epoch_begin(self, epoch, logs):
        self.events.append("epoch_begin %d" % epoch)

    def on_batch_end(self, batch, logs):
        self.batches.append(dict(logs))

    def on_epoch_end(self, epoch, logs):
        self.epochs.append(dict(logs))
        self.events.append("epoch_end %d" % epoch)

    def on_train_end(self, logs):
        self.events.append("train_end")


def _model(weights):
    from pyrapose_b200 import losses, models, optimizers
    m = models.backbone("resnet50").retinanet(num_classes=C)
    m.set_weights_dict({k: v.copy() for k, v in weights.items()})
    m.compile(loss={'3Dbox': losses.orthogonal_l1(), 'cls': losses.focal(), 'mask': losses.focal()},
              optimizer=optimizers.Adam(lr=LR, clipnorm=0.001))
    return m


@pytest.mark.parametrize("device_targets", [True, False], ids=["device_targets_uint8", "numpy_targets_float"])
def test_fit_generator_equals_a_loop_of_train_on_batch(cuda_device, device_targets):
    from pyrapose_b200.models import spec
    from pyrapose_b200.utils.anchors import (anchor_targets_bbox, anchor_targets_bbox_device, anchors_for_shape,
                                             anchors_for_shape_device)
    weights = spec.init_weights(C, seed=5, head_init="spread", randomize_bn=True)
    if device_targets:
        seq = SyntheticSequence(anchor_targets_bbox_device, anchors_for_shape_device((H, W, 3)), uint8=True)
    else:
        seq = SyntheticSequence(anchor_targets_bbox, anchors_for_shape((H, W, 3)), uint8=False)
    epochs = 2

    # the plain loop
    loop = _model(weights)
    want = []
    for _ in range(epochs):
        for k in range(N_BATCHES):
            x, y = seq[k]
            want.append(loop.train_on_batch(x, y))
    want = np.array(want)
    assert np.isfinite(want).all() and want[0, 0] > 0

    # fit_generator
    fit = _model(weights)
    rec = Recorder()
    seq.epoch_ends = 0
    hist = fit.fit_generator(seq, steps_per_epoch=len(seq), epochs=epochs, verbose=0, callbacks=[rec],
                             workers=3, use_multiprocessing=True, max_queue_size=10, shuffle=False)
    got = np.array([[l["loss"], l["3Dbox_loss"], l["cls_loss"], l["mask_loss"]] for l in rec.batches])
    assert got.shape == want.shape
    assert [l["batch"] for l in rec.batches] == list(range(N_BATCHES)) * epochs
    assert all(l["size"] == B for l in rec.batches)
    rel = np.abs(got - want) / np.abs(want)
    print("per-batch losses, fit_generator vs loop: max relative difference %.2e" % rel.max())
    assert rel.max() <= 1e-3
    # epoch logs are the size-weighted means of the batch logs; History holds them
    for e in range(epochs):
        mean = got[e * N_BATCHES:(e + 1) * N_BATCHES].mean(axis=0)
        assert abs(rec.epochs[e]["loss"] - mean[0]) <= 1e-9 * abs(mean[0]) + 1e-12
        assert hist.history["loss"][e] == rec.epochs[e]["loss"]
        assert hist.history["cls_loss"][e] == rec.epochs[e]["cls_loss"]
    assert hist.epoch == [0, 1] and seq.epoch_ends == epochs
    assert rec.events == ["train_begin", "epoch_begin 0", "epoch_end 0", "epoch_begin 1", "epoch_end 1", "train_end"]
    assert want[N_BATCHES, 0] < want[0, 0], "the same batch one epoch later has a lower loss"
    # the trained weights agree
    wa, wb = loop.get_weights_dict(), fit.get_weights_dict()
    worst = 0.0
    for name in ("res5c_branch2c/kernel", "P3/kernel", "pyramid_regression3D/kernel", "pyramid_classification/bias"):
        delta = np.linalg.norm(wa[name] - weights[name])
        assert delta > 0
        worst = max(worst, float(np.linalg.norm(wa[name] - wb[name]) / delta))
    print("trained weights: worst difference relative to the accumulated update %.2e" % worst)
    assert worst <= 1e-3


def test_fit_generator_stop_training_and_plain_generator(cuda_device):
    from pyrapose_b200.models import spec
    from pyrapose_b200.utils.anchors import anchor_targets_bbox_device, anchors_for_shape_device
    weights = spec.init_weights(C, seed=5, head_init="spread", randomize_bn=True)
    seq = SyntheticSequence(anchor_targets_bbox_device, anchors_for_shape_device((H, W, 3)))

    def forever():
        k = 0
        while True:
            yield seq[k % N_BATCHES]
            k += 1

    class StopAfter(object):
        def __init__(self, n):
            self.n, self.seen = n, 0

        def on_batch_end(self, batch, logs):
            self.seen += 1
            if self.seen >= self.n:
                self.model.stop_training = True

    m = _model(weights)
    with pytest.raises(ValueError):
        m.fit_generator(forever())                           # a plain generator needs steps_per_epoch
    cb = StopAfter(3)
    hist = m.fit_generator(forever(), steps_per_epoch=4, epochs=5, verbose=0, callbacks=[cb])
    # on_batch_end(k) runs after step k + 1 was queued: the loop stops one step after the callback asked
    assert 3 <= cb.seen <= 4 and len(hist.epoch) == 1

    def broken():
        yield seq[0]
        raise KeyError("annotations of batch 1 are missing")

    with pytest.raises(KeyError):
        m.fit_generator(broken(), steps_per_epoch=3, epochs=1, verbose=0)
    # the model is still usable afterwards
    x, y = seq[0]
    assert np.isfinite(m.train_on_batch(x, y)).all()
