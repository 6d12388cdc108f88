"""Host logic of model.fit_generator (utils/prefetch.py, models/_fit.py) without a GPU: the batch order follows Keras'
rules for a keras.utils.Sequence (PyraPose/preprocessing/generator.py:384-398 is one) and for a plain generator, and the
staging thread hands batches over in order, bounded, with errors re-raised in the consumer."""
import threading
import time

import numpy as np
import pytest

from pyrapose_b200.utils.prefetch import BatchOrder, BatchPrefetcher


class Seq(object):
    def __init__(self, n):
        self.n = n
        self.epoch_ends = 0
        self.reads = []

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        self.reads.append(i)
        return (i, [i, i, i])

    def on_epoch_end(self):
        self.epoch_ends += 1


def test_sequence_order_sequential_and_shuffled():
    s = Seq(5)
    o = BatchOrder(s, None, shuffle=False)
    assert o.steps == 5
    assert [it[0] for it in o.epoch()] == [0, 1, 2, 3, 4] and s.epoch_ends == 1
    o = BatchOrder(s, None, shuffle=True, seed=3)
    e1 = [it[0] for it in o.epoch()]
    e2 = [it[0] for it in o.epoch()]
    assert sorted(e1) == sorted(e2) == [0, 1, 2, 3, 4]      # every batch once per epoch ...
    assert e1 != e2 and s.epoch_ends == 3                    # ... in a fresh order each time
    # the reference passes steps_per_epoch = size / batch_size, a float (bin/train.py:383): truncated
    assert BatchOrder(s, 3.7, shuffle=False).steps == 3
    # more steps than batches wraps around
    assert [it[0] for it in BatchOrder(s, 7, shuffle=False).epoch()] == [0, 1, 2, 3, 4, 0, 1]


def test_plain_generator_needs_steps():
    def gen():
        k = 0
        while True:
            yield (k, [k, k, k])
            k += 1
    with pytest.raises(ValueError):
        BatchOrder(gen(), None)
    o = BatchOrder(gen(), 3)
    assert [it[0] for it in o.epoch()] == [0, 1, 2]
    assert [it[0] for it in o.epoch()] == [3, 4, 5]          # the iterator just goes on
    with pytest.raises(ValueError):
        BatchOrder(Seq(2), 0)


def test_prefetcher_keeps_order_and_is_bounded():
    staged = []

    def stage(item):
        staged.append(item)
        return item * 10

    pf = BatchPrefetcher(iter(range(20)), stage, depth=2)
    time.sleep(0.3)
    assert len(staged) <= 4          # 2 queued + 1 blocked in put (+1 being staged): never the whole epoch
    out = list(pf)
    assert out == [10 * i for i in range(20)]
    pf.close()


def test_prefetcher_runs_on_another_thread_and_overlaps():
    main = threading.get_ident()
    seen = []

    def stage(item):
        seen.append(threading.get_ident())
        time.sleep(0.02)
        return item

    t0 = time.perf_counter()
    with BatchPrefetcher(iter(range(10)), stage, depth=2) as pf:
        for _ in pf:
            time.sleep(0.02)         # the consumer's own work
    dt = time.perf_counter() - t0
    assert all(t != main for t in seen)
    assert dt < 0.36, "staging (10 x 20 ms) and consuming (10 x 20 ms) must overlap, took %.3f s" % dt


def test_errors_reach_the_consumer_and_close_is_safe():
    def items():
        yield 1
        yield 2
        raise KeyError("annotations missing")

    pf = BatchPrefetcher(items(), lambda x: x, depth=1)
    assert pf.get() == 1 and pf.get() == 2
    with pytest.raises(KeyError):
        pf.get()

    def bad_stage(x):
        if x == 1:
            raise ValueError("expected targets = [regression_3D, labels, mask]")
        return x

    pf = BatchPrefetcher(iter(range(3)), bad_stage, depth=1)
    assert pf.get() == 0
    with pytest.raises(ValueError):
        pf.get()
    # closing with a full queue and a producer blocked in put() returns promptly
    pf = BatchPrefetcher(iter(range(100)), lambda x: x, depth=1)
    time.sleep(0.1)
    t0 = time.perf_counter()
    pf.close()
    assert time.perf_counter() - t0 < 2.0 and not pf._thread.is_alive()


def test_fit_generator_argument_checks_need_no_gpu():
    from pyrapose_b200 import models
    from pyrapose_b200.models import _fit
    m = models.backbone("resnet50").retinanet(num_classes=3)
    with pytest.raises(RuntimeError):
        m.fit_generator(Seq(2))                               # not compiled
    m._compiled = True
    with pytest.raises(NotImplementedError):
        m.fit_generator(Seq(2), validation_data=Seq(1))
    logs = _fit._logs_from(np.array([1.0, 2.0, 3.0, 0.5, 9.0]))
    assert logs == {"loss": 6.5, "3Dbox_loss": 1.0, "cls_loss": 2.0, "mask_loss": 3.0}
    h = _fit.History()
    h.record(0, logs)
    h.record(1, logs)
    assert h.epoch == [0, 1] and h.history["loss"] == [6.5, 6.5]
