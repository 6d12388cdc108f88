"""Batch staging for `model.fit_generator` — the queue Keras puts between a `keras.utils.Sequence` and the training
step (`training_model.fit_generator(generator, workers, use_multiprocessing, max_queue_size)`,
PyraPose/bin/train.py:381-390; the workers run `Generator.__getitem__`, preprocessing/generator.py:391-398).

The reference's workers are forked PROCESSES because `anchor_targets_bbox` is minutes of Python per epoch. Here the
targets are a few CUDA kernels (`utils.anchors.anchor_targets_bbox_device`), so one background THREAD is enough: it
pulls batch k + 1 from the generator, packs it and copies it to the GPU on its own stream while the GPU runs step k.

  * `BatchOrder`     which item comes next (Sequence with Keras' per-epoch shuffle, or a plain iterator)
  * `BatchPrefetcher` the thread + bounded queue; the staging function is injected (the CPU tests use a plain one)
  * `DeviceStager`   numpy -> pinned slot -> device copy on a side stream; CUDA tensors pass through
"""
import queue
import threading

import numpy as np


class BatchOrder(object):
    """Item source with Keras' semantics: a Sequence (`__len__` + `__getitem__`) is read by index, in a fresh random
    order every epoch when shuffle=True (fit_generator's default) and its `on_epoch_end()` is called after each epoch;
    anything else is an iterator that is simply advanced."""

    def __init__(self, generator, steps_per_epoch=None, shuffle=True, seed=None):
        self.generator = generator
        self.is_sequence = hasattr(generator, "__getitem__") and hasattr(generator, "__len__")
        if steps_per_epoch is None:
            if not self.is_sequence:
                raise ValueError("`steps_per_epoch=None` is only valid for a generator based on the "
                                 "`keras.utils.Sequence` class. Please specify `steps_per_epoch`.")
            steps_per_epoch = len(generator)
        self.steps = int(steps_per_epoch)     # train.py:383 passes a float (size / batch_size): Keras truncates it too
        if self.steps <= 0:
            raise ValueError("steps_per_epoch must be positive, got %r" % (steps_per_epoch,))
        self.shuffle = bool(shuffle) and self.is_sequence
        self.rng = np.random.default_rng(seed)
        self._it = None if self.is_sequence else iter(generator)

    def epoch(self):
        """Yields the `steps` items of one epoch."""
        if self.is_sequence:
            n = len(self.generator)
            order = self.rng.permutation(n) if self.shuffle else np.arange(n)
            for k in range(self.steps):
                yield self.generator[int(order[k % n])]
            if hasattr(self.generator, "on_epoch_end"):
                self.generator.on_epoch_end()
        else:
            for _ in range(self.steps):
                yield next(self._it)


class BatchPrefetcher(object):
    """Runs `stage(item)` for every item of `items` (an iterable) on a background thread and hands the results out in
    order through a queue of at most `depth` staged batches. An exception in the generator or in `stage` is re-raised
    by `get()` in the consumer; `close()` stops the thread (also on a consumer-side error)."""

    _END = object()

    def __init__(self, items, stage, depth=2, name="ppb-prefetch"):
        self._items = items
        self._stage = stage
        self._q = queue.Queue(maxsize=max(1, int(depth)))
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, name=name, daemon=True)
        self._thread.start()

    def _put(self, obj):
        while not self._stop.is_set():
            try:
                self._q.put(obj, timeout=0.05)
                return True
            except queue.Full:
                continue
        return False

    def _run(self):
        try:
            for item in self._items:
                if self._stop.is_set():
                    return
                if not self._put((None, self._stage(item))):
                    return
            self._put((None, self._END))
        except BaseException as e:  # handed to the consumer
            self._put((e, None))

    def get(self):
        """Next staged batch; raises StopIteration after the last one."""
        err, obj = self._q.get()
        if err is not None:
            self.close()
            raise err
        if obj is self._END:
            raise StopIteration
        return obj

    def __iter__(self):
        while True:
            try:
                yield self.get()
            except StopIteration:
                return

    def close(self):
        self._stop.set()
        try:
            while True:
                self._q.get_nowait()
        except queue.Empty:
            pass
        if self._thread is not threading.current_thread():
            self._thread.join(timeout=5.0)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


class StagedBatch(object):
    """One batch on the device: x, y = [regression_3D, labels, mask] (CUDA tensors), the event after which they are
    complete, and the batch size."""
    __slots__ = ("x", "y", "ready", "size")

    def __init__(self, x, y, ready, size):
        self.x, self.y, self.ready, self.size = x, y, ready, size

    def tensors(self):
        return [self.x] + list(self.y)


class DeviceStager(object):
    """stage(item) for BatchPrefetcher: takes one generator item `(inputs, targets[, sample_weights])` to the GPU.

    Everything runs on `self.stream` (a side stream) of `device`, in the prefetch thread:
      * a numpy array is copied into a pinned slot (slots rotate; a slot is reused only after the copy that read it has
        completed) and from there to the device with an asynchronous copy — the copy engine works while the SMs run the
        previous step;
      * a CUDA tensor (e.g. the three targets of `anchor_targets_bbox_device`, whose kernels the generator launched on
        this thread's current stream = the side stream) passes through.
    The consumer makes its stream wait for `StagedBatch.ready` and calls `record_stream` on the tensors it uses."""

    def __init__(self, device, n_slots=3):
        import torch
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device)
        self._slots = [dict(event=None, bufs={}) for _ in range(max(2, int(n_slots)))]
        self._k = 0

    def _to_device(self, slot, key, a):
        import torch
        if isinstance(a, torch.Tensor):
            if a.is_cuda:
                return a
            if a.is_pinned() and a.dtype in (torch.uint8, torch.float32) and a.is_contiguous():
                # already page-locked: copied from where it lies; the generator must not overwrite it before the
                # batch has been consumed (numpy arrays and pageable tensors are copied into the stager's own slots)
                return a.to(self.device, non_blocking=True)
            a = a.numpy()
        a = np.asarray(a)
        if a.dtype != np.uint8:
            a = a.astype(np.float32, copy=False)
        pin = slot["bufs"].get(key)
        if pin is None or tuple(pin.shape) != a.shape or pin.numpy().dtype != a.dtype:
            pin = torch.empty(a.shape, dtype=torch.uint8 if a.dtype == np.uint8 else torch.float32).pin_memory()
            slot["bufs"][key] = pin
        np.copyto(pin.numpy(), a)
        return pin.to(self.device, non_blocking=True)

    def __call__(self, item):
        import torch
        if not isinstance(item, (tuple, list)) or len(item) not in (2, 3):
            raise ValueError("Output of generator should be a tuple `(x, y, sample_weight)` or `(x, y)`. Found: %r"
                             % (type(item),))
        if len(item) == 3 and item[2] is not None:
            raise NotImplementedError("sample weights are not used by the reference's training loop (train.py:381-390)")
        x, y = item[0], item[1]
        if isinstance(x, (list, tuple)):   # Keras allows a one-element list of inputs
            if len(x) != 1:
                raise ValueError("the model has one input, got %d" % len(x))
            x = x[0]
        if not isinstance(y, (list, tuple)) or len(y) != 3:
            raise ValueError("expected targets = [regression_3D, labels, mask]")
        slot = self._slots[self._k % len(self._slots)]
        self._k += 1
        if slot["event"] is not None:
            slot["event"].synchronize()      # the copies that read this slot's pinned buffers are done
        with torch.cuda.device(self.device), torch.cuda.stream(self.stream):
            xd = self._to_device(slot, "x", x)
            yd = [self._to_device(slot, "y%d" % i, t) for i, t in enumerate(y)]
            ev = torch.cuda.Event()
            ev.record(self.stream)
        slot["event"] = ev
        return StagedBatch(xd, yd, ev, int(xd.shape[0]))

    def iterate(self, items):
        """Wraps an item iterable so that the GENERATOR's own CUDA work (device-side target kernels) is also issued on
        the side stream: the generator is advanced inside the stream context."""
        import torch
        it = iter(items)
        while True:
            with torch.cuda.device(self.device), torch.cuda.stream(self.stream):
                try:
                    item = next(it)
                except StopIteration:
                    return
            yield item
