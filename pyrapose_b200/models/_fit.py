"""`model.fit_generator` — the training loop of the reference (PyraPose/bin/train.py:381-390:
`training_model.fit_generator(generator=train_generator, steps_per_epoch=..., epochs=..., verbose=1,
callbacks=callbacks, workers=..., use_multiprocessing=..., max_queue_size=...)`), around `model.train_on_batch`.

What Keras does per batch is kept: `on_batch_begin` / `on_batch_end(batch, logs)` with logs = {'batch', 'size', 'loss',
'3Dbox_loss', 'cls_loss', 'mask_loss'}, per-epoch means in `on_epoch_end(epoch, logs)` and in the returned
`History.history`, `generator.on_epoch_end()` after every epoch, `model.stop_training`.

What differs is WHEN things happen, so that the GPU never waits for the host:
  * batch k + 1 is fetched from the generator, packed and copied to the device by a background thread on a side stream
    (utils/prefetch.py) while step k runs;
  * the losses of step k are copied to pinned host memory asynchronously (ppb_engine_train_losses_async) and read —
    and `on_batch_end(k)` is called — right after step k + 1 has been QUEUED. Every callback still sees every batch's
    losses, one launch late; a callback that changes the learning rate in `on_batch_end(k)` takes effect from step
    k + 2 instead of k + 1 (the reference's callbacks act per epoch: callbacks/*, train.py:107-155).
`workers` / `use_multiprocessing` are accepted and ignored (one staging thread; the targets are CUDA kernels, see
prefetch.py); `max_queue_size` bounds the staged batches (each holds its targets on the device: capped at 4)."""
import time


from ..utils.prefetch import BatchOrder, BatchPrefetcher, DeviceStager

LOSS_KEYS = ("loss", "3Dbox_loss", "cls_loss", "mask_loss")


class History(object):
    """keras.callbacks.History: `.epoch` (indices) and `.history` {metric: [per-epoch mean]}."""

    def __init__(self):
        self.epoch = []
        self.history = {}
        self.model = None
        self.params = {}

    def record(self, epoch, logs):
        self.epoch.append(epoch)
        for k, v in logs.items():
            self.history.setdefault(k, []).append(v)


class _Callbacks(object):
    """Duck-typed keras.callbacks.CallbackList: calls the hooks a callback object has."""

    def __init__(self, callbacks, model, params):
        self.items = list(callbacks or [])
        for c in self.items:
            if hasattr(c, "set_model"):
                c.set_model(model)
            else:
                c.model = model
            if hasattr(c, "set_params"):
                c.set_params(params)
            else:
                c.params = params

    def call(self, hook, *args):
        for c in self.items:
            fn = getattr(c, hook, None)
            if fn is not None:
                fn(*args)


def _logs_from(raw5):
    """[3Dbox, cls, mask, l2, grad_norm^2] -> Keras' batch logs (loss includes the l2 regulariser, as Keras' does)."""
    b, c, m, l2 = float(raw5[0]), float(raw5[1]), float(raw5[2]), float(raw5[3])
    return {"loss": b + c + m + l2, "3Dbox_loss": b, "cls_loss": c, "mask_loss": m}


def fit_generator(model, generator, steps_per_epoch=None, epochs=1, verbose=1, callbacks=None, validation_data=None,
                  validation_steps=None, class_weight=None, max_queue_size=10, workers=1, use_multiprocessing=False,
                  shuffle=True, initial_epoch=0, seed=None):
    import torch
    if validation_data is not None or validation_steps is not None:
        raise NotImplementedError("validation runs through the evaluation callbacks in the reference "
                                  "(train.py:107-155), not through fit_generator's validation_data")
    if class_weight is not None:
        raise NotImplementedError("class_weight is not used by the reference's training loop")
    if not getattr(model, "_compiled", False):
        raise RuntimeError("You must compile a model before training/testing. Use `model.compile(optimizer, loss)`.")
    order = BatchOrder(generator, steps_per_epoch, shuffle=shuffle, seed=seed)
    steps = order.steps
    dev = getattr(model, "device", None)
    if dev is None:
        dev = torch.cuda.current_device()
    device = torch.device("cuda", dev) if isinstance(dev, int) else torch.device(dev)
    # the staging stream and the pinned slots (cudaHostAlloc costs milliseconds per slot) are kept across calls
    stager = model.__dict__.get("_stager")
    if stager is None or stager.device != device:
        stager = model.__dict__["_stager"] = DeviceStager(device, n_slots=3)
    depth = max(1, min(int(max_queue_size), 4))
    history = History()
    history.model = model
    params = {"epochs": int(epochs), "steps": steps, "verbose": verbose, "metrics": list(LOSS_KEYS)}
    history.params = params
    cbs = _Callbacks(callbacks, model, params)
    model.stop_training = False
    model.history = history

    main = torch.cuda.current_stream(device)
    loss_slots = model.__dict__.get("_loss_slots")
    if loss_slots is None:
        loss_slots = model.__dict__["_loss_slots"] = [
            dict(buf=torch.zeros(5, dtype=torch.float64).pin_memory(), ev=torch.cuda.Event()) for _ in range(2)]

    def all_items():
        for _ in range(initial_epoch, int(epochs)):
            for item in order.epoch():
                yield item

    cbs.call("on_train_begin", {})
    pf = BatchPrefetcher(stager.iterate(all_items()), stager, depth=depth)
    try:
        for epoch in range(initial_epoch, int(epochs)):
            cbs.call("on_epoch_begin", epoch, {})
            sums = dict.fromkeys(LOSS_KEYS, 0.0)
            seen = 0
            pending = None      # (batch index, size, loss slot) of the step whose losses are still in flight
            t_epoch = time.perf_counter()

            def resolve(p):
                nonlocal seen
                k, size, slot = p
                slot["ev"].synchronize()
                logs = _logs_from(slot["buf"].numpy())
                for key in LOSS_KEYS:
                    sums[key] += logs[key] * size
                seen += size
                model.last_step = logs
                logs = dict(logs, batch=k, size=size)
                cbs.call("on_batch_end", k, logs)

            for k in range(steps):
                staged = pf.get()
                cbs.call("on_batch_begin", k, {"batch": k, "size": staged.size})
                main.wait_event(staged.ready)
                model.train_on_batch(staged.x, staged.y, sync_losses=False)
                for t in staged.tensors():
                    t.record_stream(main)        # allocated on the staging stream, last used on this one
                slot = loss_slots[k & 1]
                model._train_engine.train_losses_async(slot["buf"])
                slot["ev"].record(main)
                if pending is not None:
                    resolve(pending)             # step k is already queued: the GPU does not wait for this read
                pending = (k, staged.size, slot)
                if model.stop_training:
                    break
            if pending is not None:
                resolve(pending)
            logs = {key: (sums[key] / seen if seen else float("nan")) for key in LOSS_KEYS}
            history.record(epoch, logs)
            cbs.call("on_epoch_end", epoch, logs)
            if verbose:
                dt = time.perf_counter() - t_epoch
                print("Epoch %d/%d - %d steps - %.1fs - %s" % (epoch + 1, int(epochs), steps, dt, " - ".join(
                    "%s: %.4f" % (key, logs[key]) for key in LOSS_KEYS)))
            if model.stop_training:
                break
    finally:
        pf.close()
        torch.cuda.current_stream(device).synchronize()
    cbs.call("on_train_end", {})
    return history
