"""PFPN model objects — mirror of the builders in PyraPose/models/retinanet.py.

  retinanet(inputs, backbone_layers, num_classes, ...)   -> training model, outputs ['3Dbox', 'cls', 'mask'] (:260-299)
  retinanet_bbox(model, nms, class_specific_filter, ...) -> inference model, outputs [boxes3D, cls, mask]      (:302-335)

The returned PFPNModel answers the calls the reference's callers make on a keras.Model for this path:
predict_on_batch, output_names, load_weights(by_name=True), save / save_weights, summary, get_layer(name).weights.
The graph itself runs in the native engine (tcgen05 implicit-GEMM convolutions, one CUDA graph per batch size).
"""
import os

import numpy as np

from . import assert_training_model, spec
from ..utils.anchors import AnchorParameters, generate_anchors


class _LayerView(object):
    def __init__(self, model, name):
        self.name = name
        self._model = model

    def get_weights(self):
        keys = [k for k in self._model.weights if k.split("/")[0] == self.name]
        order = {"kernel": 0, "bias": 1, "gamma": 0, "beta": 1, "moving_mean": 2, "moving_variance": 3}
        return [self._model.weights[k] for k in sorted(keys, key=lambda k: order[k.split("/")[1]])]


class PFPNModel(object):
    def __init__(self, num_classes, num_anchors=None, name='retinanet', inference=False, anchor_params=None,
                 weights=None, seed=None, device=None, precision='bf16', alias_arena=False):
        if precision not in ('bf16', 'fp32'):
            raise ValueError("precision must be 'bf16' (production path, tolerance 1e-2) or 'fp32' (three-part bf16 "
                             "split, six MMAs per product, tolerance 1e-4; inference only), got %r" % (precision,))
        self.precision = precision
        # production arena: backbone intermediates share buffers (6.4 -> 3.7 GB at batch 32); inference models only,
        # and model.activation() can then only return tensors that survive the pass (outputs, C3-C5, the pyramid)
        self.alias_arena = bool(alias_arena)
        self.num_classes = int(num_classes)
        self.num_anchors = int(num_anchors if num_anchors is not None else AnchorParameters.default.num_anchors())
        self.name = name
        self.inference = bool(inference)
        self.anchor_params = anchor_params if anchor_params is not None else AnchorParameters.default
        self.device = device
        self.layers = spec.all_layers(self.num_classes, self.num_anchors)
        self.weights = weights if weights is not None else spec.init_weights(
            self.num_classes, self.num_anchors, seed=1 if seed is None else seed)
        self._engines = {}         # (H, W) -> Engine; SHARED with the prediction model convert_model builds
        self._parent = None        # prediction model: the training model whose layers (engines, weights) it shares
        self._host_out = {}        # predict_on_batch: persistent pinned output buffers per output shape
        self.regress_mean = None   # RegressBoxes3D(mean, std) defaults: 0 / 0.2 (layers/_misc.py:176-179)
        self.regress_std = None

    # ------------------------------------------------------------------ keras.Model-like surface
    @property
    def output_names(self):
        return ['boxes3D', 'cls', 'mask'] if self.inference else ['3Dbox', 'cls', 'mask']

    def get_layer(self, name):
        if not any(k.split("/")[0] == name for k in self.weights):
            raise ValueError('No such layer: ' + name)
        return _LayerView(self, name)

    def summary(self, print_fn=print):
        total = 0
        print_fn('Model: "%s"' % self.name)
        for l in self.layers:
            n = l.ksize * l.ksize * l.cin * l.cout + (l.cout if l.use_bias else 0) + (4 * l.cout if l.bn else 0)
            total += n
            print_fn("%-28s conv %dx%d/%d  %4d -> %4d  params %d" % (l.name, l.ksize, l.ksize, l.stride, l.cin, l.cout, n))
        print_fn("Total params: {:,}".format(total))

    def _root(self):
        return self if self._parent is None else self._parent._root()

    def get_weights_dict(self):
        self.sync_weights()
        return self.weights

    def set_weights_dict(self, weights, skip_mismatch=False):
        self.sync_weights()   # tensors NOT named in `weights` keep their trained values
        for k, v in weights.items():
            if k not in self.weights:
                continue  # by_name semantics: unknown names are ignored
            v = np.asarray(v, dtype=np.float32)
            if v.shape != self.weights[k].shape:
                if skip_mismatch:
                    continue
                raise ValueError("shape mismatch for %s: %s vs %s" % (k, v.shape, self.weights[k].shape))
            self.weights[k] = v
        for e in self._engines.values():
            e.load_weights(self.weights, self._bn_names())
            e._wver = self._root().__dict__.get('_wver', 0)

    def load_weights(self, filepath, by_name=True, skip_mismatch=False):
        """`model.load_weights(path, by_name=True, skip_mismatch=...)` (PyraPose/bin/train.py:65-68).
        Keras HDF5 files (what the reference writes and what keras-resnet's ImageNet checkpoint is,
        PyraPose/models/resnet.py:89-98) are read by the pure-Python reader in utils/hdf5.py and mapped to this model's
        layers by name where the reference names them and by graph order where it does not (models/keras_io.py);
        .npz containers keyed '<layer>/<param>' are accepted as well. Layers the file does not hold keep their values
        (by_name semantics); a shape mismatch raises unless skip_mismatch."""
        from . import keras_io
        if keras_io.is_hdf5(filepath):
            w, _unused = keras_io.load_h5(filepath, self.weights.keys())
            return self.set_weights_dict(w, skip_mismatch=skip_mismatch)
        with np.load(filepath) as z:
            self.set_weights_dict({k: z[k] for k in z.files if not k.startswith("__")}, skip_mismatch=skip_mismatch)

    @staticmethod
    def _wants_h5(filepath):
        return str(filepath).lower().endswith((".h5", ".hdf5", ".keras.h5"))

    def save_weights(self, filepath):
        """Keras HDF5 layout with the reference's layer names for *.h5 / *.hdf5 (loadable by the reference's
        model.load_weights(by_name=True)); npz otherwise."""
        self.sync_weights()
        if self._wants_h5(filepath):
            from . import keras_io
            return keras_io.save_h5(filepath, self.weights, self.layers)
        np.savez(filepath, **self.weights)

    def save(self, filepath):
        """ModelCheckpoint's full-model save (bin/train.py:131-142): *.h5 gets Keras' 'model_weights' group; the
        architecture travels as (num_classes, num_anchors, inference) instead of Keras' model_config JSON."""
        self.sync_weights()
        meta = np.array([self.num_classes, self.num_anchors, int(self.inference)], dtype=np.int64)
        if self._wants_h5(filepath):
            from . import keras_io
            return keras_io.save_h5(filepath, self.weights, self.layers, meta=meta, full_model=True)
        with open(filepath, "wb") as f:
            np.savez(f, __meta__=meta, **self.weights)

    @classmethod
    def load(cls, filepath):
        from . import keras_io
        if keras_io.is_hdf5(filepath):
            layers, attrs = keras_io.read_layers(filepath)
            if "pyrapose_b200_meta" in attrs:
                meta = [int(v) for v in np.atleast_1d(attrs["pyrapose_b200_meta"])]
            else:   # a snapshot written by the reference: classes from the head shapes, inference model
                meta = [keras_io.infer_classes(layers), 9, 1]
            model = cls(meta[0], meta[1], inference=bool(meta[2]))
            w, _unused = keras_io.to_engine_weights(layers, model.weights.keys())
            missing = [k for k in model.weights if k not in w]
            if missing:
                raise ValueError("%s: %d tensors of the model are missing, e.g. %s" % (filepath, len(missing), missing[:3]))
            model.set_weights_dict(w)
            return model
        with np.load(filepath) as z:
            meta = z["__meta__"]
            weights = {k: z[k] for k in z.files if not k.startswith("__")}
        return cls(int(meta[0]), int(meta[1]), inference=bool(meta[2]), weights=weights)

    # ------------------------------------------------------------------ engine plumbing
    def _bn_names(self):
        return {l.name: l.bn for l in self.layers if l.bn}

    def _engine(self, batch, height, width):
        import torch
        from ._engine import Engine
        fp32 = self.precision == 'fp32'
        key = (height, width, 'fp32') if fp32 else (height, width)
        if self.alias_arena:
            key = key + ('alias',)
        eng = self._engines.get(key)
        if eng is None or eng.max_batch < batch:
            if eng is not None:
                # a larger batch needs a larger arena: the trained master weights go back into self.weights first so
                # the new engine starts from them (the Adam moments restart)
                self._root().sync_weights()
                if self._root().__dict__.get('_train_engine') is eng:
                    self._root()._train_engine = None
                eng.close()
            dev = self.device if self.device is not None else torch.cuda.current_device()
            if fp32:
                self._root().sync_weights()   # a trained parent's master weights first
            if self.alias_arena:
                self._root().sync_weights()
            eng = Engine(dev, batch, height, width, self.num_classes, self.num_anchors, precision=1 if fp32 else 0,
                         alias_arena=self.alias_arena)
            eng.load_weights(self.weights, self._bn_names())
            eng._anchor_owner = None
            eng._wver = self._root().__dict__.get('_wver', 0)
            self._engines[key] = eng
        elif (fp32 or self.alias_arena) and eng._wver != self._root().__dict__.get('_wver', 0):
            # engines of the fp32 path / the production arena are not the one the training step updates in place:
            # reload them when the model has stepped since they were filled
            self._root().sync_weights()
            eng.load_weights(self.weights, self._bn_names())
            eng._wver = self._root().__dict__.get('_wver', 0)
        if getattr(eng, '_anchor_owner', None) is not self:
            p = self.anchor_params
            # Anchors layer: float64 base anchors cast to float32 (layers/_misc.py:52-56)
            base = np.stack([generate_anchors(p.sizes[i], p.ratios, p.scales).astype(np.float32) for i in range(3)])
            eng.set_anchors(base, np.asarray(p.strides[:3], np.float32), self.regress_mean, self.regress_std)
            eng._anchor_owner = self
        return eng

    @staticmethod
    def _as_input(images):
        """-> (torch tensor, is_uint8). float arrays are BGR mean-subtracted images (what the reference's callers pass
        after preprocess_image); uint8 arrays are RAW BGR images: preprocess_image(mode='caffe') then runs fused in the
        stem's loader on the device (utils/image.py:35-62), with identical results and a quarter of the copy."""
        import torch
        if isinstance(images, np.ndarray):
            if images.ndim != 4 or images.shape[3] != 3:
                raise ValueError("expected images of shape (B, H, W, 3), got %s" % (images.shape,))
            if images.dtype == np.uint8:
                return torch.from_numpy(np.ascontiguousarray(images)), True
            return torch.from_numpy(np.ascontiguousarray(images, dtype=np.float32)), False
        if images.dim() != 4 or images.shape[3] != 3:
            raise ValueError("expected images of shape (B, H, W, 3), got %s" % (tuple(images.shape),))
        if images.dtype == torch.uint8:
            return images, True
        return (images.float() if images.dtype != torch.float32 else images), False

    def forward_device(self, images, want_hull=False, eager=False):
        """images: (B,H,W,3) numpy (pinned or not) or CUDA tensor; float32 BGR mean-subtracted, or raw uint8 BGR.
        Returns the engine's output tensors (valid until the next call): dict boxes3D / cls / mask / hull."""
        import torch
        src, u8 = self._as_input(images)
        B, H, W = int(src.shape[0]), int(src.shape[1]), int(src.shape[2])
        eng = self._engine(B, H, W)
        buf = eng.bind(B, decode=self.inference, want_mask=True, want_hull=want_hull, input_u8=u8)
        with torch.cuda.device(eng.device):
            buf["images"].copy_(src, non_blocking=True)
            eng.run(eager=eager)
        return buf

    def predict_on_batch(self, x):
        """ Returns [boxes3D (B,N,16), classification (B,N,C), mask (B,H/8*W/8,C)] float32 numpy for an inference
        model ([3Dbox, cls, mask] for a training model), as the reference's model.predict_on_batch does
        (PyraPose/utils/linemod_eval.py:303)."""
        import torch
        import weakref
        src, u8 = self._as_input(x)
        B, H, W = int(src.shape[0]), int(src.shape[1]), int(src.shape[2])
        eng = self._engine(B, H, W)
        buf = eng.bind(B, decode=self.inference, want_mask=True, want_hull=False, input_u8=u8)
        outs = [buf["boxes3D"], buf["cls"], buf["mask"]]
        key = tuple(tuple(o.shape) for o in outs)
        # Output buffers are persistent PINNED host tensors, so the three device->host copies (234 MB at batch 32) are
        # asynchronous DMAs at PCIe rate, and the arrays handed to the caller are views of them: no host memcpy, no
        # page faults on 234 MB of fresh memory per call (that alone was 45 ms of a 60 ms call). The caller still OWNS
        # what it gets, like the fresh arrays Keras returns: a buffer set is only reused once every array handed out
        # from it (and every view derived from one) has been dropped — tracked with weak references; while the caller
        # holds on to results, new sets are allocated (up to 4 per output shape, then plain copies).
        pool = self._host_out.setdefault(key, [])
        if len(self._host_out) > 4:
            self._host_out = {key: pool}
        slot = None
        for s in pool:
            if all(r() is None for r in s["refs"]):
                slot = s
                break
        if slot is None and len(pool) < 4:
            slot = {"host": [torch.empty(o.shape, dtype=o.dtype).pin_memory() for o in outs], "refs": []}
            pool.append(slot)
        with torch.cuda.device(outs[0].device):
            main = torch.cuda.current_stream()
            buf["images"].copy_(src, non_blocking=True)
            if slot is None:   # every set is still referenced by the caller: private pageable copies
                eng.run()
                return [o.cpu().numpy() for o in outs]
            # The forward pass runs as three graphs cut where an output is complete (classification head, regression
            # head, mask head: ppb_engine_run_part), and each output's device->host DMA starts on a copy stream as soon
            # as its part is done: the 109 MB of 'cls' travel while the regression head computes, the 116 MB of
            # 'boxes3D' while the mask head does.
            copy = self.__dict__.get('_copy_stream')
            if copy is None or copy.device != outs[0].device:
                copy = self._copy_stream = torch.cuda.Stream(device=outs[0].device)
            for part, oi in ((0, 1), (1, 0), (2, 2)):
                eng.run_part(part)
                ev = torch.cuda.Event()
                ev.record(main)
                copy.wait_event(ev)
                with torch.cuda.stream(copy):
                    slot["host"][oi].copy_(outs[oi], non_blocking=True)
            copy.synchronize()
        arrays = [h.numpy() for h in slot["host"]]
        slot["refs"] = [weakref.ref(a) for a in arrays]
        return arrays

    def predict_on_images(self, raw_images, min_side=480, max_side=640):
        """ The evaluators' per-image sequence read_image_bgr -> preprocess_image -> resize_image -> predict_on_batch
        (PyraPose/utils/linemod_eval.py:283-303) for a list of RAW uint8 BGR images of any sizes: preprocessing,
        bilinear resize and zero padding run on the device straight into the engine's input tensor
        (utils.image.stage_batch). Returns (predict_on_batch outputs, scales)."""
        from ..utils import image as ui
        scales = [ui.compute_resize_scale(im.shape, min_side, max_side) for im in raw_images]
        shapes = [ui._resized_shape(im.shape, s) for im, s in zip(raw_images, scales)]
        B, H, W = len(raw_images), max(s[0] for s in shapes), max(s[1] for s in shapes)
        eng = self._engine(B, H, W)
        buf = eng.bind(B, decode=self.inference, want_mask=True, want_hull=False, input_u8=False)
        ui.stage_batch(raw_images, min_side, max_side, out=buf["images"])
        eng.run()
        return [buf[k].cpu().numpy() for k in ("boxes3D", "cls", "mask")], scales

    def detect_on_batch(self, x, score_threshold=0.05, max_detections=300, nms=True, nms_threshold=0.5,
                        class_specific_filter=True):
        """Forward + decode + FilterDetections entirely on the device; only the (B, max_detections, .) results are
        copied to the host. FilterDetections' 2-D `boxes` input is the axis-aligned hull of the decoded corners
        (the PyraPose heads produce no 2-D boxes). Returns [boxes, boxes3D, scores, labels] numpy."""
        from .. import _kernels
        if not self.inference:
            raise ValueError("detect_on_batch needs an inference model (models.convert_model)")
        buf = self.forward_device(x, want_hull=True)
        out = _kernels.filter_detections(buf["hull"], buf["boxes3D"], buf["cls"],
                                         class_specific_filter=class_specific_filter, nms=nms,
                                         score_threshold=score_threshold, max_detections=max_detections,
                                         nms_threshold=nms_threshold)
        return [o.cpu().numpy() for o in out]

    def estimate_poses_on_batch(self, x, threeD_boxes, camera, threshold=0.5, min_votes=10, iterations=300,
                                reprojection_error=5.0):
        """ predict_on_batch + the evaluators' vote filter + pose from votes, all on the device
        (PyraPose/utils/linemod_eval.py:303-334, 427-430, 479-485): per image a list of
        {cls, rvec, R, t, votes, inliers, confidence, rmse}. See utils/pose.py."""
        from ..utils import pose
        if not self.inference:
            raise ValueError("estimate_poses_on_batch needs an inference model (models.convert_model)")
        buf = self.forward_device(x)
        return pose.estimate_poses(buf["cls"], buf["boxes3D"], threeD_boxes, camera, threshold, min_votes, iterations,
                                   reprojection_error)

    def detect_stream(self, batches, score_threshold=0.05, max_detections=300, nms=True, nms_threshold=0.5,
                      class_specific_filter=True):
        """Pipelined detect_on_batch over an iterable of host batches (the loop keras.Model.predict /
        predict_generator runs over its batches): the host->device copy of batch i+1 runs on a copy stream while
        batch i is in the engine, and the (B, max_detections, .) results come back through pinned buffers while
        batch i+1 computes. Yields [boxes, boxes3D, scores, labels] numpy arrays per batch, in order.
        Every batch still pays its own H2D and D2H; they are overlapped, not skipped."""
        import torch
        from .. import _kernels
        if not self.inference:
            raise ValueError("detect_stream needs an inference model (models.convert_model)")
        kw = dict(class_specific_filter=class_specific_filter, nms=nms, score_threshold=score_threshold,
                  max_detections=max_detections, nms_threshold=nms_threshold)
        copy_stream = None
        staging, staged_ev, consumed_ev, host_out, pending = [None, None], [None, None], [None, None], [None, None], []

        def finish(item):
            ev, outs = item
            ev.synchronize()
            return [o.numpy().copy() for o in outs]

        for i, x in enumerate(batches):
            src, u8 = self._as_input(x)
            B, H, W = int(src.shape[0]), int(src.shape[1]), int(src.shape[2])
            eng = self._engine(B, H, W)
            buf = eng.bind(B, decode=True, want_mask=True, want_hull=True, input_u8=u8)
            k = i & 1
            with torch.cuda.device(eng.device):
                main = torch.cuda.current_stream()
                if copy_stream is None:
                    copy_stream = torch.cuda.Stream()
                if staging[k] is None or staging[k].shape != buf["images"].shape or staging[k].dtype != buf["images"].dtype:
                    staging[k] = torch.empty_like(buf["images"])
                # the staging buffer was last read by the device-side copy of batch i-2: wait for THAT copy only, not
                # for the forward pass of batch i-1 that is still running on `main`
                if consumed_ev[k] is not None:
                    copy_stream.wait_event(consumed_ev[k])
                with torch.cuda.stream(copy_stream):
                    staging[k].copy_(src, non_blocking=True)
                    staged_ev[k] = torch.cuda.Event()
                    staged_ev[k].record(copy_stream)
                main.wait_event(staged_ev[k])
                buf["images"].copy_(staging[k], non_blocking=True)
                consumed_ev[k] = torch.cuda.Event()
                consumed_ev[k].record(main)
                eng.run()
                outs = _kernels.filter_detections(buf["hull"], buf["boxes3D"], buf["cls"], **kw)
                if host_out[k] is None or host_out[k][0].shape[0] != B or host_out[k][0].shape[1] != max_detections:
                    host_out[k] = [torch.empty(o.shape, dtype=o.dtype).pin_memory() for o in outs]
                for h, o in zip(host_out[k], outs):
                    h.copy_(o, non_blocking=True)
                done = torch.cuda.Event()
                done.record(main)
            pending.append((done, host_out[k]))
            if len(pending) == 2:  # results of batch i-1 are read while batch i runs
                yield finish(pending.pop(0))
        while pending:
            yield finish(pending.pop(0))

    def activation(self, name, height, width):
        key = (height, width, 'fp32') if self.precision == 'fp32' else (height, width)
        if self.alias_arena:
            key = key + ('alias',)
        return self._engines[key].activation(name)

    # ------------------------------------------------------------------ training (keras.Model.compile / train_on_batch)
    def compile(self, loss=None, optimizer=None, dp_normalizer='global', bucket_bytes=25 << 20, data_parallel=True,
                exchange='auto', **kwargs):
        """ Mirror of `training_model.compile(loss={'3Dbox': orthogonal_l1(), 'cls': focal(), 'mask': focal()},
        optimizer=Adam(lr, clipnorm=0.001))` (PyraPose/bin/train.py:95-102). Only that configuration exists in the
        native step; anything else raises ValueError.

        Data parallel (one process per GPU, torch.distributed initialised; the reference's
        keras.utils.multi_gpu_model, bin/train.py:82-86):
          dp_normalizer='global'  — the losses are normalised by the positives of the GLOBAL batch and the replica
                                    gradients are summed: exactly the single-replica step on the concatenated batch,
                                    which is what multi_gpu_model computes (it concatenates the replica outputs and
                                    evaluates each loss once);
          dp_normalizer='replica' — every replica normalises by its own positives (losses.py:61-66 on the local
                                    batch) and the gradients are averaged.
          bucket_bytes            — size of the gradient buckets whose NCCL all-reduce overlaps the rest of the
                                    backward pass (0: one all-reduce after the backward pass);
          data_parallel=False     — this model steps alone even when torch.distributed is initialised;
          exchange                — how the replica gradients are summed: 'nccl' (= 'auto') all-reduces the buckets
                                    on NCCL's stream while the backward pass runs; 'p2p' pushes every finished bucket
                                    into the peers' memory with copy engines over NVLink and adds the slices in one
                                    kernel after the backward pass (utils/exchange.py; single node). Measured r02:
                                    equal at 2 GPUs (13.30-13.42 vs 13.31 ms per step), slower at 8 (15.08 vs 13.63 ms:
                                    an all-gather by copies moves N-1 times the gradient), hence not the default."""
        if exchange not in ('auto', 'p2p', 'nccl'):
            raise ValueError("exchange must be 'auto', 'p2p' or 'nccl'")
        self.exchange = os.environ.get("PPB_EXCHANGE", exchange)
        if dp_normalizer not in ('global', 'replica'):
            raise ValueError("dp_normalizer must be 'global' or 'replica'")
        self.dp_normalizer = dp_normalizer
        self.bucket_bytes = int(bucket_bytes)
        if os.environ.get("PPB_BUCKET_MB"):   # A/B runs of the bucket size without touching the caller
            self.bucket_bytes = int(float(os.environ["PPB_BUCKET_MB"]) * (1 << 20))
        self.data_parallel = bool(data_parallel)
        from .. import losses, optimizers
        if self.inference:
            raise ValueError("compile() needs a training model")
        loss = loss if loss is not None else {'3Dbox': losses.orthogonal_l1(), 'cls': losses.focal(),
                                              'mask': losses.focal()}
        want = {'3Dbox': ('orthogonal_l1', {'weight': 0.125, 'sigma': 3.0}),
                'cls': ('focal', {'alpha': 0.25, 'gamma': 2.0}), 'mask': ('focal', {'alpha': 0.25, 'gamma': 2.0})}
        if set(loss) != set(want):
            raise ValueError("loss must name the outputs %s" % sorted(want))
        for k, (kind, params) in want.items():
            if getattr(loss[k], 'kind', None) != kind or loss[k].params != params:
                raise ValueError("output %s: the native training step implements %s(%s) only" % (k, kind, params))
        self.optimizer = optimizer if optimizer is not None else optimizers.Adam(lr=1e-5, clipnorm=0.001)
        if not isinstance(self.optimizer, optimizers.Adam):
            raise ValueError("optimizer must be pyrapose_b200.optimizers.Adam")
        self._compiled = True
        self._train_engine = getattr(self, '_train_engine', None)
        self.metrics_names = ['loss', '3Dbox_loss', 'cls_loss', 'mask_loss']

    def train_on_batch(self, x, y, sync_losses=True):
        """ One optimisation step on a batch, as `fit_generator` runs it for every batch of the reference generator.

        x: images (B,H,W,3) float32 (mean-subtracted) or raw uint8 BGR; y: [regression_3D (B,N,17), labels (B,N,C+1), mask (B,H/8*W/8,C+1)] — the three
        arrays anchor_targets_bbox returns. numpy or CUDA tensors. With torch.distributed initialised (one replica per
        GPU) the gradient buckets are all-reduced over NCCL while the rest of the backward pass runs; see compile().
        Returns [loss, 3Dbox_loss, cls_loss, mask_loss] (loss includes the l2 regulariser) — the local replica's
        share (with dp_normalizer='global' the shares of all replicas add up to the loss of the global batch) — or
        None when sync_losses=False (no host synchronisation)."""
        import torch
        if not getattr(self, '_compiled', False):
            raise RuntimeError("You must compile a model before training/testing. Use `model.compile(optimizer, loss)`.")
        if not isinstance(y, (list, tuple)) or len(y) != 3:
            raise ValueError("expected y = [regression_3D, labels, mask]")
        if self.precision == 'fp32':
            raise RuntimeError("the fp32 path (precision='fp32') is inference only; train with precision='bf16'")
        x, u8 = self._as_input(x)
        B, H, W = int(x.shape[0]), int(x.shape[1]), int(x.shape[2])
        dist = torch.distributed
        world = dist.get_world_size() if (self.data_parallel and dist.is_available() and dist.is_initialized()) else 1
        eng = self._engine(B, H, W)
        buf = eng.train_bind(B, input_u8=u8, bucket_bytes=self.bucket_bytes if world > 1 else 0)
        self._train_engine = eng

        def put(dst, src):
            if isinstance(src, np.ndarray):
                src = torch.from_numpy(np.ascontiguousarray(src, dtype=np.uint8 if src.dtype == np.uint8 else np.float32))
            if tuple(src.shape) != tuple(dst.shape):
                raise ValueError("expected an array of shape %s, got %s" % (tuple(dst.shape), tuple(src.shape)))
            dst.copy_(src, non_blocking=True)

        opt = self.optimizer
        with torch.cuda.device(eng.device):
            put(buf["images"], x)
            put(buf["y_reg"], y[0])
            put(buf["y_cls"], y[1])
            put(buf["y_mask"], y[2])
            scale = 1.0
            if world > 1:
                # bucketed exchange: the all-reduce of gradient bucket k runs on NCCL's stream while segment k + 1 of
                # the backward pass runs on this one (async_op=True: NCCL's stream waits for the work queued so far)
                eng.train_count()
                counts_work = None
                if self.dp_normalizer == 'global':
                    # a few bytes, but a full NCCL latency: in flight during the forward pass (segment 0), awaited
                    # before the loss kernels (the first launches of segment 1)
                    counts_work = dist.all_reduce(eng.train_counts_tensor(), async_op=True)
                else:
                    scale = 1.0 / world
                grads, works = eng.grads_tensor(), []
                p2p = self._p2p_exchange(eng, grads)
                for i, (off, cnt, _) in enumerate(eng.train_segments()):
                    eng.train_run_segment(i)
                    if i == 0 and counts_work is not None:
                        counts_work.wait()
                    if cnt > 0:
                        if p2p is not None:
                            p2p.push(grads, off, cnt)      # copy engines, no SM
                        else:
                            works.append(dist.all_reduce(grads[off:off + cnt], async_op=True))
                if p2p is not None:
                    p2p.finish(grads)                      # "all pushes landed" signal + one sum kernel
                for w in works:
                    w.wait()
            else:
                eng.train_forward_backward()
            eng.train_apply(opt.lr, opt.beta_1, opt.beta_2, opt.epsilon, opt.clipnorm, scale)
        self._weights_stale = True
        self._wver = self.__dict__.get('_wver', 0) + 1
        if not sync_losses:
            return None
        l = eng.train_losses()
        self.last_step = l
        return [l["3Dbox"] + l["cls"] + l["mask"] + l["l2"], l["3Dbox"], l["cls"], l["mask"]]

    def fit_generator(self, generator, steps_per_epoch=None, epochs=1, verbose=1, callbacks=None,
                      validation_data=None, validation_steps=None, class_weight=None, max_queue_size=10, workers=1,
                      use_multiprocessing=False, shuffle=True, initial_epoch=0):
        """ The reference's training loop (PyraPose/bin/train.py:381-390) around train_on_batch: `generator` is a
        keras.utils.Sequence-like object (`__len__`, `__getitem__` -> (inputs, [regression_3D, labels, mask]),
        `on_epoch_end`) or a plain Python generator; arrays may be numpy or CUDA tensors (inject
        `utils.anchors.anchor_targets_bbox_device` as the Generator's `compute_anchor_targets`,
        preprocessing/generator.py:58-59, and the targets never leave the GPU). The next batch is staged by a
        background thread while the GPU runs the current step, and per-batch losses reach the callbacks one launch
        late (models/_fit.py). Returns a History (`.history['loss']`, `'3Dbox_loss'`, `'cls_loss'`, `'mask_loss'`)."""
        from ._fit import fit_generator
        return fit_generator(self, generator, steps_per_epoch=steps_per_epoch, epochs=epochs, verbose=verbose,
                             callbacks=callbacks, validation_data=validation_data, validation_steps=validation_steps,
                             class_weight=class_weight, max_queue_size=max_queue_size, workers=workers,
                             use_multiprocessing=use_multiprocessing, shuffle=shuffle, initial_epoch=initial_epoch)

    def _p2p_exchange(self, eng, grads):
        """The engine's P2PExchange (created collectively on the first data-parallel step), or None for NCCL."""
        import torch
        import torch.distributed as dist
        ex = eng.__dict__.get('_p2p', False)
        if ex is not False:
            return ex
        from ..utils import exchange as xch
        want = getattr(self, 'exchange', 'auto')
        ex = None
        if want == 'p2p' and xch.available():
            ex = xch.P2PExchange(grads.numel(), grads.device)   # collective; never raises on one replica alone
            # every replica must take the same path (this all-reduce is also the barrier before the first push)
            ok = torch.tensor([1.0 if ex.ok else 0.0], device=grads.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if float(ok.item()) < 1.0:
                err = ex.error
                ex.close()
                ex = None
                if want == 'p2p':
                    raise RuntimeError("exchange='p2p': peer-to-peer staging could not be set up (%r)" % (err,))
                import sys
                sys.stderr.write("pyrapose_b200: peer-to-peer gradient exchange unavailable (%r): using NCCL\n" % (err,))
        elif want == 'p2p':
            raise RuntimeError("exchange='p2p' needs torch.distributed with all replicas on this node")
        eng._p2p = ex
        return ex

    def sync_weights(self):
        """Copies the trained master weights from the engine back into `self.weights` (Keras HWIO layout)."""
        if self._parent is not None:
            return self._parent.sync_weights()   # the prediction model shares the training model's dict
        eng = getattr(self, '_train_engine', None)
        if eng is None or not getattr(self, '_weights_stale', False):
            return
        for l in self.layers:
            k = eng.train_layer_tensor("weights", l.name, l.ksize, l.cin, l.cout)
            if k is None:
                continue
            self.weights[l.name + "/kernel"] = k
            if l.use_bias:
                self.weights[l.name + "/bias"] = eng.train_layer_tensor("weights", l.name, l.ksize, l.cin, l.cout, bias=True)
        self._weights_stale = False


def retinanet(inputs=None, backbone_layers=None, num_classes=None, num_anchors=None, create_pyramid_features=None,
              submodels=None, name='retinanet', seed=None, device=None, precision='bf16'):
    """ Construct a PFPN training model (outputs '3Dbox', 'cls', 'mask'). `inputs`/`backbone_layers` are accepted for
    signature compatibility; the ResNet-50 backbone and __create_sparceFPN are built into the engine."""
    if num_classes is None:
        raise ValueError("num_classes is required")
    if create_pyramid_features is not None or submodels is not None:
        raise NotImplementedError("custom pyramid builders / submodels are not part of the hot path")
    return PFPNModel(num_classes, num_anchors, name=name, seed=seed, device=device, precision=precision)


def retinanet_bbox(model=None, nms=True, class_specific_filter=True, name='retinanet-bbox', anchor_params=None,
                   precision=None, alias_arena=False, **kwargs):
    """ Inference model: training model + Anchors + RegressBoxes3D. Like the reference (:302-335), `nms` and
    `class_specific_filter` are accepted and not wired into the graph; use layers.FilterDetections or
    model.detect_on_batch for filtering."""
    if anchor_params is None:
        anchor_params = AnchorParameters.default
    if model is None:
        model = retinanet(num_anchors=anchor_params.num_anchors(), **kwargs)
    else:
        assert_training_model(model)
    out = PFPNModel(model.num_classes, model.num_anchors, name=name, inference=True, anchor_params=anchor_params,
                    weights=model.weights, device=model.device,
                    precision=precision if precision is not None else model.precision, alias_arena=alias_arena)
    # Like the reference (bin/train.py:86-92: prediction_model = retinanet_bbox(model=model) is built BEFORE training
    # and evaluated after every epoch), the prediction model shares the training model's layers: same weight dict,
    # same engines. The Adam kernel rewrites the bf16 forward matrices in place, so a predict_on_batch after
    # train_on_batch runs on the trained weights with no copy.
    out._engines = model._engines
    out._parent = model
    return out
