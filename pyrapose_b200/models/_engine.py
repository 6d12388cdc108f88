"""Python handle of the native PFPN engine (ppb_engine_* in include/pyrapose_b200.h).

Owns the torch tensors the engine is bound to (input staging + outputs), so that predict_on_batch is:
pinned host -> device copy, one CUDA-graph launch, device -> host copy."""
import ctypes

import numpy as np
import torch

from .. import _native as nat

_f32p = ctypes.POINTER(ctypes.c_float)


def _fp(a):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a.ctypes.data_as(_f32p), a  # keep the array alive alongside the pointer


class Engine(object):
    def __init__(self, device, max_batch, height, width, num_classes, num_anchors, precision=0, alias_arena=False):
        self.lib = nat.load()
        self.lib.ppb_engine_device_bytes.restype = ctypes.c_size_t
        self.device = torch.device("cuda", device if device is not None else torch.cuda.current_device())
        self.max_batch = int(max_batch)
        self.height, self.width = int(height), int(width)
        self.num_classes, self.num_anchors = int(num_classes), int(num_anchors)
        self._h = ctypes.c_void_p()
        self.precision = int(precision)   # 0: bf16 production path; 1: fp32 path (three bf16 parts, six MMAs)
        nat.check(self.lib.ppb_engine_create_ex(ctypes.byref(self._h), self.device.index, self.max_batch, self.height,
                                                self.width, self.num_classes, self.num_anchors, self.precision))
        self.alias_arena = bool(alias_arena)
        if self.alias_arena:   # production arena: backbone intermediates share buffers (ppb_engine_set_arena_alias)
            nat.check(self.lib.ppb_engine_set_arena_alias(self._h, 1))
        self.levels = [((self.height + (8 << i) - 1) // (8 << i), (self.width + (8 << i) - 1) // (8 << i))
                       for i in range(3)]
        self.num_locations = sum(h * w for h, w in self.levels)
        self.num_boxes = self.num_locations * self.num_anchors
        self._bound = None
        self.buffers = {}

    def close(self):
        if self._h:
            self.lib.ppb_engine_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ weights
    def layer_table(self):
        out = []
        name = ctypes.create_string_buffer(128)
        vals = [ctypes.c_int() for _ in range(6)]
        for i in range(self.lib.ppb_engine_num_layers(self._h)):
            nat.check(self.lib.ppb_engine_layer_info(self._h, i, name, 128, *[ctypes.byref(v) for v in vals]))
            out.append((name.value.decode(),) + tuple(v.value for v in vals))
        return out

    def load_weights(self, weights, bn_names, bn_eps=1e-5):
        """weights: {"<layer>/kernel": HWIO, "<layer>/bias", "<bn>/gamma|beta|moving_mean|moving_variance"};
        bn_names: {conv layer name -> its BatchNorm layer name}."""
        for (name, k, cin, cout, stride, has_bn, has_bias) in self.layer_table():
            kern = weights[name + "/kernel"]
            if tuple(kern.shape) != (k, k, cin, cout):
                raise ValueError("layer %s: kernel shape %s, expected %s" % (name, tuple(kern.shape), (k, k, cin, cout)))
            keep = []
            args = []
            for arr in ([kern, weights.get(name + "/bias") if has_bias else None] +
                        ([weights[bn_names[name] + s] for s in ("/gamma", "/beta", "/moving_mean", "/moving_variance")]
                         if has_bn else [None] * 4)):
                if arr is None:
                    args.append(None)
                else:
                    p, a = _fp(arr)
                    keep.append(a)
                    args.append(p)
            nat.check(self.lib.ppb_engine_set_layer(self._h, name.encode(), *args, ctypes.c_float(bn_eps)))
        self._bound = None

    def set_anchors(self, base_anchors_f32, strides, mean=None, std=None):
        b, keep_b = _fp(base_anchors_f32)
        s, keep_s = _fp(strides)
        m = _fp(mean) if mean is not None else (None, None)
        d = _fp(std) if std is not None else (None, None)
        nat.check(self.lib.ppb_engine_set_anchors(self._h, b, s, m[0], d[0]))
        self._bound = None

    # ------------------------------------------------------------------ execution
    def _set_input(self, input_u8):
        nat.check(self.lib.ppb_engine_set_input_u8(self._h, int(bool(input_u8)), None))

    def bind(self, batch, decode=True, want_mask=True, want_hull=False, input_u8=False):
        """input_u8: `images` holds raw uint8 BGR pixels; preprocess_image('caffe') runs inside the stem loader."""
        key = (int(batch), bool(decode), bool(want_mask), bool(want_hull), bool(input_u8))
        if self._bound == key:
            return self.buffers
        dev = self.device
        B, N, C = int(batch), self.num_boxes, self.num_classes
        mh, mw = self.levels[0]
        buf = {
            "images": torch.empty((B, self.height, self.width, 3), dtype=torch.uint8 if input_u8 else torch.float32,
                                  device=dev),
            "boxes3D": torch.empty((B, N, 16), dtype=torch.float32, device=dev),
            "cls": torch.empty((B, N, C), dtype=torch.float32, device=dev),
            "mask": torch.empty((B, mh * mw, C), dtype=torch.float32, device=dev) if want_mask else None,
            "hull": torch.empty((B, N, 4), dtype=torch.float32, device=dev) if want_hull else None,
        }
        with torch.cuda.device(dev):
            torch.cuda.current_stream().synchronize()
            self._set_input(input_u8)
            nat.check(self.lib.ppb_engine_bind(self._h, B, nat.ptr(buf["images"]), nat.ptr(buf["boxes3D"]),
                                               nat.ptr(buf["cls"]), nat.ptr(buf["mask"]), nat.ptr(buf["hull"]),
                                               int(decode), None))
        self.buffers = buf
        self._bound = key
        return buf

    def run(self, eager=False):
        with torch.cuda.device(self.device):
            fn = self.lib.ppb_engine_run_eager if eager else self.lib.ppb_engine_run
            nat.check(fn(self._h, nat.current_stream_ptr()))

    def run_part(self, part):
        """Part 0 / 1 / 2 of the forward pass (through the classification head / regression head / mask head); the
        three in order equal run(). See ppb_engine_run_part."""
        with torch.cuda.device(self.device):
            nat.check(self.lib.ppb_engine_run_part(self._h, int(part), nat.current_stream_ptr()))

    def kernels_per_run(self):
        return int(self.lib.ppb_engine_kernels_per_run(self._h))

    def device_bytes(self):
        return int(self.lib.ppb_engine_device_bytes(self._h))

    def activation(self, name):
        shape = (ctypes.c_int * 4)()
        nat.check(self.lib.ppb_engine_activation(self._h, name.encode(), None, shape))
        out = np.empty(tuple(shape), dtype=np.float32)
        nat.check(self.lib.ppb_engine_activation(self._h, name.encode(), out.ctypes.data_as(_f32p), shape))
        return out

    def profile(self):
        """[(kind, ms)] for every launch of one eager forward pass (kinds: 0 conv, 1 stem-s2d, 2 pool, 3 planes, 4 add)."""
        n = self.kernels_per_run()
        ms = (ctypes.c_float * n)()
        kinds = (ctypes.c_int * n)()
        with torch.cuda.device(self.device):
            nat.check(self.lib.ppb_engine_profile(self._h, nat.current_stream_ptr(), ms, kinds, n))
        return [(int(kinds[i]), float(ms[i])) for i in range(n)]

    def trace_op(self, index, cta=0, cap=4096):
        """Cycle-stamped trace of CTA `cta` of convolution launch `index` (ppb_engine_trace_op). Returns (info dict,
        {role: int64 array [n, 4] of (event, tile, sub, clock)})."""
        buf = torch.zeros(8 * cap, 2, dtype=torch.int32, device=self.device)
        info = (ctypes.c_int * 12)()
        with torch.cuda.device(self.device):
            nat.check(self.lib.ppb_engine_trace_op(self._h, int(index), int(cta), nat.ptr(buf), int(cap), info,
                                                   nat.current_stream_ptr()))
            torch.cuda.synchronize()
        raw = buf.cpu().numpy().astype(np.int64) & 0xFFFFFFFF
        keys = ("m_tiles", "n_tiles", "block_n", "taps", "cin_blocks", "stages", "ring_slots", "pair", "slab3", "bres",
                "staged", "epi_groups")
        roles = {}
        for r, name in enumerate(("producer", "mma", "residual", "epi0", "epi1", "epi2")):
            reg = raw[r * cap:(r + 1) * cap]
            n = int(reg[0, 0])
            if n == 0:
                continue
            tag, clk = reg[1:1 + n, 0], reg[1:1 + n, 1]
            roles[name] = np.stack([tag >> 24, (tag >> 8) & 0xFFFF, tag & 0xFF, clk], axis=1)
        return dict(zip(keys, [int(v) for v in info])), roles

    def op_table(self):
        """[(name, algorithmic flops)] per launch of the bound plan."""
        out = []
        name = ctypes.create_string_buffer(64)
        fl = ctypes.c_double()
        for i in range(self.kernels_per_run()):
            nat.check(self.lib.ppb_engine_op_info(self._h, i, name, 64, ctypes.byref(fl)))
            out.append((name.value.decode(), fl.value))
        return out

    # ------------------------------------------------------------------ training step (ppb_engine_train_*)
    def train_bind(self, batch, input_u8=False, bucket_bytes=0):
        """Allocates the step's device buffers (inputs, targets, the three model outputs) and captures
        forward + losses + backward as one CUDA graph — or, with bucket_bytes > 0, as one graph per gradient bucket
        (train_num_segments / train_run_segment) for the overlapped data-parallel exchange. Returns the buffer dict."""
        key = ("train", int(batch), bool(input_u8), int(bucket_bytes))
        if self._bound == key:
            return self.buffers
        dev = self.device
        B, N, C = int(batch), self.num_boxes, self.num_classes
        mh, mw = self.levels[0]
        f32 = dict(dtype=torch.float32, device=dev)
        buf = {
            "images": torch.empty((B, self.height, self.width, 3), dtype=torch.uint8 if input_u8 else torch.float32,
                                  device=dev),
            "y_reg": torch.zeros((B, N, 17), **f32),
            "y_cls": torch.zeros((B, N, C + 1), **f32),
            "y_mask": torch.zeros((B, mh * mw, C + 1), **f32),
            "3Dbox": torch.empty((B, N, 16), **f32),
            "cls": torch.empty((B, N, C), **f32),
            "mask": torch.empty((B, mh * mw, C), **f32),
        }
        with torch.cuda.device(dev):
            torch.cuda.current_stream().synchronize()
            self._set_input(input_u8)
            nat.check(self.lib.ppb_engine_train_set_bucket_bytes(self._h, ctypes.c_longlong(int(bucket_bytes))))
            nat.check(self.lib.ppb_engine_train_bind(self._h, B, nat.ptr(buf["images"]), nat.ptr(buf["y_reg"]),
                                                     nat.ptr(buf["y_cls"]), nat.ptr(buf["y_mask"]),
                                                     nat.ptr(buf["3Dbox"]), nat.ptr(buf["cls"]), nat.ptr(buf["mask"]),
                                                     None))
        self.buffers = buf
        self._bound = key
        self._grads = None
        self._counts = None
        return buf

    def train_count(self):
        """Positives of the three outputs -> the device counters (train_counts_tensor); targets only, no forward."""
        with torch.cuda.device(self.device):
            nat.check(self.lib.ppb_engine_train_count(self._h, nat.current_stream_ptr()))

    def train_counts_tensor(self):
        """The engine's int32[4] positive counters as a torch tensor sharing the device memory."""
        if getattr(self, "_counts", None) is None:
            p = ctypes.POINTER(ctypes.c_int)()
            nat.check(self.lib.ppb_engine_train_counts(self._h, ctypes.byref(p)))
            addr = ctypes.cast(p, ctypes.c_void_p).value

            class _Cnt(object):
                __cuda_array_interface__ = {"shape": (4,), "typestr": "<i4", "data": (int(addr), False),
                                            "version": 2, "strides": None}
            with torch.cuda.device(self.device):
                self._counts = torch.as_tensor(_Cnt(), device=self.device)
        return self._counts

    def train_segments(self):
        """[(grad_offset, grad_count, launches)] per segment graph of the bound step."""
        out = []
        off, cnt, nl = ctypes.c_longlong(), ctypes.c_longlong(), ctypes.c_int()
        for i in range(self.lib.ppb_engine_train_num_segments(self._h)):
            nat.check(self.lib.ppb_engine_train_segment_info(self._h, i, ctypes.byref(off), ctypes.byref(cnt),
                                                             ctypes.byref(nl)))
            out.append((off.value, cnt.value, nl.value))
        return out

    def train_run_segment(self, i):
        with torch.cuda.device(self.device):
            nat.check(self.lib.ppb_engine_train_run_segment(self._h, int(i), nat.current_stream_ptr()))

    def train_forward_backward(self, eager=False):
        with torch.cuda.device(self.device):
            nat.check(self.lib.ppb_engine_train_forward_backward(self._h, int(eager), nat.current_stream_ptr()))

    def grads_tensor(self):
        """The engine's flat fp32 gradient buffer as a torch tensor sharing the device memory (for NCCL all-reduce)."""
        if getattr(self, "_grads", None) is None:
            p = ctypes.c_void_p()
            n = ctypes.c_longlong()
            nat.check(self.lib.ppb_engine_train_grads(self._h, ctypes.byref(p), ctypes.byref(n)))

            class _Flat(object):
                __cuda_array_interface__ = {"shape": (int(n.value),), "typestr": "<f4", "data": (int(p.value), False),
                                            "version": 2, "strides": None}
            with torch.cuda.device(self.device):
                self._grads = torch.as_tensor(_Flat(), device=self.device)
        return self._grads

    def train_apply(self, lr, beta_1=0.9, beta_2=0.999, epsilon=1e-7, clipnorm=0.0, grad_scale=1.0):
        f = ctypes.c_float
        with torch.cuda.device(self.device):
            nat.check(self.lib.ppb_engine_train_apply(self._h, f(lr), f(beta_1), f(beta_2), f(epsilon), f(clipnorm),
                                                      f(grad_scale), nat.current_stream_ptr()))

    def train_losses(self):
        """Synchronises. -> dict 3Dbox / cls / mask / l2 / grad_norm and the positive counts."""
        out = (ctypes.c_double * 5)()
        cnt = (ctypes.c_int * 3)()
        with torch.cuda.device(self.device):
            nat.check(self.lib.ppb_engine_train_losses(self._h, out, cnt, nat.current_stream_ptr()))
        return {"3Dbox": out[0], "cls": out[1], "mask": out[2], "l2": out[3], "grad_norm": float(np.sqrt(out[4])),
                "positives": (cnt[0], cnt[1], cnt[2])}

    def train_losses_async(self, out5_pinned):
        """No synchronisation: queues the copy of [3Dbox, cls, mask, l2, grad_norm^2] (float64) into the PINNED host
        tensor `out5_pinned` on the current stream; the caller orders its read with an event."""
        if not (out5_pinned.is_pinned() and out5_pinned.dtype == torch.float64 and out5_pinned.numel() >= 5):
            raise ValueError("train_losses_async needs a pinned float64 host tensor of at least 5 elements")
        with torch.cuda.device(self.device):
            nat.check(self.lib.ppb_engine_train_losses_async(self._h, ctypes.c_void_p(out5_pinned.data_ptr()), None,
                                                             nat.current_stream_ptr()))

    def train_param_info(self, layer):
        ko, bo = ctypes.c_longlong(), ctypes.c_longlong()
        tr = ctypes.c_int()
        nat.check(self.lib.ppb_engine_train_param_info(self._h, layer.encode(), ctypes.byref(ko), ctypes.byref(bo),
                                                       ctypes.byref(tr)))
        return ko.value, bo.value, bool(tr.value)

    def train_read(self, which, offset, count):
        """which: 'grads' | 'weights' | 'm' | 'v' -> float32 numpy copy of [offset, offset+count) of that flat buffer."""
        idx = {"grads": 0, "weights": 1, "m": 2, "v": 3}[which]
        out = np.empty(int(count), dtype=np.float32)
        nat.check(self.lib.ppb_engine_train_read(self._h, idx, ctypes.c_longlong(offset), ctypes.c_longlong(count),
                                                 out.ctypes.data_as(_f32p)))
        return out

    def train_layer_tensor(self, which, name, ksize, cin, cout, bias=False):
        """One layer's kernel (as Keras HWIO) or bias from a flat training buffer."""
        ko, bo, _ = self.train_param_info(name)
        if bias:
            return self.train_read(which, bo, cout) if bo >= 0 else None
        if ko < 0:
            return None
        flat = self.train_read(which, ko, cout * ksize * ksize * cin)
        return np.ascontiguousarray(flat.reshape(cout, ksize, ksize, cin).transpose(1, 2, 3, 0))  # O,HWI -> HWIO

    def train_op_table(self):
        out = []
        name = ctypes.create_string_buffer(64)
        fl = ctypes.c_double()
        for i in range(self.lib.ppb_engine_train_num_ops(self._h)):
            nat.check(self.lib.ppb_engine_train_op_info(self._h, i, name, 64, ctypes.byref(fl)))
            out.append((name.value.decode(), fl.value))
        return out

    def train_profile(self):
        """[(op name, algorithmic flops, ms)] for one eager training step (forward, losses, backward)."""
        ops = self.train_op_table()
        ms = (ctypes.c_float * len(ops))()
        with torch.cuda.device(self.device):
            nat.check(self.lib.ppb_engine_train_profile(self._h, nat.current_stream_ptr(), ms, len(ops)))
        return [(n, f, float(ms[i])) for i, (n, f) in enumerate(ops)]
