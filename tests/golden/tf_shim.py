"""numpy stand-ins for the few `tensorflow` / `keras` calls the reference's hot-path modules make — TEST INFRASTRUCTURE,
used only by tests/golden/make_golden_shim.py (in the build container, where /root/reference exists) to IMPORT AND RUN
the reference's own Python modules

    PyraPose/backend/common.py, backend/tf_backend.py, losses.py, layers/_misc.py, layers/filter_detections.py,
    utils/anchors.py

without TensorFlow 2.1.1 / Keras 2.3.1 (not installable here). The reference code is executed unmodified through the
normal import system; what is restated HERE is only the behaviour of the library primitives it calls, each a thin
numpy call on float32 / int64 arrays:

  elementwise arithmetic, stack / reshape / transpose / tile / concatenate, where, gather, gather_nd, pad, arange,
  meshgrid, shape, cast, max / argmax / sum / abs / pow / maximum / minimum, comparisons                 (one numpy call)
  keras.backend.binary_crossentropy   Keras 2.3.1 tensorflow_backend: clip to [eps, 1 - eps], eps = 1e-7, convert to
                                      logits log(p / (1 - p)), tf.nn.sigmoid_cross_entropy_with_logits
                                      = max(x, 0) - x z + log1p(exp(-|x|))
  keras.losses.mean_absolute_error    mean(|a - b|, axis=-1)
  tf.nn.top_k                         descending, ties -> lower index first
  tf.image.non_max_suppression        V3: candidates in descending score, suppress when IoU > threshold, IoU from
                                      min/max-normalised corners, 0 when an area is <= 0, float32
  tf.image.resize NEAREST             TF 2.x nearest kernel with half-pixel centres — see note
  tf.map_fn                           a Python loop over axis 0
  transforms3d.quaternions.quat2mat   transforms3d 0.3.1 (the published algorithm; the package is not installed)

So a golden vector made this way pins everything the reference's OWN text decides — operand order, which anchor
coordinate a corner is measured from, the class loop and concatenation order of filter_detections, the padding, the
normalisers of the losses, the 24 edge differences of orthogonal_l1, the whole of anchor_targets_bbox — and leaves
unpinned only the primitives listed above (for which the oracle has its own cross-checks: torchvision NMS / focal).

Note on resize: the reference's UpsampleLike calls tensorflow.image.resize(images, size, method, align_corners) with
TF 2.1, whose NEAREST kernel uses half_pixel_centers=True: src = min(floor((i + 0.5) * in / out), in - 1). That is what
this shim implements (and what oracle/pfpn.py and the CUDA kernel implement)."""
import sys
import types

import numpy as np

F32 = np.float32


class T(np.ndarray):
    """ndarray standing in for a tensor: tolerates the static-shape calls the reference makes, and converts SCALAR
    operands (Python numbers, numpy scalars such as the float64 entries of RegressBoxes3D's `std` array) to the
    tensor's own dtype the way TensorFlow's operator overloads do (convert_to_tensor with the tensor's dtype as hint),
    where plain numpy would promote a float32 array times a numpy float64 scalar to float64."""

    def set_shape(self, shape):
        assert tuple(int(s) for s in shape) == tuple(self.shape), (shape, self.shape)

    def __array_ufunc__(self, ufunc, method, *inputs, **kwargs):
        dt = next(x.dtype for x in inputs if isinstance(x, T))
        conv = []
        for x in inputs:
            if isinstance(x, T):
                conv.append(x.view(np.ndarray))
            elif isinstance(x, (bool, int, float, np.generic)) or (isinstance(x, np.ndarray) and x.ndim == 0):
                same_kind = dt.kind == "f" or np.asarray(x).dtype.kind != "f"
                conv.append(np.asarray(x, dtype=dt) if same_kind else x)
            else:
                conv.append(x)
        if "out" in kwargs:
            kwargs["out"] = tuple(o.view(np.ndarray) if isinstance(o, T) else o for o in kwargs["out"])
        out = getattr(ufunc, method)(*conv, **kwargs)
        return t(out) if isinstance(out, np.ndarray) else out


def t(a, dtype=None):
    return np.asarray(a, dtype=dtype).view(T)


def _dtype(d):
    if d is None:
        return None
    return np.dtype(getattr(d, "name", d))


# ------------------------------------------------------------------ tensorflow
def _where(condition, x=None, y=None):
    if x is None and y is None:
        return t(np.argwhere(np.asarray(condition)).astype(np.int64))
    return t(np.where(np.asarray(condition), x, y))


def _gather_nd(params, indices):
    indices = np.asarray(indices)
    return t(np.asarray(params)[tuple(indices[..., k] for k in range(indices.shape[-1]))])


def _top_k(values, k=1, sorted=True):
    values = np.asarray(values)
    k = int(k)
    order = np.argsort(-values, kind="stable")[:k]          # descending, ties: lower index first
    return t(values[order]), t(order.astype(np.int32))


def _iou(b, i, j):
    ymin_i, xmin_i = min(b[i, 0], b[i, 2]), min(b[i, 1], b[i, 3])
    ymax_i, xmax_i = max(b[i, 0], b[i, 2]), max(b[i, 1], b[i, 3])
    ymin_j, xmin_j = min(b[j, 0], b[j, 2]), min(b[j, 1], b[j, 3])
    ymax_j, xmax_j = max(b[j, 0], b[j, 2]), max(b[j, 1], b[j, 3])
    area_i = F32(ymax_i - ymin_i) * F32(xmax_i - xmin_i)
    area_j = F32(ymax_j - ymin_j) * F32(xmax_j - xmin_j)
    if area_i <= 0 or area_j <= 0:
        return F32(0)
    ih = max(F32(min(ymax_i, ymax_j) - max(ymin_i, ymin_j)), F32(0))
    iw = max(F32(min(xmax_i, xmax_j) - max(xmin_i, xmin_j)), F32(0))
    inter = F32(ih * iw)
    return F32(inter / F32(F32(area_i + area_j) - inter))


def _non_max_suppression(boxes, scores, max_output_size, iou_threshold=0.5, score_threshold=float("-inf")):
    b = np.asarray(boxes, dtype=F32)
    s = np.asarray(scores, dtype=F32)
    order = np.argsort(-s, kind="stable")
    keep = []
    thr = F32(iou_threshold)
    for i in order:
        if len(keep) >= int(max_output_size):
            break
        if s[i] <= score_threshold:
            break
        if all(not (_iou(b, i, j) > thr) for j in keep):
            keep.append(int(i))
    return t(np.array(keep, dtype=np.int32))


def _resize(images, size, method="nearest", *unused, **unused_kw):
    assert method == "nearest", "only the method the path uses"
    images = np.asarray(images)
    in_h, in_w = images.shape[1:3]
    out_h, out_w = int(size[0]), int(size[1])
    ys = np.minimum(np.floor((np.arange(out_h) + 0.5) * (in_h / out_h)).astype(np.int64), in_h - 1)
    xs = np.minimum(np.floor((np.arange(out_w) + 0.5) * (in_w / out_w)).astype(np.int64), in_w - 1)
    return t(images[:, ys][:, :, xs])


def _map_fn(fn, elems, dtype=None, parallel_iterations=None, **kw):
    def item(e, i):
        return [item(x, i) for x in e] if isinstance(e, (list, tuple)) else t(np.asarray(e)[i])

    first = elems
    while isinstance(first, (list, tuple)):
        first = first[0]
    outs = [fn(item(elems, i)) for i in range(np.asarray(first).shape[0])]
    return [t(np.stack([np.asarray(o[k]) for o in outs])) for k in range(len(outs[0]))]


def make_tensorflow():
    tf = types.ModuleType("tensorflow")
    tf.ones = lambda shape, dtype="float32": t(np.ones(tuple(int(s) for s in shape), dtype=_dtype(dtype)))
    tf.transpose = lambda a, perm=None: t(np.transpose(np.asarray(a), perm))
    tf.map_fn = _map_fn
    tf.pad = lambda a, paddings, mode="CONSTANT", constant_values=0: t(np.pad(
        np.asarray(a), [(int(p[0]), int(p[1])) for p in paddings], constant_values=constant_values))
    tf.clip_by_value = lambda a, lo, hi: t(np.minimum(np.maximum(np.asarray(a), lo), hi))
    tf.range = lambda *a, **k: t(np.arange(*a, **{kk: _dtype(v) if kk == "dtype" else v for kk, v in k.items()}))
    tf.gather_nd = _gather_nd
    tf.where = _where
    tf.meshgrid = lambda *a, **k: [t(m) for m in np.meshgrid(*[np.asarray(x) for x in a], **k)]
    tf.nn = types.SimpleNamespace(top_k=_top_k)
    tf.image = types.SimpleNamespace(
        non_max_suppression=_non_max_suppression, resize=_resize,
        ResizeMethod=types.SimpleNamespace(BILINEAR="bilinear", NEAREST_NEIGHBOR="nearest", BICUBIC="bicubic",
                                           AREA="area"))

    def scatter_nd(indices, updates, shape):
        out = np.zeros(tuple(int(s) for s in shape), dtype=np.asarray(updates).dtype)
        np.add.at(out, tuple(np.asarray(indices).T), np.asarray(updates))
        return t(out)

    tf.scatter_nd = scatter_nd
    return tf


# ------------------------------------------------------------------ keras
def _binary_crossentropy(target, output, from_logits=False):
    target = np.asarray(target, dtype=F32)
    output = np.asarray(output, dtype=F32)
    if not from_logits:
        eps = F32(1e-7)
        output = np.minimum(np.maximum(output, eps), F32(1) - eps)
        output = np.log(output / (F32(1) - output))
    x, z = output, target
    return t(np.maximum(x, F32(0)) - x * z + np.log1p(np.exp(-np.abs(x))))


class Layer(object):
    """keras.layers.Layer, as far as the reference's layer classes use it."""

    def __init__(self, **kwargs):
        self.name = kwargs.get("name")

    def __call__(self, inputs, **kwargs):
        return self.call(inputs, **kwargs)

    def get_config(self):
        return {"name": self.name}


def make_keras():
    keras = types.ModuleType("keras")
    K = types.ModuleType("keras.backend")
    K.floatx = lambda: "float32"
    K.image_data_format = lambda: "channels_last"
    K.shape = lambda a: np.array(np.asarray(a).shape, dtype=np.int64)
    K.int_shape = lambda a: tuple(int(s) for s in np.asarray(a).shape)
    K.stack = lambda xs, axis=0: t(np.stack([np.asarray(x) for x in xs], axis=axis))
    K.concatenate = lambda xs, axis=-1: t(np.concatenate([np.asarray(x) for x in xs], axis=axis))
    K.arange = lambda start, stop=None, step=1, dtype="int32": t(np.arange(start, stop, step, dtype=_dtype(dtype)))
    K.constant = lambda v, dtype=None, shape=None: t(np.asarray(v, dtype=_dtype(dtype) or F32))
    K.variable = lambda v, dtype=None, name=None: t(np.asarray(v, dtype=_dtype(dtype) or F32))
    K.reshape = lambda a, shape: t(np.reshape(np.asarray(a), tuple(int(s) for s in shape)))
    K.transpose = lambda a: t(np.transpose(np.asarray(a)))
    K.tile = lambda a, n: t(np.tile(np.asarray(a), tuple(int(s) for s in n)))
    K.expand_dims = lambda a, axis=-1: t(np.expand_dims(np.asarray(a), axis))
    K.cast = lambda a, dtype: t(np.asarray(a).astype(_dtype(dtype)))
    K.cast_to_floatx = lambda v: np.asarray(v, dtype=F32)
    K.not_equal = lambda a, b: t(np.not_equal(a, b))
    K.equal = lambda a, b: t(np.equal(a, b))
    K.greater = lambda a, b: t(np.greater(a, b))
    K.less = lambda a, b: t(np.less(a, b))
    K.ones_like = lambda a: t(np.ones_like(np.asarray(a)))
    K.binary_crossentropy = _binary_crossentropy
    K.sum = lambda a, axis=None: t(np.sum(np.asarray(a), axis=axis, dtype=np.asarray(a).dtype))
    K.max = lambda a, axis=None: t(np.max(np.asarray(a), axis=axis))
    K.argmax = lambda a, axis=-1: t(np.argmax(np.asarray(a), axis=axis).astype(np.int64))
    K.maximum = lambda a, b: t(np.maximum(a, b))
    K.minimum = lambda a, b: np.minimum(a, b)
    K.abs = lambda a: t(np.abs(np.asarray(a)))
    K.pow = lambda a, p: t(np.power(np.asarray(a), np.asarray(p, dtype=np.asarray(a).dtype)))
    K.gather = lambda ref, idx: t(np.asarray(ref)[np.asarray(idx)])
    keras.backend = K
    keras.losses = types.SimpleNamespace(
        mean_absolute_error=lambda a, b: t(np.mean(np.abs(np.asarray(b) - np.asarray(a)), axis=-1,
                                                   dtype=np.asarray(a).dtype)))
    layers = types.ModuleType("keras.layers")
    layers.Layer = Layer
    keras.layers = layers
    keras.initializers = types.SimpleNamespace(Initializer=object)
    keras.utils = types.SimpleNamespace(Sequence=object)
    return keras, K, layers


def quat2mat(q):
    """transforms3d.quaternions.quat2mat (transforms3d 0.3.1): (w, x, y, z) -> 3x3 rotation matrix, float64."""
    w, x, y, z = q
    Nq = w * w + x * x + y * y + z * z
    if Nq < np.finfo(np.float64).eps:
        return np.eye(3)
    s = 2.0 / Nq
    X, Y, Z = x * s, y * s, z * s
    wX, wY, wZ = w * X, w * Y, w * Z
    xX, xY, xZ = x * X, x * Y, x * Z
    yY, yZ, zZ = y * Y, y * Z, z * Z
    return np.array([[1.0 - (yY + zZ), xY - wZ, xZ + wY],
                     [xY + wZ, 1.0 - (xX + zZ), yZ - wX],
                     [xZ - wY, yZ + wX, 1.0 - (xX + yY)]])


def install(reference_root="/root/reference", compute_overlap_module=None):
    """Puts the stand-ins into sys.modules and makes `PyraPose` importable from `reference_root` WITHOUT running its
    package __init__ chains that need the real libraries (utils/__init__ pulls in the evaluation code). Returns the
    list of sys.modules keys to remove again (uninstall)."""
    import importlib.machinery
    import os
    sys.dont_write_bytecode = True            # /root/reference is read-only
    added = []

    def put(name, mod):
        if name not in sys.modules:
            added.append(name)
        sys.modules[name] = mod

    keras, K, layers = make_keras()
    tf = make_tensorflow()
    put("tensorflow", tf)
    put("keras", keras)
    put("keras.backend", K)
    put("keras.layers", layers)
    tf3d = types.ModuleType("transforms3d")
    tf3d.quaternions = types.SimpleNamespace(quat2mat=quat2mat)
    put("transforms3d", tf3d)
    if "imgaug" not in sys.modules:          # utils/image.py imports the augmentation library at module level only
        ia = types.ModuleType("imgaug")
        ia.augmenters = types.ModuleType("imgaug.augmenters")
        put("imgaug", ia)
        put("imgaug.augmenters", ia.augmenters)
    root = os.path.join(reference_root, "PyraPose")
    for name, sub in (("PyraPose", ""), ("PyraPose.utils", "utils")):   # layers/, backend/, models/ import normally
        pkg = types.ModuleType(name)
        pkg.__path__ = [os.path.join(root, sub)]
        pkg.__spec__ = importlib.machinery.ModuleSpec(name, None, is_package=True)
        put(name, pkg)
    if compute_overlap_module is not None:
        put("PyraPose.utils.compute_overlap", compute_overlap_module)
    return added


def uninstall(added):
    for name in list(sys.modules):
        if name in added or name.startswith("PyraPose."):
            sys.modules.pop(name, None)
    sys.modules.pop("PyraPose", None)
