"""Keras HDF5 weight files <-> the engine's weight dict.

What the reference reads and writes:
  model.load_weights(weights, by_name=True, skip_mismatch=...)     PyraPose/bin/train.py:65-68
  keras-resnet's ImageNet file ResNet-50-model.keras.h5            PyraPose/models/resnet.py:89-98
  ModelCheckpoint snapshots '{backbone}_{dataset}_{epoch:02d}.h5'  PyraPose/bin/train.py:131-142 (full-model save)
  models.load_model(path) -> keras.models.load_model               PyraPose/models/__init__.py:68-71

Layout Keras 2.3.1 writes (keras/engine/saving.py save_weights_to_hdf5_group): root (or the 'model_weights' group of a
full-model file) carries attrs `layer_names`, `backend`, `keras_version`; every layer is a group with attr
`weight_names` (e.g. b'conv1/kernel:0') and its arrays at <layer>/<weight name>. A nested sub-model (the regression /
classification / mask heads) is ONE layer group whose weight_names list all of its layers' variables in layer order.

THE NAME MAP. The reference names very little:
  named      keras-resnet layers (conv1, bn_conv1, res2a_branch2a, bn2a_branch2a, ...), the pyramid outputs 'P3', 'P4',
             'P5' (PyraPose/models/retinanet.py:204,209,212) and the mask sub-model 'mask' (:98)
  auto-named everything else: the head convolutions (their names are commented out at retinanet.py:34,44,122,126), the
             lateral / mid / down convolutions of __create_sparceFPN (:192-202, 207), the two unnamed sub-models.
Keras numbers auto-named layers per class in CREATION order, which for `backbone.retinanet(num_classes)` in a fresh
process is (retinanet.py:260-299: default_submodels -> mask head -> pyramid features):
  model_1 (3Dbox)  conv2d_1..4 = pyramid_regression3D_0..3, conv2d_5 = pyramid_regression3D      (:101-131)
  model_2 (cls)    conv2d_6..9 = pyramid_classification_0..3, conv2d_10 = pyramid_classification (:9-54)
  mask             conv2d_11..14 = mask_0..3, conv2d_15 = mask_out                               (:57-98)
  top level        conv2d_16 = P3_lateral, 17 = P4_lateral, 18 = P5_lateral (:192-194), 19 = P4_mid (:199),
                   20 = P3_mid (:201), 21 = P3_down (:202), 22 = P4_down (:207)
A process that built other models first shifts the numbers, so loading does not trust them: it maps by GRAPH ORDER and
SHAPE — sub-models are told apart by their kernel shapes (512-wide towers = regression, output width A*C = cls,
C = mask), their convolutions are taken in weight_names order; the top-level auto-named convolutions are sorted by
their number, the three 1x1 laterals are identified by their input width (512 / 1024 / 2048) and the four 3x3
256->256 ones are taken in creation order P4_mid, P3_mid, P3_down, P4_down. Saving writes the fresh-process numbers
above, which is what `model.load_weights(path, by_name=True)` of the reference expects.
"""
import re

import numpy as np

from ..utils import hdf5

_BN_PARAMS = ("gamma", "beta", "moving_mean", "moving_variance")
_AUTO = re.compile(r"^(conv2d|model|separable_conv2d)_(\d+)$")


def is_hdf5(path):
    try:
        with open(path, "rb") as f:
            head = f.read(8 + 2048)
    except (IOError, OSError):
        return False
    return head[:8] == hdf5.SIGNATURE or head[512:520] == hdf5.SIGNATURE or head[1024:1032] == hdf5.SIGNATURE


def keras_layer_names(num_classes=None, num_anchors=9):
    """engine layer name -> (top-level Keras layer / group name, Keras layer name the variable belongs to), with the
    fresh-process auto-numbering documented above."""
    del num_classes, num_anchors
    out = {}
    for i in range(4):
        out["pyramid_regression3D_%d" % i] = ("model_1", "conv2d_%d" % (1 + i))
        out["pyramid_classification_%d" % i] = ("model_2", "conv2d_%d" % (6 + i))
        out["mask_%d" % i] = ("mask", "conv2d_%d" % (11 + i))
    out["pyramid_regression3D"] = ("model_1", "conv2d_5")
    out["pyramid_classification"] = ("model_2", "conv2d_10")
    out["mask_out"] = ("mask", "conv2d_15")
    for k, name in enumerate(("P3_lateral", "P4_lateral", "P5_lateral", "P4_mid", "P3_mid", "P3_down", "P4_down")):
        out[name] = ("conv2d_%d" % (16 + k), "conv2d_%d" % (16 + k))
    return out


def _attr_list(attrs, name):
    """Keras splits attribute lists that exceed the 64 KB object-header limit into name0, name1, ..."""
    if name in attrs:
        vals = attrs[name]
    else:
        vals, k = [], 0
        while "%s%d" % (name, k) in attrs:
            vals += list(np.atleast_1d(attrs["%s%d" % (name, k)]))
            k += 1
    return [v.decode("utf-8") if isinstance(v, bytes) else str(v) for v in np.atleast_1d(vals)]


def read_layers(path):
    """-> [(layer name, [(weight name, float32 array)])] in file (= model layer) order, + root attrs."""
    root = hdf5.read_h5(path)
    g = root["model_weights"] if "model_weights" in root else root
    names = _attr_list(g.attrs, "layer_names") or [k for k, v in g.items() if isinstance(v, hdf5.Group)]
    layers = []
    for ln in names:
        if ln not in g:
            continue
        lg = g[ln]
        wn = _attr_list(lg.attrs, "weight_names")
        layers.append((ln, [(w, np.asarray(lg.get_path(w).data, dtype=np.float32)) for w in wn]))
    return layers, root.attrs


def _param(weight_name):
    return weight_name.split("/")[-1].split(":")[0]


def to_engine_weights(layers, known):
    """layers: read_layers() output; known: the engine's weight keys ('<layer>/<param>'). Returns ({key: array},
    [names of file layers that were not recognised])."""
    known_layers = set(k.split("/")[0] for k in known)
    out, unused = {}, []
    auto_top, submodels = [], []
    for ln, ws in layers:
        if not ws:
            continue
        if ln in known_layers:                                   # keras-resnet names, P3 / P4 / P5, or our own names
            for w, a in ws:
                out["%s/%s" % (ln, _param(w))] = a
            continue
        sub = {}
        for w, a in ws:                                          # variables grouped by the layer they belong to
            owner = w.split("/")[-2] if w.count("/") >= 1 else ln
            sub.setdefault(owner, []).append((_param(w), a))
        if len(sub) > 1 or (len(sub) == 1 and list(sub)[0] != ln):      # a nested sub-model
            convs = [(o, dict(p)) for o, p in sub.items() if "kernel" in dict(p)]
            if len(convs) != 5:
                unused.append(ln)
                continue
            submodels.append((ln, convs))
            continue
        m = _AUTO.match(ln)
        if m and "kernel" in dict(sub[ln]):
            auto_top.append((int(m.group(2)), ln, dict(sub[ln])))
        else:
            unused.append(ln)
    # sub-models: 512-wide towers = 3D regression (retinanet.py:101-131); of the two 256-wide ones the mask head is the
    # one named 'mask' (:98) — or, if a file renamed it, the one with the narrower output (C < A * C)
    narrow = [sm for sm in submodels if sm[1][0][1]["kernel"].shape[-1] != 512]
    mask_ln = None
    if any(ln == "mask" for ln, _ in narrow):
        mask_ln = "mask"
    elif len(narrow) == 2:
        mask_ln = min(narrow, key=lambda sm: sm[1][-1][1]["kernel"].shape[-1])[0]
    for ln, convs in submodels:
        if convs[0][1]["kernel"].shape[-1] == 512:
            names = ["pyramid_regression3D_%d" % i for i in range(4)] + ["pyramid_regression3D"]
        elif ln == mask_ln:
            names = ["mask_%d" % i for i in range(4)] + ["mask_out"]
        else:
            names = ["pyramid_classification_%d" % i for i in range(4)] + ["pyramid_classification"]
        for nm, (_o, p) in zip(names, convs):
            for k, a in p.items():
                out["%s/%s" % (nm, k)] = a
    # top-level auto-named convolutions of __create_sparceFPN, in creation order
    auto_top.sort()
    mids = []
    for _n, ln, p in auto_top:
        k = p["kernel"]
        if k.shape[:2] == (1, 1) and k.shape[2] in (512, 1024, 2048) and k.shape[3] == 256:
            nm = {512: "P3_lateral", 1024: "P4_lateral", 2048: "P5_lateral"}[k.shape[2]]
            for kk, a in p.items():
                out["%s/%s" % (nm, kk)] = a
        elif k.shape == (3, 3, 256, 256):
            mids.append(p)
        else:
            unused.append(ln)
    if mids:
        if len(mids) != 4:
            raise ValueError("expected the four unnamed 3x3 convolutions of __create_sparceFPN (P4_mid, P3_mid, P3_down, "
                             "P4_down), found %d auto-named 3x3 256->256 layers" % len(mids))
        for nm, p in zip(("P4_mid", "P3_mid", "P3_down", "P4_down"), mids):
            for kk, a in p.items():
                out["%s/%s" % (nm, kk)] = a
    return out, unused


def infer_classes(layers, num_anchors=9):
    """num_classes from the head shapes of a weight file (for models.load_model on a reference snapshot)."""
    for ln, ws in layers:
        kernels = [a for w, a in ws if _param(w) == "kernel"]
        if ln == "mask" and kernels:
            return int(kernels[-1].shape[-1])
    for ln, ws in layers:
        kernels = [a for w, a in ws if _param(w) == "kernel"]
        if len(kernels) == 5 and kernels[0].shape[-1] == 256 and kernels[-1].shape[-1] % num_anchors == 0:
            return int(kernels[-1].shape[-1] // num_anchors)
    raise ValueError("cannot infer the number of classes: no classification / mask sub-model in the file")


def load_h5(path, known):
    layers, _ = read_layers(path)
    return to_engine_weights(layers, known)


def save_h5(path, weights, layers_spec, meta=None, full_model=False):
    """Writes the weight dict in Keras' layout with the reference's layer names. layers_spec: models.spec.all_layers().
    full_model=True nests everything under 'model_weights' like keras.Model.save (ModelCheckpoint)."""
    kmap = keras_layer_names()
    root = hdf5.Group()
    g = root
    if full_model:
        g = root["model_weights"] = hdf5.Group()
    order = []

    def layer_group(top):
        if top not in g:
            g[top] = hdf5.Group()
            g[top].attrs["weight_names"] = []
            order.append(top)
        return g[top]

    def put(top, owner, param, arr):
        lg = layer_group(top)
        sub = lg.setdefault(owner, hdf5.Group())
        sub[param + ":0"] = hdf5.Dataset(np.ascontiguousarray(arr, dtype=np.float32))
        lg.attrs["weight_names"].append(("%s/%s:0" % (owner, param)).encode("utf-8"))

    # model layer order: backbone (conv, then its BatchNorm), sub-models are created before the pyramid but the
    # Keras layer list is topological: pyramid convolutions come before the sub-model layers
    for l in layers_spec:
        top, owner = kmap.get(l.name, (l.name, l.name))
        if top in ("model_1", "model_2", "mask"):
            continue
        put(top, owner, "kernel", weights[l.name + "/kernel"])
        if l.use_bias:
            put(top, owner, "bias", weights[l.name + "/bias"])
        if l.bn is not None:
            for p in _BN_PARAMS:
                put(l.bn, l.bn, p, weights["%s/%s" % (l.bn, p)])
    for l in layers_spec:
        top, owner = kmap.get(l.name, (l.name, l.name))
        if top not in ("model_1", "model_2", "mask"):
            continue
        put(top, owner, "kernel", weights[l.name + "/kernel"])
        put(top, owner, "bias", weights[l.name + "/bias"])
    for top in order:
        g[top].attrs["weight_names"] = np.array(g[top].attrs["weight_names"])
    g.attrs["layer_names"] = np.array([n.encode("utf-8") for n in order])
    g.attrs["backend"] = b"tensorflow"
    g.attrs["keras_version"] = b"2.3.1"
    if meta is not None:
        root.attrs["pyrapose_b200_meta"] = np.asarray(meta, dtype=np.int64)
    hdf5.write_h5(path, root)
