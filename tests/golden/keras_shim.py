"""A miniature Keras functional API (define-by-run over concrete probe tensors, graphs re-executable on new inputs) —
TEST INFRASTRUCTURE, used only by tests/golden/make_golden_model.py in the build container to IMPORT AND RUN the
reference's own model-building code

    PyraPose/models/__init__.py (backbone), models/resnet.py (resnet_retinanet), models/retinanet.py
    (default_classification_model, default_mask_model, default_3Dregression_model, __create_sparceFPN,
    __build_pyramid, __build_anchors, retinanet, retinanet_bbox), initializers.py, layers/_misc.py

without Keras 2.3.1 / TensorFlow 2.1.1 / keras-resnet 0.1.0 (not installable here). It extends tests/golden/tf_shim.py.

What running the reference over it pins: the WIRING — which layer feeds which in the sparse pyramid and the three
heads, strides / kernel sizes / paddings / activations as the reference passes them, sub-models shared across pyramid
levels, the Reshape + Concatenate(axis=1) output order, Anchors + RegressBoxes3D of the prediction model — and Keras'
AUTO-NAMING of the layers the reference leaves unnamed (conv2d_<k>, model_<k> in creation order), i.e. the layer-name
map a real .h5 checkpoint is loaded by (pyrapose_b200/models/keras_io.py).

What is restated HERE (the library behaviour, not the reference's text):
  Conv2D            torch.nn.functional.conv2d on float32 CPU; padding='same' as TensorFlow computes it: out = ceil(in /
                    stride), total = max((out - 1) * stride + k - in, 0), before = total // 2 (the extra pixel goes
                    bottom / right)
  MaxPooling2D      'same': the same padding rule, padded cells never win (-inf)
  BatchNormalization (keras_resnet.layers, freeze=True -> inference mode): (x - mean) / sqrt(var + eps) * gamma + beta
  layer naming      keras.backend.get_uid per snake-cased class name, 'input_<k>' for Input, 'model_<k>' for Model
  keras_resnet.models.ResNet50   keras-resnet 0.1.0's published structure: ZeroPadding2D(3) + 7x7/2 conv + BN + ReLU +
                    3x3/2 'same' max-pool, then [3, 4, 6, 3] bottleneck_2d blocks (stride on the first 1x1 of the
                    first block of stages 3-5, ZeroPadding2D(1) + 'valid' 3x3, projection shortcut with BN in block 0,
                    Add then ReLU), outputs = the four stage outputs, layers in creation order
"""
import re
import sys
import types
from collections import defaultdict

import numpy as np

import tf_shim
from tf_shim import t

_uids = defaultdict(int)


def reset_uids():
    _uids.clear()


def get_uid(prefix):
    _uids[prefix] += 1
    return _uids[prefix]


def to_snake_case(name):
    intermediate = re.sub("(.)([A-Z][a-z0-9]+)", r"\1_\2", name)
    insecure = re.sub("([a-z])([A-Z])", r"\1_\2", intermediate).lower()
    return insecure if insecure[0] != "_" else "private" + insecure


class KT(object):
    """A tensor of the graph: its concrete probe value and where it came from."""

    def __init__(self, value, layer, inputs, index, listed):
        self.value = value
        self.layer, self.inputs, self.index, self.listed = layer, inputs, index, listed

    @property
    def shape(self):
        return tuple(self.value.shape)

    @property
    def dtype(self):
        return self.value.dtype


def _evaluate(out, memo):
    """Value of graph tensor `out` given memo {id(KT): value} for the substituted inputs."""
    if id(out) in memo:
        return memo[id(out)]
    if not out.inputs:
        raise RuntimeError("graph input %r was not fed" % (out.layer.name,))
    vals = [_evaluate(i, memo) for i in out.inputs]
    res = out.layer.call(vals if out.listed else vals[0])
    outs = res if isinstance(res, (list, tuple)) else [res]
    for sibling, v in zip(out.layer._node_outputs(out), outs):
        memo[id(sibling)] = v
    return memo[id(out)]


class Layer(object):
    def __init__(self, name=None, trainable=True, **kwargs):
        if name is None:
            prefix = to_snake_case(self.__class__.__name__)
            name = prefix + "_" + str(get_uid(prefix))
        self.name = name
        self.trainable = trainable
        self.weights = {}
        self._nodes = []

    def _node_outputs(self, kt):
        for node in self._nodes:
            if any(o is kt for o in node):
                return node
        raise KeyError(self.name)

    def __call__(self, inputs, **kwargs):
        listed = isinstance(inputs, (list, tuple))
        ins = list(inputs) if listed else [inputs]
        if not any(isinstance(x, KT) for x in ins):
            return self.call(inputs, **kwargs)            # plain arrays: behave like tf_shim's Layer
        vals = [x.value for x in ins]
        res = self.call(vals if listed else vals[0], **kwargs)
        outs = list(res) if isinstance(res, (list, tuple)) else [res]
        node = [KT(t(v), self, ins, i, listed) for i, v in enumerate(outs)]
        self._nodes.append(node)
        return node if isinstance(res, (list, tuple)) else node[0]

    @property
    def output(self):
        assert len(self._nodes) == 1, "layer %s has %d inbound nodes" % (self.name, len(self._nodes))
        return self._nodes[0][0] if len(self._nodes[0]) == 1 else self._nodes[0]

    def get_config(self):
        return {"name": self.name, "trainable": self.trainable}


class InputLayer(Layer):
    pass


def Input(shape=None, name=None, **kwargs):
    layer = InputLayer(name=name or "input_" + str(get_uid("input")))
    c = int(shape[-1])
    h = w = 64 if c == 3 else 8                          # probe size: the graphs are fully convolutional
    kt = KT(t(np.zeros((1, h, w, c), np.float32)), layer, [], 0, False)
    layer._nodes.append([kt])
    return kt


def _same_pad(size, k, stride):
    out = -(-size // stride)
    total = max((out - 1) * stride + k - size, 0)
    return total // 2, total - total // 2


def _pair(v):
    return (int(v), int(v)) if np.isscalar(v) else (int(v[0]), int(v[1]))


class Conv2D(Layer):
    def __init__(self, filters, kernel_size, strides=1, padding="valid", activation=None, use_bias=True,
                 kernel_initializer=None, bias_initializer=None, kernel_regularizer=None, name=None, **kwargs):
        super(Conv2D, self).__init__(name=name)
        self.filters, self.kernel_size, self.strides = int(filters), _pair(kernel_size), _pair(strides)
        self.padding, self.activation, self.use_bias = padding, activation, use_bias
        self.kernel_initializer, self.bias_initializer = kernel_initializer, bias_initializer
        self.kernel_regularizer = kernel_regularizer

    def call(self, x):
        import torch
        import torch.nn.functional as F
        x = np.asarray(x, dtype=np.float32)
        if "kernel" not in self.weights:
            self.weights["kernel"] = np.zeros(self.kernel_size + (x.shape[3], self.filters), np.float32)
            if self.use_bias:
                self.weights["bias"] = np.zeros(self.filters, np.float32)
        xt = torch.from_numpy(np.ascontiguousarray(x)).permute(0, 3, 1, 2)
        if self.padding == "same":
            pt, pb = _same_pad(x.shape[1], self.kernel_size[0], self.strides[0])
            pl, pr = _same_pad(x.shape[2], self.kernel_size[1], self.strides[1])
            xt = F.pad(xt, (pl, pr, pt, pb))
        else:
            assert self.padding == "valid"
        w = torch.from_numpy(self.weights["kernel"]).permute(3, 2, 0, 1).contiguous()
        b = torch.from_numpy(self.weights["bias"]) if self.use_bias else None
        y = F.conv2d(xt, w, b, stride=self.strides)
        if self.activation == "relu":
            y = torch.relu(y)
        elif self.activation == "sigmoid":
            y = torch.sigmoid(y)
        else:
            assert self.activation is None, self.activation
        return t(y.permute(0, 2, 3, 1).contiguous().numpy())


class ZeroPadding2D(Layer):
    def __init__(self, padding=1, name=None, **kwargs):
        super(ZeroPadding2D, self).__init__(name=name)
        self.padding = _pair(padding)

    def call(self, x):
        p = self.padding
        return t(np.pad(np.asarray(x), ((0, 0), (p[0], p[0]), (p[1], p[1]), (0, 0))))


class MaxPooling2D(Layer):
    def __init__(self, pool_size=(2, 2), strides=None, padding="valid", name=None, **kwargs):
        super(MaxPooling2D, self).__init__(name=name)
        self.pool_size, self.strides, self.padding = _pair(pool_size), _pair(strides or pool_size), padding

    def call(self, x):
        import torch
        import torch.nn.functional as F
        xt = torch.from_numpy(np.ascontiguousarray(np.asarray(x, np.float32))).permute(0, 3, 1, 2)
        if self.padding == "same":
            pt, pb = _same_pad(xt.shape[2], self.pool_size[0], self.strides[0])
            pl, pr = _same_pad(xt.shape[3], self.pool_size[1], self.strides[1])
            xt = F.pad(xt, (pl, pr, pt, pb), value=float("-inf"))
        y = F.max_pool2d(xt, self.pool_size, self.strides)
        return t(y.permute(0, 2, 3, 1).contiguous().numpy())


class BatchNormalization(Layer):
    """keras_resnet.layers.BatchNormalization(freeze=True): always inference mode."""

    def __init__(self, axis=3, epsilon=1e-3, freeze=False, name=None, **kwargs):
        super(BatchNormalization, self).__init__(name=name)
        assert axis in (3, -1)
        self.epsilon, self.freeze = float(epsilon), freeze

    def call(self, x):
        x = np.asarray(x, np.float32)
        if "gamma" not in self.weights:
            c = x.shape[3]
            self.weights.update(gamma=np.ones(c, np.float32), beta=np.zeros(c, np.float32),
                                moving_mean=np.zeros(c, np.float32), moving_variance=np.ones(c, np.float32))
        w = self.weights
        inv = (w["gamma"] / np.sqrt(w["moving_variance"] + np.float32(self.epsilon))).astype(np.float32)
        return t((x - w["moving_mean"]) * inv + w["beta"])


class Activation(Layer):
    def __init__(self, activation, name=None, **kwargs):
        super(Activation, self).__init__(name=name)
        self.activation = activation

    def call(self, x):
        x = np.asarray(x, np.float32)
        if self.activation == "relu":
            return t(np.maximum(x, np.float32(0)))
        assert self.activation == "sigmoid", self.activation
        import torch
        return t(torch.sigmoid(torch.from_numpy(np.ascontiguousarray(x))).numpy())


class Add(Layer):
    def call(self, xs):
        out = np.asarray(xs[0], np.float32)
        for x in xs[1:]:
            out = out + np.asarray(x, np.float32)
        return t(out)


class Reshape(Layer):
    def __init__(self, target_shape, name=None, **kwargs):
        super(Reshape, self).__init__(name=name)
        self.target_shape = tuple(target_shape)

    def call(self, x):
        x = np.asarray(x)
        return t(x.reshape((x.shape[0],) + self.target_shape))


class Concatenate(Layer):
    def __init__(self, axis=-1, name=None, **kwargs):
        super(Concatenate, self).__init__(name=name)
        self.axis = axis

    def call(self, xs):
        return t(np.concatenate([np.asarray(x) for x in xs], axis=self.axis))


class Model(Layer):
    def __init__(self, inputs=None, outputs=None, name=None, **kwargs):
        super(Model, self).__init__(name=name or "model_" + str(get_uid("model")))
        self._in_listed = isinstance(inputs, (list, tuple))
        self._out_listed = isinstance(outputs, (list, tuple))
        self.inputs = list(inputs) if self._in_listed else [inputs]
        self.outputs = list(outputs) if self._out_listed else [outputs]
        # layers in creation order of first use, inputs first (Keras sorts by graph depth; inside one residual block
        # the projection branch can sit elsewhere, which no caller on the path depends on)
        seen, order = set(), []

        def visit(kt):
            for i in kt.inputs:
                visit(i)
            if id(kt.layer) not in seen:
                seen.add(id(kt.layer))
                order.append(kt.layer)

        for o in self.outputs:
            visit(o)
        self.layers = order

    @property
    def output_names(self):
        return [o.layer.name for o in self.outputs]

    def get_layer(self, name=None, index=None):
        if index is not None:
            return self.layers[index]
        for layer in self.layers:
            if layer.name == name:
                return layer
        raise ValueError("No such layer: " + str(name))

    def call(self, xs):
        xs = list(xs) if isinstance(xs, (list, tuple)) else [xs]
        memo = {id(i): t(np.asarray(x)) for i, x in zip(self.inputs, xs)}
        outs = [_evaluate(o, memo) for o in self.outputs]
        return outs if self._out_listed else outs[0]

    def __call__(self, inputs, **kwargs):
        listed = isinstance(inputs, (list, tuple))
        ins = list(inputs) if listed else [inputs]
        if not any(isinstance(x, KT) for x in ins):
            return self.call(inputs)
        res = self.call([x.value for x in ins])
        outs = list(res) if isinstance(res, (list, tuple)) else [res]
        node = [KT(t(v), self, ins, i, True) for i, v in enumerate(outs)]
        self._nodes.append(node)
        return node if isinstance(res, (list, tuple)) else node[0]

    def predict_on_batch(self, x):
        res = self.call(x)
        return [np.array(np.asarray(r)) for r in res] if isinstance(res, list) else np.array(np.asarray(res))

    def load_weights(self, *args, **kwargs):     # resnet.py:98 loads the ImageNet file: weights are set by the caller
        return None

    def all_layers(self, prefix=""):
        """(path, layer) of every weighted layer, nested models flattened: path = 'model_1/conv2d_3'."""
        for layer in self.layers:
            if isinstance(layer, Model):
                for item in layer.all_layers(prefix + layer.name + "/"):
                    yield item
            elif layer.weights:
                yield prefix + layer.name, layer


def _keras_resnet():
    """keras-resnet 0.1.0 (not installed): models.ResNet50 and layers.BatchNormalization, see the module header."""
    pkg = types.ModuleType("keras_resnet")
    pkg.custom_objects = {"BatchNormalization": BatchNormalization}
    pkg.layers = types.ModuleType("keras_resnet.layers")
    pkg.layers.BatchNormalization = BatchNormalization
    pkg.models = types.ModuleType("keras_resnet.models")

    def bottleneck_2d(filters, stage, block, freeze_bn):
        stride = 1 if (block != 0 or stage == 0) else 2
        tag = "%d%s" % (stage + 2, chr(ord("a") + block))

        def f(x):
            y = Conv2D(filters, (1, 1), strides=stride, use_bias=False, name="res%s_branch2a" % tag)(x)
            y = BatchNormalization(axis=3, epsilon=1e-5, freeze=freeze_bn, name="bn%s_branch2a" % tag)(y)
            y = Activation("relu", name="res%s_branch2a_relu" % tag)(y)
            y = ZeroPadding2D(padding=1, name="padding%s_branch2b" % tag)(y)
            y = Conv2D(filters, (3, 3), use_bias=False, name="res%s_branch2b" % tag)(y)
            y = BatchNormalization(axis=3, epsilon=1e-5, freeze=freeze_bn, name="bn%s_branch2b" % tag)(y)
            y = Activation("relu", name="res%s_branch2b_relu" % tag)(y)
            y = Conv2D(filters * 4, (1, 1), use_bias=False, name="res%s_branch2c" % tag)(y)
            y = BatchNormalization(axis=3, epsilon=1e-5, freeze=freeze_bn, name="bn%s_branch2c" % tag)(y)
            if block == 0:
                s = Conv2D(filters * 4, (1, 1), strides=stride, use_bias=False, name="res%s_branch1" % tag)(x)
                s = BatchNormalization(axis=3, epsilon=1e-5, freeze=freeze_bn, name="bn%s_branch1" % tag)(s)
            else:
                s = x
            y = Add(name="res%s" % tag)([y, s])
            return Activation("relu", name="res%s_relu" % tag)(y)

        return f

    def ResNet50(inputs, blocks=None, include_top=True, classes=1000, freeze_bn=False, *args, **kwargs):
        assert not include_top
        blocks = blocks or [3, 4, 6, 3]
        x = ZeroPadding2D(padding=3, name="padding_conv1")(inputs)
        x = Conv2D(64, (7, 7), strides=(2, 2), use_bias=False, name="conv1")(x)
        x = BatchNormalization(axis=3, epsilon=1e-5, freeze=freeze_bn, name="bn_conv1")(x)
        x = Activation("relu", name="conv1_relu")(x)
        x = MaxPooling2D((3, 3), strides=(2, 2), padding="same", name="pool1")(x)
        features, outputs = 64, []
        for stage, n in enumerate(blocks):
            for block in range(n):
                x = bottleneck_2d(features, stage, block, freeze_bn)(x)
            features *= 2
            outputs.append(x)
        return Model(inputs=inputs, outputs=outputs, name="resnet50")

    pkg.models.ResNet50 = ResNet50
    return pkg


def install(reference_root="/root/reference", compute_overlap_module=None):
    """tf_shim.install + the functional API above + keras_resnet, and `PyraPose.models` importable."""
    import importlib.machinery
    import os
    reset_uids()
    added = tf_shim.install(reference_root, compute_overlap_module)
    keras = sys.modules["keras"]
    layers = sys.modules["keras.layers"]
    for cls in (Layer, Conv2D, ZeroPadding2D, MaxPooling2D, BatchNormalization, Activation, Add, Reshape, Concatenate):
        setattr(layers, cls.__name__, cls)
    layers.Input = Input
    keras.layers = layers
    keras.models = types.SimpleNamespace(Model=Model)
    keras.regularizers = types.SimpleNamespace(l2=lambda v: ("l2", float(v)))
    keras.initializers = types.SimpleNamespace(
        Initializer=object, normal=lambda mean=0.0, stddev=0.05, seed=None: ("normal", mean, stddev),
        zeros=lambda: "zeros")
    utils = types.ModuleType("keras.utils")
    utils.get_file = lambda fname, origin=None, **kw: os.path.join("/nonexistent", fname)
    utils.Sequence = object
    keras.utils = utils
    kr = _keras_resnet()

    def put(name, mod):
        if name not in sys.modules:
            added.append(name)
        sys.modules[name] = mod

    put("keras.utils", utils)
    put("keras_resnet", kr)
    put("keras_resnet.models", kr.models)
    put("keras_resnet.layers", kr.layers)
    pkg = types.ModuleType("PyraPose.models")
    pkg.__path__ = [os.path.join(reference_root, "PyraPose", "models")]
    pkg.__spec__ = importlib.machinery.ModuleSpec("PyraPose.models", None, is_package=True)
    # the package's own __init__ (Backbone, backbone()) is executed from the reference's file into this module
    src = os.path.join(reference_root, "PyraPose", "models", "__init__.py")
    pkg.__file__ = src
    pkg.__package__ = "PyraPose.models"
    put("PyraPose.models", pkg)
    exec(compile(open(src).read(), src, "exec"), pkg.__dict__)
    return added


uninstall = tf_shim.uninstall
