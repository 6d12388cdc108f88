"""Writes tests/golden/keras_partial.h5: a small Keras-2.3.1-layout weight file assembled by hand with
pyrapose_b200.utils.hdf5 (h5py is not available here): the first layers of keras-resnet's ResNet50 under their own
names, one auto-named lateral convolution of __create_sparceFPN (PyraPose/models/retinanet.py:192) under the name a
process that had already built 30 other Conv2D layers would give it, and a layer this model does not have.
Values are seeded; tests/test_hdf5_cpu.py regenerates them to compare.

  python tests/golden/make_keras_h5.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from pyrapose_b200.utils import hdf5  # noqa: E402


def arrays():
    rng = np.random.default_rng(2026)
    f = lambda *s: rng.normal(0, 0.05, s).astype(np.float32)
    return [
        ("conv1", [("conv1/kernel:0", f(7, 7, 3, 64))]),
        ("bn_conv1", [("bn_conv1/gamma:0", f(64) + 1), ("bn_conv1/beta:0", f(64)), ("bn_conv1/moving_mean:0", f(64)),
                      ("bn_conv1/moving_variance:0", np.abs(f(64)) + 1)]),
        ("res2a_branch2a", [("res2a_branch2a/kernel:0", f(1, 1, 64, 64))]),
        ("conv2d_46", [("conv2d_46/kernel:0", f(1, 1, 512, 256)), ("conv2d_46/bias:0", f(256))]),
        ("fc1000", [("fc1000/kernel:0", f(8, 4)), ("fc1000/bias:0", f(4))]),
        ("activation_49", []),
    ]


def main():
    root = hdf5.Group()
    names = []
    for layer, ws in arrays():
        g = root[layer] = hdf5.Group()
        g.attrs["weight_names"] = np.array([w.encode() for w, _ in ws]) if ws else np.zeros((0,), "S1")
        for w, a in ws:
            owner, leaf = w.split("/")
            g.setdefault(owner, hdf5.Group())[leaf] = hdf5.Dataset(a)
        names.append(layer.encode())
    root.attrs["layer_names"] = np.array(names)
    root.attrs["backend"] = b"tensorflow"
    root.attrs["keras_version"] = b"2.3.1"
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "keras_partial.h5")
    hdf5.write_h5(out, root)
    print(out, os.path.getsize(out))


if __name__ == "__main__":
    main()
