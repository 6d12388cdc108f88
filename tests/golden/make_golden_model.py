"""Golden vectors of the PFPN forward pass made by running the REFERENCE'S OWN model-building code
(PyraPose/models/__init__.py backbone / convert_model, models/resnet.py resnet_retinanet, models/retinanet.py,
initializers.py, layers/_misc.py) over the miniature Keras of tests/golden/keras_shim.py — see its header for what is
the reference's text (the wiring, the layer arguments, the output order, the auto-naming) and what is restated library
behaviour (Conv2D / 'same' padding / MaxPooling2D / BatchNormalization arithmetic, keras-resnet's ResNet50).

    python tests/golden/make_golden_model.py      -> tests/golden/reference_model.npz, reference_model_layers.txt

The weights are this repo's seeded generator (pyrapose_b200.models.spec.init_weights), loaded into the reference-built
model THROUGH the Keras layer-name map of pyrapose_b200/models/keras_io.py — so the vectors also verify that map: a
wrong conv2d_<k> assignment would put a kernel into a layer of another shape or another place and change the outputs."""
import importlib
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, HERE)

C, B, H, W = 4, 2, 96, 136          # 136 / 8 = 17: odd feature widths exercise the asymmetric 'same' padding
SEED = 5


def make_inputs():
    from pyrapose_b200.models import spec
    weights = spec.init_weights(C, seed=SEED, head_init="spread", randomize_bn=True)
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, (B, H, W, 3)).astype(np.float32) - np.array([103.939, 116.779, 123.68], np.float32)
    return weights, img


def main():
    import keras_shim
    from oracle import build as obuild
    from pyrapose_b200.models import keras_io
    co = obuild.load_ref_compute_overlap()
    com = types.ModuleType("PyraPose.utils.compute_overlap")
    com.compute_overlap = co
    added = keras_shim.install(compute_overlap_module=com)
    try:
        models = importlib.import_module("PyraPose.models")
        training = models.backbone("resnet50").retinanet(num_classes=C)         # bin/train.py:71-104 create_models
        prediction = models.convert_model(training)                             # models/__init__.py convert_model
        weights, img = make_inputs()
        # ---- weights in, through the name map under test
        name_map = keras_io.keras_layer_names()
        by_path = dict(training.all_layers())
        used = set()
        for key, arr in weights.items():
            layer, param = key.split("/")
            group, kname = name_map.get(layer, (layer, layer))
            path = kname if group == kname else group + "/" + kname
            tgt = by_path[path].weights
            assert param in tgt, (key, path, list(tgt))
            assert tgt[param].shape == arr.shape, (key, path, tgt[param].shape, arr.shape)
            tgt[param] = np.ascontiguousarray(arr, np.float32)
            used.add((path, param))
        every = set((p, k) for p, l in by_path.items() for k in l.weights)
        assert used == every, sorted(every - used)[:5]
        # ---- forward
        out = {}
        reg, cls, msk = training.predict_on_batch(img)
        boxes3D, cls2, msk2 = prediction.predict_on_batch(img)
        assert np.array_equal(cls, cls2) and np.array_equal(msk, msk2)
        out["train_3Dbox"], out["train_cls"], out["train_mask"] = reg, cls, msk
        out["pred_boxes3D"] = boxes3D
        Model = sys.modules["keras"].models.Model
        resnet_outs = [training.get_layer(n).output for n in ("res3d_relu", "res4f_relu", "res5c_relu")]
        feats = [training.get_layer(n).output for n in ("P3", "P4", "P5")]
        taps = Model(inputs=training.inputs, outputs=resnet_outs + feats).predict_on_batch(img)
        for name, v in zip(("C3", "C4", "C5", "P3", "P4", "P5"), taps):
            out["tap_" + name] = v
        out["image"] = img
        out["config"] = np.array([C, B, H, W, SEED])
        np.savez_compressed(os.path.join(HERE, "reference_model.npz"), **out)
        # ---- the layer table of the reference-built model: path, class, weight shapes, trainable after resnet.py:100-103
        with open(os.path.join(HERE, "reference_model_layers.txt"), "w") as f:
            f.write("# output_names training %s prediction %s\n" % (training.output_names, prediction.output_names))
            for path, layer in training.all_layers():
                shapes = " ".join("%s=%s" % (k, "x".join(str(d) for d in v.shape)) for k, v in layer.weights.items())
                f.write("%s %s trainable=%d %s\n" % (path, type(layer).__name__, int(layer.trainable), shapes))
        print("wrote reference_model.npz (%d bytes) and reference_model_layers.txt (%d layers)" %
              (os.path.getsize(os.path.join(HERE, "reference_model.npz")), len(by_path)))
        for k, v in out.items():
            print("%-14s %s" % (k, v.shape))
    finally:
        keras_shim.uninstall(added)


if __name__ == "__main__":
    main()
