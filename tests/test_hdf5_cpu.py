"""Keras HDF5 weight files without h5py: the pure-Python reader / writer (utils/hdf5.py), the
Keras-name -> engine-name map by graph order (models/keras_io.py) and the model-level calls the reference makes:
model.load_weights(path, by_name=True, skip_mismatch=...) (PyraPose/bin/train.py:65-68), model.save / save_weights
(ModelCheckpoint, bin/train.py:131-142) and models.load_model (PyraPose/models/__init__.py:68-71)."""
import glob
import os

import numpy as np
import pytest

from pyrapose_b200 import models
from pyrapose_b200.models import keras_io, spec
from pyrapose_b200.utils import hdf5

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_reader_on_a_genuine_libhdf5_file():
    """scipy ships one MATLAB v7.3 file = an HDF5 file written by libhdf5 behind a 512-byte user block: superblock 0,
    symbol-table root group, version-1 object headers, an attribute and a float64 dataset."""
    import scipy
    hits = glob.glob(os.path.join(os.path.dirname(scipy.__file__), "io", "matlab", "tests", "data", "testhdf5_7.4_GLNX86.mat"))
    if not hits:
        pytest.skip("scipy's HDF5 test file is not installed")
    root = hdf5.read_h5(hits[0])
    assert list(root.keys()) == ["testdouble"]
    ds = root["testdouble"]
    assert ds.data.shape == (9, 1) and ds.data.dtype == np.float64
    assert np.allclose(ds.data[:, 0], np.arange(9) * np.pi / 4, rtol=0, atol=1e-15)
    assert bytes(ds.attrs["MATLAB_class"]) == b"double"


def test_write_read_round_trip_of_every_supported_type(tmp_path):
    rng = np.random.default_rng(0)
    root = hdf5.Group()
    root.attrs["title"] = b"round trip"
    root.attrs["names"] = np.array([b"a", b"bcd", b"ef"])
    root.attrs["version"] = np.int64(7)
    root.attrs["floats"] = np.array([1.5, -2.25], np.float32)
    g = root["outer"] = hdf5.Group()
    g["f32"] = hdf5.Dataset(rng.normal(size=(3, 4, 5)).astype(np.float32), {"note": b"x"})
    g["f64"] = hdf5.Dataset(rng.normal(size=(7,)))
    g["i32"] = hdf5.Dataset(rng.integers(-5, 5, (2, 2)).astype(np.int32))
    g["u8"] = hdf5.Dataset(rng.integers(0, 255, (11,)).astype(np.uint8))
    g["empty"] = hdf5.Dataset(np.zeros((0, 3), np.float32))
    inner = g["inner"] = hdf5.Group()
    inner["scalar"] = hdf5.Dataset(np.float32(3.25))
    many = root["many"] = hdf5.Group()                      # more entries than libhdf5's default leaf node holds
    for i in range(300):
        many["layer_%03d" % i] = hdf5.Dataset(np.full((2,), i, np.float32))
    path = str(tmp_path / "t.h5")
    hdf5.write_h5(path, root)
    back = hdf5.read_h5(path)
    assert bytes(back.attrs["title"]) == b"round trip" and int(back.attrs["version"]) == 7
    assert [bytes(v) for v in back.attrs["names"]] == [b"a", b"bcd", b"ef"]
    assert np.array_equal(back.attrs["floats"], root.attrs["floats"])
    for k in ("f32", "f64", "i32", "u8", "empty"):
        a, b = back["outer"][k].data, g[k].data
        assert a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a, b), k
    assert bytes(back["outer"]["f32"].attrs["note"]) == b"x"
    assert float(back["outer"]["inner"]["scalar"].data) == 3.25
    assert len(back["many"]) == 300 and float(back["many"]["layer_299"].data[0]) == 299.0
    assert sorted(back["many"].keys()) == sorted(many.keys())


def test_rejects_what_is_not_hdf5(tmp_path):
    p = tmp_path / "x.h5"
    p.write_bytes(b"not an hdf5 file at all" * 10)
    with pytest.raises(ValueError):
        hdf5.read_h5(str(p))
    assert not keras_io.is_hdf5(str(p))


def test_committed_fixture_by_name_and_graph_order():
    """tests/golden/keras_partial.h5 (written by tests/golden/make_keras_h5.py): named backbone layers load by name,
    the auto-named 1x1 512->256 convolution is recognised as P3_lateral although its number is not the fresh-process
    one, unknown layers are reported and ignored (by_name semantics)."""
    import importlib.util
    sp = importlib.util.spec_from_file_location("make_keras_h5", os.path.join(GOLD, "make_keras_h5.py"))
    mk = importlib.util.module_from_spec(sp)
    sp.loader.exec_module(mk)
    want = {l: dict((w.split("/")[1].split(":")[0], a) for w, a in ws) for l, ws in mk.arrays()}
    model = models.backbone("resnet50").retinanet(num_classes=4, seed=2)
    before = {k: v.copy() for k, v in model.weights.items()}
    got, unused = keras_io.load_h5(os.path.join(GOLD, "keras_partial.h5"), model.weights.keys())
    assert sorted(unused) == ["fc1000"]
    assert set(got) == {"conv1/kernel", "bn_conv1/gamma", "bn_conv1/beta", "bn_conv1/moving_mean",
                        "bn_conv1/moving_variance", "res2a_branch2a/kernel", "P3_lateral/kernel", "P3_lateral/bias"}
    model.load_weights(os.path.join(GOLD, "keras_partial.h5"), by_name=True)
    assert np.array_equal(model.weights["conv1/kernel"], want["conv1"]["kernel"])
    assert np.array_equal(model.weights["bn_conv1/moving_variance"], want["bn_conv1"]["moving_variance"])
    assert np.array_equal(model.weights["P3_lateral/kernel"], want["conv2d_46"]["kernel"])
    assert np.array_equal(model.weights["P3_lateral/bias"], want["conv2d_46"]["bias"])
    untouched = [k for k in before if k not in got]
    assert len(untouched) > 290 and all(np.array_equal(model.weights[k], before[k]) for k in untouched)


@pytest.fixture(scope="module")
def full_file(tmp_path_factory):
    C = 6
    w = spec.init_weights(C, seed=3, randomize_bn=True)
    path = str(tmp_path_factory.mktemp("h5") / "weights.h5")
    m = models.backbone("resnet50").retinanet(num_classes=C)
    m.set_weights_dict({k: v.copy() for k, v in w.items()})
    m.save_weights(path)
    return C, w, path


def test_save_weights_writes_the_reference_layer_names(full_file):
    C, w, path = full_file
    root = hdf5.read_h5(path)
    names = [bytes(n).decode() for n in root.attrs["layer_names"]]
    # what a reference model built in a fresh process holds (retinanet.py:260-299 creation order)
    for n in ("conv1", "bn_conv1", "res5c_branch2c", "bn5c_branch2c", "P3", "P4", "P5", "model_1", "model_2", "mask") + \
            tuple("conv2d_%d" % k for k in range(16, 23)):
        assert n in names, n
    assert not any(n.startswith("pyramid_") or n.endswith("_lateral") for n in names)
    wn = [bytes(n).decode() for n in root["model_1"].attrs["weight_names"]]
    assert wn == [x for k in range(1, 6) for x in ("conv2d_%d/kernel:0" % k, "conv2d_%d/bias:0" % k)]
    assert root["model_1"]["conv2d_5"]["kernel:0"].data.shape == (3, 3, 512, 9 * 16)
    assert root["model_2"]["conv2d_10"]["kernel:0"].data.shape == (3, 3, 256, 9 * C)
    assert root["mask"]["conv2d_15"]["bias:0"].data.shape == (C,)
    assert root["conv2d_18"]["conv2d_18"]["kernel:0"].data.shape == (1, 1, 2048, 256)      # P5_lateral
    assert np.array_equal(root["conv2d_19"]["conv2d_19"]["kernel:0"].data, w["P4_mid/kernel"])
    assert np.array_equal(root["conv2d_22"]["conv2d_22"]["kernel:0"].data, w["P4_down/kernel"])
    assert bytes(root.attrs["keras_version"]) == b"2.3.1"


def test_load_weights_round_trip_and_shifted_auto_numbers(full_file, tmp_path):
    C, w, path = full_file
    m = models.backbone("resnet50").retinanet(num_classes=C, seed=9)
    assert not np.array_equal(m.weights["P3_mid/kernel"], w["P3_mid/kernel"])
    m.load_weights(path, by_name=True)
    assert all(np.array_equal(m.weights[k], w[k]) for k in w)
    # the same file as a process that had built other models first would have written it: every auto number shifted
    root = hdf5.read_h5(path)
    shifted = hdf5.Group()
    ren = {}
    for name in root:
        new = name
        if name.startswith("conv2d_"):
            new = "conv2d_%d" % (int(name[7:]) + 37)
        elif name.startswith("model_"):
            new = "model_%d" % (int(name[6:]) + 4)
        ren[name] = new
    for name, g in root.items():
        ng = shifted[ren[name]] = hdf5.Group()
        wn = []
        for n in g.attrs["weight_names"]:
            owner, leaf = bytes(n).decode().split("/")
            o2 = "conv2d_%d" % (int(owner[7:]) + 37) if owner.startswith("conv2d_") else owner
            ng.setdefault(o2, hdf5.Group())[leaf] = g[owner][leaf]
            wn.append(("%s/%s" % (o2, leaf)).encode())
        ng.attrs["weight_names"] = np.array(wn)
    shifted.attrs["layer_names"] = np.array([ren[bytes(n).decode()].encode() for n in root.attrs["layer_names"]])
    p2 = str(tmp_path / "shifted.h5")
    hdf5.write_h5(p2, shifted)
    m2 = models.backbone("resnet50").retinanet(num_classes=C, seed=10)
    m2.load_weights(p2, by_name=True)
    assert all(np.array_equal(m2.weights[k], w[k]) for k in w)


def test_skip_mismatch_and_full_model_snapshot(full_file, tmp_path):
    C, w, path = full_file
    other = models.backbone("resnet50").retinanet(num_classes=C + 2, seed=4)
    with pytest.raises(ValueError, match="shape mismatch"):
        other.load_weights(path, by_name=True)
    keep = other.weights["pyramid_classification/kernel"].copy()
    other.load_weights(path, by_name=True, skip_mismatch=True)            # train.py:67 passes skip_mismatch=True
    assert np.array_equal(other.weights["pyramid_classification/kernel"], keep)
    assert np.array_equal(other.weights["res4c_branch2b/kernel"], w["res4c_branch2b/kernel"])
    # ModelCheckpoint-style full-model file -> models.load_model
    m = models.backbone("resnet50").retinanet(num_classes=C)
    m.set_weights_dict({k: v.copy() for k, v in w.items()})
    snap = str(tmp_path / "resnet50_linemod_01.h5")
    models.convert_model(m).save(snap)
    assert "model_weights" in hdf5.read_h5(snap)
    back = models.load_model(snap, backbone_name="resnet50")
    assert back.inference and back.num_classes == C
    assert all(np.array_equal(back.weights[k], w[k]) for k in w)
    # a snapshot without this package's meta attribute (i.e. one the reference wrote): classes from the head shapes
    root = hdf5.read_h5(snap)
    root.attrs.pop("pyrapose_b200_meta")
    snap2 = str(tmp_path / "reference_snapshot.h5")
    hdf5.write_h5(snap2, root)
    back2 = models.load_model(snap2)
    assert back2.num_classes == C and all(np.array_equal(back2.weights[k], w[k]) for k in w)
