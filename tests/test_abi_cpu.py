"""CPU-box checks of the drop-in boundary: the C-ABI library loads, exports every symbol the header declares,
rejects bad arguments with PPB_ERR_INVALID before touching CUDA, and FAILS LOUDLY (no CPU fallback) when a
device is needed and absent. No kernels are launched here."""
import ctypes
import os
import re

import numpy as np
import pytest

from pyrapose_b200 import _native as nat

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pyrapose_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"PPB_API\s+[\w\s\*]+?\b(ppb_\w+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(nat.lib_path()):
        from pyrapose_b200 import build
        build.build()
    return nat.load()


def test_header_declares_the_expected_surface():
    syms = declared_symbols()
    for must in ("ppb_compute_overlap_f64", "ppb_gt_assign_f64", "ppb_filter_detections", "ppb_engine_create",
                 "ppb_engine_run", "ppb_box3d_transform_inv_f32", "ppb_conv2d_bf16", "ppb_score_mask",
                 "ppb_vote_gather", "ppb_anchors_shift_f64", "ppb_anchors_shift_f32"):
        assert must in syms
    assert len(syms) >= 25


def test_library_exports_every_declared_symbol(lib):
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, "declared in include/pyrapose_b200.h but not exported: %s" % missing


def test_header_cites_reference_for_each_group():
    text = open(HEADER).read()
    for cite in ("compute_overlap.pyx:13-53", "anchors.py:290-318", "filter_detections.py:21-118",
                 "backend/common.py:25-90", "linemod_eval.py:317-319", "retinanet.py:260-299"):
        assert cite in text, cite


def test_version_and_error_string(lib):
    assert lib.ppb_version() == 100
    assert isinstance(lib.ppb_last_error(), bytes)
    assert nat.launch_count() >= 0


def test_invalid_arguments_are_rejected_before_cuda(lib):
    rc = lib.ppb_conv2d_bf16(None, None, None, None, None, 1, 8, 8, 64, 64, 3, 1, None)
    assert rc == nat.PPB_ERR_INVALID
    with pytest.raises(ValueError):
        nat.check(rc)
    a = np.zeros((4, 4))
    rc = lib.ppb_conv2d_bf16(nat.ptr(a), nat.ptr(a), None, None, nat.ptr(a), 1, 8, 8, 60, 64, 3, 1, None)
    assert rc == nat.PPB_ERR_INVALID and b"multiple of 64" in lib.ppb_last_error()
    rc = lib.ppb_compute_overlap_f64(None, None, None, ctypes.c_longlong(-1), 3, 1, 1, None)
    assert rc == nat.PPB_ERR_INVALID
    rc = lib.ppb_filter_detections(None, None, None, 1, 10, 3, 1, 1, ctypes.c_float(0.05), ctypes.c_float(0.5), 5000,
                                   None, None, None, None, None, None, ctypes.c_size_t(0), None)
    assert rc == nat.PPB_ERR_INVALID and b"max_detections" in lib.ppb_last_error()
    h = ctypes.c_void_p()
    rc = lib.ppb_engine_create(ctypes.byref(h), 0, 1, 480, 640, 300, 9)  # 2700 head channels > 2000
    assert rc == nat.PPB_ERR_INVALID


def test_empty_problems_are_no_ops(lib):
    # n == 0 / k == 0 / batch == 0 return PPB_OK without a device, like the reference returns empty arrays
    assert lib.ppb_compute_overlap_f64(None, None, None, ctypes.c_longlong(0), 5, 1, 1, None) == 0
    assert lib.ppb_clip_boxes_f32(None, None, ctypes.c_longlong(0), ctypes.c_float(1), ctypes.c_float(1), None) == 0
    lib.ppb_filter_workspace_bytes.restype = ctypes.c_size_t
    assert lib.ppb_filter_workspace_bytes(2, 56700, 15, 1, 300) > 2 * 15 * 56700 // 8


def test_no_cpu_fallback_without_a_device(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("this check is for the CPU box")
    h = ctypes.c_void_p()
    rc = lib.ppb_engine_create(ctypes.byref(h), 0, 1, 480, 640, 15, 9)
    assert rc == -2, "engine creation must fail with PPB_ERR_CUDA when there is no GPU"
    with pytest.raises(RuntimeError):
        nat.check(rc)
    from pyrapose_b200 import _kernels
    with pytest.raises(ValueError, match="no CPU path"):
        _kernels.compute_overlap(torch.zeros(3, 4, dtype=torch.float64), torch.zeros(2, 4, dtype=torch.float64))


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "pyrapose_b200")
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh")):
                src = open(os.path.join(base, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), os.path.join(base, f)
                assert "oracle/" not in src or f == "build.py", os.path.join(base, f)


def test_missing_library_raises(monkeypatch):
    monkeypatch.setattr(nat, "_lib", None)
    monkeypatch.setattr(nat, "_LIB_PATH", "/nonexistent/libpyrapose_b200.so")
    with pytest.raises(nat.NativeLibraryMissing):
        nat.load()
