"""ppb_filter_detections_static: FilterDetections with its scratch in a file-scope `static __device__` array of the
library (16 MiB) instead of a caller workspace — for callers that cannot call cudaMalloc (CUDA graph capture, tight
serving loops). No counterpart in the reference.

Checked here, through the C-ABI:
  * results are bit-identical (the tolerance every filter test uses: integer / index work and copied floats) to
    ppb_filter_detections with a caller workspace, on two calls in a row whose inputs differ (the second call finds
    the first call's contents in the array) and on a third, full-size call (one 480x640 image, 56 700 anchors);
  * the device's free memory does not move across the calls (nothing is allocated);
  * shapes beyond the array are rejected with PPB_ERR_INVALID and a message that names both sizes (no GPU needed).
The file name sorts last on purpose."""
import ctypes

import numpy as np
import pytest

from pyrapose_b200 import _native as nat


def _args(boxes, b3, cls, class_specific, nms, thr, nms_thr, D, outs, stream):
    B, N, C = cls.shape
    return [nat.ptr(boxes), nat.ptr(b3), nat.ptr(cls), B, N, C, int(class_specific), int(nms), ctypes.c_float(thr),
            ctypes.c_float(nms_thr), D] + [nat.ptr(o) for o in outs]


def _outputs(torch, dev, B, D, fill):
    return [torch.full((B, D, 4), fill, dtype=torch.float32, device=dev),
            torch.full((B, D, 16), fill, dtype=torch.float32, device=dev),
            torch.full((B, D), fill, dtype=torch.float32, device=dev),
            torch.full((B, D), int(fill), dtype=torch.int32, device=dev),
            torch.full((B, D), int(fill), dtype=torch.int32, device=dev)]


def test_static_scratch_rejects_sizes_beyond_the_array():
    lib = nat.load()
    lib.ppb_filter_static_capacity_bytes.restype = ctypes.c_size_t
    lib.ppb_filter_workspace_bytes.restype = ctypes.c_size_t
    cap = lib.ppb_filter_static_capacity_bytes()
    assert cap == 16 << 20
    assert lib.ppb_filter_workspace_bytes(1, 56700, 31, 1, 300) <= cap < lib.ppb_filter_workspace_bytes(1, 56700, 32, 1, 300)
    need = lib.ppb_filter_workspace_bytes(4, 56700, 15, 1, 300)
    assert need > cap
    # rejected on the size alone, before any pointer is looked at or CUDA is touched
    rc = lib.ppb_filter_detections_static(None, None, None, 4, 56700, 15, 1, 1, ctypes.c_float(0.05),
                                          ctypes.c_float(0.5), 300, None, None, None, None, None, None)
    assert rc == nat.PPB_ERR_INVALID
    msg = lib.ppb_last_error().decode()
    assert str(need) in msg and str(cap) in msg, msg
    with pytest.raises(ValueError):
        nat.check(rc)
    rc = lib.ppb_filter_detections_static(None, None, None, 1, 100, 3, 1, 1, ctypes.c_float(0.05), ctypes.c_float(0.5),
                                          5000, None, None, None, None, None, None)
    assert rc == nat.PPB_ERR_INVALID   # max_detections beyond the kernel's limit, as in ppb_filter_detections


@pytest.mark.gpu
def test_static_scratch_matches_the_workspace_entry_point(cuda_device):
    import torch
    from test_kernels_gpu import _synthetic_dets
    lib = nat.load()
    lib.ppb_filter_workspace_bytes.restype = ctypes.c_size_t
    dev = cuda_device
    cases = [  # (seed, B, N, C, logit mean, class_specific, nms, score thr, nms thr, D)
        (11, 2, 9000, 15, -1.0, 1, 1, 0.05, 0.5, 300),    # dense: thousands of candidates per class
        (12, 1, 5000, 7, -6.0, 1, 1, 0.05, 0.5, 300),     # sparse and smaller: finds the first call's keys in the array
        (13, 1, 56700, 15, -5.0, 1, 1, 0.05, 0.5, 300),   # one 480x640 image
        (14, 3, 7000, 5, -3.0, 0, 1, 0.3, 0.4, 100),      # class-agnostic (class maxima live in the scratch too)
    ]
    staged = []
    for seed, B, N, C, mean, cs, nms, thr, nthr, D in cases:
        boxes, b3, cls = (torch.from_numpy(a).to(dev) for a in _synthetic_dets(seed, B, N, C, mean))
        need = lib.ppb_filter_workspace_bytes(B, N, C, cs, D)
        ws = torch.empty(int(need), dtype=torch.uint8, device=dev)
        ref = _outputs(torch, dev, B, D, -7.0)
        got = _outputs(torch, dev, B, D, -9.0)
        staged.append((boxes, b3, cls, cs, nms, thr, nthr, D, ws, need, ref, got))
    stream = torch.cuda.current_stream(dev)
    sp = ctypes.c_void_p(stream.cuda_stream)
    with torch.cuda.device(dev):
        # reference results from the existing entry point (caller workspace)
        for boxes, b3, cls, cs, nms, thr, nthr, D, ws, need, ref, got in staged:
            nat.check(lib.ppb_filter_detections(*(_args(boxes, b3, cls, cs, nms, thr, nthr, D, ref, sp)
                                                  + [nat.ptr(ws), ctypes.c_size_t(need), sp])))
        torch.cuda.synchronize(dev)
        # one untimed call first: module loading / the symbol look-up happen here, not an allocation of ours
        b0 = staged[1]
        nat.check(lib.ppb_filter_detections_static(*(_args(*b0[:8], _outputs(torch, dev, b0[2].shape[0], b0[7], 0.0), sp)
                                                     + [sp])))
        torch.cuda.synchronize(dev)
        free_before = torch.cuda.mem_get_info(dev)[0]
        torch_before = torch.cuda.memory_allocated(dev)
        # the calls under test, back to back on one stream, nothing in between
        for boxes, b3, cls, cs, nms, thr, nthr, D, ws, need, ref, got in staged:
            nat.check(lib.ppb_filter_detections_static(*(_args(boxes, b3, cls, cs, nms, thr, nthr, D, got, sp) + [sp])))
        torch.cuda.synchronize(dev)
        assert torch.cuda.mem_get_info(dev)[0] >= free_before, "device memory was allocated during the calls"
        assert torch.cuda.memory_allocated(dev) == torch_before
    names = ["boxes", "boxes3D", "scores", "labels", "indices"]
    for i, st in enumerate(staged):
        ref, got = st[10], st[11]
        kept = int((ref[3] >= 0).sum())
        print("call %d: %d detections kept, outputs identical" % (i, kept))
        assert kept > 0
        for nm, r, g in zip(names, ref, got):
            assert torch.equal(r, g), "call %d: %s differ between the two entry points" % (i, nm)
