"""Thin torch-tensor front-ends of the C-ABI kernels (device memory and streams come from torch;
the arithmetic is all in libpyrapose_b200.so). Every function takes CUDA tensors and launches on the
current torch stream. No CPU fallback: a CPU tensor raises ValueError."""
import ctypes

import torch

from . import _native as nat

_c_float_p = ctypes.POINTER(ctypes.c_float)


def _dev(t, dtype, name):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise ValueError("%s must be a CUDA tensor (pyrapose_b200 has no CPU path)" % name)
    if t.dtype != dtype:
        raise ValueError("%s must have dtype %s, got %s" % (name, dtype, t.dtype))
    return t.contiguous()


def _stream():
    return nat.current_stream_ptr()


def _host16(v, n):
    if v is None:
        return None
    arr = (ctypes.c_float * n)(*[float(x) for x in v])
    return arr


# --------------------------------------------------------------------------- IoU / assignment
def compute_overlap(boxes, query):
    """boxes f64 [N,4] or [B,N,4]; query f64 [K,4] or [B,K,4] -> f64 [N,K] / [B,N,K]."""
    boxes = _dev(boxes, torch.float64, "boxes")
    query = _dev(query, torch.float64, "query_boxes")
    batched = query.dim() == 3
    q = query if batched else query.unsqueeze(0)
    B, K = q.shape[0], q.shape[1]
    shared = boxes.dim() == 2
    N = boxes.shape[-2]
    out = torch.empty((B, N, K), dtype=torch.float64, device=boxes.device)
    with torch.cuda.device(boxes.device):
        nat.check(nat.load().ppb_compute_overlap_f64(nat.ptr(boxes), nat.ptr(q), nat.ptr(out), ctypes.c_longlong(N),
                                                     K, B, int(shared), _stream()))
    return out if batched else out[0]


def gt_assign(boxes, query, query_count=None, negative_overlap=0.4, positive_overlap=0.5, want_max=False):
    """Fused compute_gt_annotations. boxes f64 [N,4] (shared) or [B,N,4]; query f64 [B,K,4] or [K,4].
    Returns (positive bool, ignore bool, argmax int64[, max f64]) shaped [B,N] (or [N])."""
    boxes = _dev(boxes, torch.float64, "anchors")
    query = _dev(query, torch.float64, "annotations")
    batched = query.dim() == 3
    q = query if batched else query.unsqueeze(0)
    B, K = q.shape[0], q.shape[1]
    shared = boxes.dim() == 2
    N = boxes.shape[-2]
    dev = boxes.device
    argmax = torch.empty((B, N), dtype=torch.int64, device=dev)
    pos = torch.empty((B, N), dtype=torch.uint8, device=dev)
    ign = torch.empty((B, N), dtype=torch.uint8, device=dev)
    mx = torch.empty((B, N), dtype=torch.float64, device=dev) if want_max else None
    qc = _dev(query_count, torch.int32, "query_count") if query_count is not None else None
    with torch.cuda.device(dev):
        nat.check(nat.load().ppb_gt_assign_f64(nat.ptr(boxes), nat.ptr(q), nat.ptr(qc), N, K, B, int(shared),
                                               ctypes.c_double(negative_overlap), ctypes.c_double(positive_overlap),
                                               nat.ptr(argmax), nat.ptr(pos), nat.ptr(ign), nat.ptr(mx), _stream()))
    res = (pos.bool(), ign.bool(), argmax) + ((mx,) if want_max else ())
    return res if batched else tuple(r[0] for r in res)



def anchor_targets(anchors, query, counts, labels, boxes16, image_wh, num_classes, negative_overlap=0.4,
                   positive_overlap=0.5):
    """Per-anchor targets of anchor_targets_bbox on the device: fused IoU assignment (ppb_gt_assign_f64) followed by
    ppb_anchor_targets. anchors f64 [N,4]; query f64 [B,K,4]; counts i32 [B]; labels i32 [B,K]; boxes16 f64 [B,K,16];
    image_wh i32 [B,2]. Returns CUDA tensors regression_3D f32 [B,N,17], labels f32 [B,N,C+1]."""
    anchors = _dev(anchors, torch.float64, "anchors")
    query = _dev(query, torch.float64, "bboxes")
    counts = _dev(counts, torch.int32, "counts")
    labels = _dev(labels, torch.int32, "labels")
    boxes16 = _dev(boxes16, torch.float64, "boxes16")
    image_wh = _dev(image_wh, torch.int32, "image_wh")
    B, K = int(query.shape[0]), int(query.shape[1])
    N = int(anchors.shape[0])
    dev = anchors.device
    argmax = torch.empty((B, N), dtype=torch.int64, device=dev)
    pos = torch.empty((B, N), dtype=torch.uint8, device=dev)
    ign = torch.empty((B, N), dtype=torch.uint8, device=dev)
    reg = torch.empty((B, N, 17), dtype=torch.float32, device=dev)
    cls = torch.empty((B, N, num_classes + 1), dtype=torch.float32, device=dev)
    lib = nat.load()
    with torch.cuda.device(dev):
        nat.check(lib.ppb_gt_assign_f64(nat.ptr(anchors), nat.ptr(query), nat.ptr(counts), N, K, B, 1,
                                        ctypes.c_double(negative_overlap), ctypes.c_double(positive_overlap),
                                        nat.ptr(argmax), nat.ptr(pos), nat.ptr(ign), None, _stream()))
        nat.check(lib.ppb_anchor_targets(nat.ptr(anchors), nat.ptr(argmax), nat.ptr(pos), nat.ptr(ign), nat.ptr(labels),
                                         nat.ptr(boxes16), nat.ptr(counts), nat.ptr(image_wh), N, K, B,
                                         int(num_classes), nat.ptr(reg), nat.ptr(cls), _stream()))
    return reg, cls


def mask_targets(masks, mask_ids, labels, counts, mask_h, mask_w, num_classes):
    """Mask targets of anchor_targets_bbox on the device. masks u8 [B,H,W] instance-id images; mask_ids / labels
    i32 [B,K]; counts i32 [B]. Returns a CUDA tensor f32 [B, mask_h*mask_w, C+1]."""
    masks = _dev(masks, torch.uint8, "mask")
    mask_ids = _dev(mask_ids, torch.int32, "mask_ids")
    labels = _dev(labels, torch.int32, "labels")
    counts = _dev(counts, torch.int32, "counts")
    B, H, W = int(masks.shape[0]), int(masks.shape[1]), int(masks.shape[2])
    K = int(mask_ids.shape[1])
    dev = masks.device
    cells = torch.empty((B, max(K, 1)), dtype=torch.int32, device=dev)
    out = torch.empty((B, mask_h * mask_w, num_classes + 1), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        nat.check(nat.load().ppb_mask_targets(nat.ptr(masks), H, W, nat.ptr(mask_ids), nat.ptr(labels), nat.ptr(counts),
                                              K, B, int(mask_h), int(mask_w), int(num_classes), nat.ptr(cells),
                                              nat.ptr(out), _stream()))
    return out

# --------------------------------------------------------------------------- anchors
def anchors_shift(base, fh, fw, stride, out=None):
    """base [A,4] f32/f64 CUDA -> [fh*fw*A, 4] shifted anchors of one level."""
    if base.dtype not in (torch.float32, torch.float64):
        raise ValueError("base anchors must be float32 or float64")
    base = _dev(base, base.dtype, "base anchors")
    A = base.shape[0]
    if out is None:
        out = torch.empty((fh * fw * A, 4), dtype=base.dtype, device=base.device)
    lib = nat.load()
    with torch.cuda.device(base.device):
        if base.dtype == torch.float64:
            nat.check(lib.ppb_anchors_shift_f64(nat.ptr(base), A, int(fh), int(fw), ctypes.c_double(stride),
                                                nat.ptr(out), _stream()))
        else:
            nat.check(lib.ppb_anchors_shift_f32(nat.ptr(base), A, int(fh), int(fw), ctypes.c_float(stride),
                                                nat.ptr(out), _stream()))
    return out


# --------------------------------------------------------------------------- decode
def box3d_transform_inv(anchors, deltas, mean=None, std=None, denormalize=False, want_hull=False):
    anchors = _dev(anchors, torch.float32, "anchors")
    deltas = _dev(deltas, torch.float32, "regression")
    if deltas.dim() != 3 or deltas.shape[2] != 16:
        raise ValueError("regression must have shape (B, N, 16), got %s" % (tuple(deltas.shape),))
    B, N = deltas.shape[0], deltas.shape[1]
    batched = anchors.dim() == 3
    if anchors.shape[-2] != N or anchors.shape[-1] != 4 or (batched and anchors.shape[0] != B):
        raise ValueError("anchors %s do not match regression %s" % (tuple(anchors.shape), tuple(deltas.shape)))
    out = torch.empty_like(deltas)
    hull = torch.empty((B, N, 4), dtype=torch.float32, device=deltas.device) if want_hull else None
    with torch.cuda.device(deltas.device):
        nat.check(nat.load().ppb_box3d_transform_inv_f32(nat.ptr(anchors), nat.ptr(deltas), nat.ptr(out), nat.ptr(hull),
                                                         B, ctypes.c_longlong(N), int(batched), _host16(mean, 16),
                                                         _host16(std, 16), int(denormalize), _stream()))
    return (out, hull) if want_hull else out


def bbox_transform_inv(anchors, deltas, mean=None, std=None):
    anchors = _dev(anchors, torch.float32, "anchors")
    deltas = _dev(deltas, torch.float32, "regression")
    if deltas.dim() != 3 or deltas.shape[2] != 4:
        raise ValueError("regression must have shape (B, N, 4), got %s" % (tuple(deltas.shape),))
    B, N = deltas.shape[0], deltas.shape[1]
    batched = anchors.dim() == 3
    out = torch.empty_like(deltas)
    with torch.cuda.device(deltas.device):
        nat.check(nat.load().ppb_bbox_transform_inv_f32(nat.ptr(anchors), nat.ptr(deltas), nat.ptr(out), B,
                                                        ctypes.c_longlong(N), int(batched), _host16(mean, 4),
                                                        _host16(std, 4), _stream()))
    return out


def clip_boxes(boxes, width, height):
    boxes = _dev(boxes, torch.float32, "boxes")
    out = torch.empty_like(boxes)
    with torch.cuda.device(boxes.device):
        nat.check(nat.load().ppb_clip_boxes_f32(nat.ptr(boxes), nat.ptr(out), ctypes.c_longlong(boxes.numel() // 4),
                                                ctypes.c_float(width), ctypes.c_float(height), _stream()))
    return out


def upsample_like(source, dst_h, dst_w):
    """UpsampleLike: nearest resize of an NHWC float32 tensor to (dst_h, dst_w)."""
    source = _dev(source, torch.float32, "source")
    if source.dim() != 4:
        raise ValueError("expected source of shape (B, H, W, C), got %s" % (tuple(source.shape),))
    B, ih, iw, C = (int(v) for v in source.shape)
    out = torch.empty((B, int(dst_h), int(dst_w), C), dtype=torch.float32, device=source.device)
    with torch.cuda.device(source.device):
        nat.check(nat.load().ppb_upsample_like_f32(nat.ptr(source), nat.ptr(out), B, ih, iw, int(dst_h), int(dst_w), C,
                                                   _stream()))
    return out


# --------------------------------------------------------------------------- filter
_workspaces = {}


def _workspace(dev, nbytes):
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(int(nbytes), dtype=torch.uint8, device=dev)
        _workspaces[key] = ws
    return ws


def filter_detections(boxes, boxes3D, classification, class_specific_filter=True, nms=True, score_threshold=0.05,
                      max_detections=300, nms_threshold=0.5, want_indices=False):
    boxes = _dev(boxes, torch.float32, "boxes")
    boxes3D = _dev(boxes3D, torch.float32, "boxes3D")
    cls = _dev(classification, torch.float32, "classification")
    if boxes.dim() != 3 or boxes.shape[2] != 4:
        raise ValueError("boxes must have shape (B, N, 4), got %s" % (tuple(boxes.shape),))
    B, N = boxes.shape[0], boxes.shape[1]
    if tuple(boxes3D.shape) != (B, N, 16) or cls.dim() != 3 or tuple(cls.shape[:2]) != (B, N):
        raise ValueError("boxes3D %s / classification %s do not match boxes %s" %
                         (tuple(boxes3D.shape), tuple(cls.shape), tuple(boxes.shape)))
    C = cls.shape[2]
    D = int(max_detections)
    dev = boxes.device
    lib = nat.load()
    lib.ppb_filter_workspace_bytes.restype = ctypes.c_size_t
    need = lib.ppb_filter_workspace_bytes(B, N, C, int(class_specific_filter), D)
    ws = _workspace(dev, need)
    ob = torch.empty((B, D, 4), dtype=torch.float32, device=dev)
    o3 = torch.empty((B, D, 16), dtype=torch.float32, device=dev)
    os_ = torch.empty((B, D), dtype=torch.float32, device=dev)
    ol = torch.empty((B, D), dtype=torch.int32, device=dev)
    oi = torch.empty((B, D), dtype=torch.int32, device=dev) if want_indices else None
    with torch.cuda.device(dev):
        nat.check(lib.ppb_filter_detections(nat.ptr(boxes), nat.ptr(boxes3D), nat.ptr(cls), B, N, C,
                                            int(class_specific_filter), int(nms), ctypes.c_float(score_threshold),
                                            ctypes.c_float(nms_threshold), D, nat.ptr(ob), nat.ptr(o3), nat.ptr(os_),
                                            nat.ptr(ol), nat.ptr(oi), nat.ptr(ws), ctypes.c_size_t(ws.numel()),
                                            _stream()))
    return (ob, o3, os_, ol) + ((oi,) if want_indices else ())


def gather_rows(src, indices):
    """src f32 [B,N,...], indices i32 [B,D] (-1 = padding) -> f32 [B,D,...] (-1 filled where padded)."""
    src = _dev(src, torch.float32, "other")
    indices = _dev(indices, torch.int32, "indices")
    B, N = src.shape[0], src.shape[1]
    D = indices.shape[1]
    row = int(src.numel() // max(B * N, 1))
    out = torch.empty((B, D) + tuple(src.shape[2:]), dtype=torch.float32, device=src.device)
    with torch.cuda.device(src.device):
        nat.check(nat.load().ppb_gather_rows_f32(nat.ptr(src), nat.ptr(indices), nat.ptr(out), B, N, D, row, _stream()))
    return out


def score_mask(scores, threshold, want_counts=True):
    """scores f32 [B,N,C] -> (mask u32-as-i32 [B,C,ceil(N/32)], counts i32 [B,C] or None)."""
    scores = _dev(scores, torch.float32, "scores")
    B, N, C = scores.shape
    W = (N + 31) // 32
    mask = torch.empty((B, C, W), dtype=torch.int32, device=scores.device)
    counts = torch.empty((B, C), dtype=torch.int32, device=scores.device) if want_counts else None
    with torch.cuda.device(scores.device):
        nat.check(nat.load().ppb_score_mask(nat.ptr(scores), B, N, C, ctypes.c_float(threshold), nat.ptr(mask),
                                            nat.ptr(counts), _stream()))
    return mask, counts


def vote_gather(mask, scores, boxes3D, offsets, total):
    scores = _dev(scores, torch.float32, "scores")
    boxes3D = _dev(boxes3D, torch.float32, "boxes3D")
    B, N, C = scores.shape
    dev = scores.device
    indices = torch.empty((total,), dtype=torch.int64, device=dev)
    votes = torch.empty((total, 16), dtype=torch.float32, device=dev)
    conf = torch.empty((B, C), dtype=torch.float32, device=dev)
    offsets = _dev(offsets, torch.int64, "offsets")
    with torch.cuda.device(dev):
        nat.check(nat.load().ppb_vote_gather(nat.ptr(mask), nat.ptr(scores), nat.ptr(boxes3D), B, N, C,
                                             nat.ptr(offsets), nat.ptr(indices), nat.ptr(votes), nat.ptr(conf),
                                             _stream()))
    return indices, votes, conf


# --------------------------------------------------------------------------- convolution
def conv2d_bf16(x_padded, weight, bias=None, residual=None, ksize=3, relu=False):
    """x_padded bf16 [B,H+2,W+2,Cin] (zero border), weight bf16 [Cout, k*k*Cin] -> bf16 [B,H+2,W+2,Cout]."""
    x = _dev(x_padded, torch.bfloat16, "x")
    w = _dev(weight, torch.bfloat16, "weight")
    B, Hp, Wp, Cin = x.shape
    Cout = w.shape[0]
    y = torch.empty((B, Hp, Wp, Cout), dtype=torch.bfloat16, device=x.device)
    b = _dev(bias, torch.float32, "bias") if bias is not None else None
    r = _dev(residual, torch.bfloat16, "residual") if residual is not None else None
    with torch.cuda.device(x.device):
        nat.check(nat.load().ppb_conv2d_bf16(nat.ptr(x), nat.ptr(w), nat.ptr(b), nat.ptr(r), nat.ptr(y), B, Hp - 2,
                                             Wp - 2, Cin, Cout, int(ksize), int(relu), _stream()))
    return y


# --------------------------------------------------------------------------- pose from votes
def pnp_ransac_votes(votes, offsets, obj_points, cam, iterations=300, reprojection_error=5.0, min_inliers=5):
    """Batched cv2.solvePnPRansac replacement over vote groups (ppb_pnp_ransac_votes). votes f32 [T,16];
    offsets i64 [G+1]; obj_points f32 [G,8,3]; cam f32 [G,4] (fx, fy, cx, cy).
    Returns rt f32 [G,6] (Rodrigues vector, translation), inliers i32 [G] (0 = no pose), rmse f32 [G]."""
    votes = _dev(votes, torch.float32, "votes")
    offsets = _dev(offsets, torch.int64, "offsets")
    obj_points = _dev(obj_points, torch.float32, "obj_points")
    cam = _dev(cam, torch.float32, "cam")
    G = int(offsets.shape[0]) - 1
    T = int(votes.shape[0])
    dev = offsets.device
    rt = torch.zeros((max(G, 0), 6), dtype=torch.float32, device=dev)
    inl = torch.zeros((max(G, 0),), dtype=torch.int32, device=dev)
    rmse = torch.zeros((max(G, 0),), dtype=torch.float32, device=dev)
    if G <= 0:
        return rt, inl, rmse
    if tuple(obj_points.shape) != (G, 8, 3) or tuple(cam.shape) != (G, 4):
        raise ValueError("obj_points must be [G,8,3] and cam [G,4] for G = %d groups" % G)
    lib = nat.load()
    lib.ppb_pnp_workspace_bytes.restype = ctypes.c_size_t
    need = int(lib.ppb_pnp_workspace_bytes(G, ctypes.c_longlong(T), int(iterations)))
    ws = _workspace(dev, need)
    with torch.cuda.device(dev):
        nat.check(lib.ppb_pnp_ransac_votes(nat.ptr(votes), nat.ptr(offsets), nat.ptr(obj_points), nat.ptr(cam), G,
                                           ctypes.c_longlong(T), int(iterations), ctypes.c_float(reprojection_error),
                                           int(min_inliers), nat.ptr(rt), nat.ptr(inl), nat.ptr(rmse), nat.ptr(ws),
                                           ctypes.c_size_t(ws.numel()), _stream()))
    return rt, inl, rmse
