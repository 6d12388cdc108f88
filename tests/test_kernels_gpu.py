"""GPU parity of the bandwidth-class kernels (through the C-ABI) against the CPU oracle.
Bit-exact: IoU matrix, assignment, anchors, kept indices, labels, gathered values.
Tolerance 1e-4 relative (fp32 path, north_star): decoded corners (the only difference allowed is none in
practice — the kernels round every op separately like the oracle — but the bound is stated here)."""
import hashlib
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _t(a, dev):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def gt_boxes(seed, k=100, w=640, h=480):
    rng = np.random.default_rng(seed)
    x1 = rng.uniform(0, w - 16, k)
    y1 = rng.uniform(0, h - 16, k)
    bw = rng.uniform(16, 256, k)
    bh = rng.uniform(16, 256, k)
    return np.stack([x1, y1, np.minimum(x1 + bw, w), np.minimum(y1 + bh, h)], axis=1).astype(np.float64)


# ------------------------------------------------------------------ IoU
def test_compute_overlap_bit_exact_full_size(cuda_device):
    from oracle import anchors as oa, overlap as oo
    from pyrapose_b200 import _kernels as K
    sha = dict(l.split() for l in open(os.path.join(GOLD, "sha256.txt")))
    a = oa.anchors_for_shape((480, 640, 3))
    for seed in (1000, 1001):
        q = gt_boxes(seed)
        got = K.compute_overlap(_t(a, cuda_device), _t(q, cuda_device)).cpu().numpy()
        assert hashlib.sha256(got.tobytes()).hexdigest() == sha["iou_full_seed%d" % seed]  # reference Cython output
        assert np.array_equal(got, oo.compute_overlap(a, q))


@pytest.mark.parametrize("n,k", [(1, 1), (7, 3), (1000, 37), (513, 600), (5, 0), (0, 4)])
def test_compute_overlap_ragged_shapes(cuda_device, n, k):
    from oracle import overlap as oo
    from pyrapose_b200 import _kernels as K
    rng = np.random.default_rng(n * 31 + k)
    b = rng.uniform(-50, 700, (n, 4))
    b[:, 2:] = b[:, :2] + rng.uniform(-5, 300, (n, 2))   # includes inverted (degenerate) boxes
    q = rng.uniform(0, 600, (k, 4))
    q[:, 2:] = q[:, :2] + rng.uniform(0, 200, (k, 2))
    got = K.compute_overlap(_t(b, cuda_device), _t(q, cuda_device)).cpu().numpy()
    assert got.shape == (n, k)
    assert np.array_equal(got, oo.compute_overlap(b, q))


def test_compute_overlap_known_answers(cuda_device):
    from pyrapose_b200 import _kernels as K
    g = np.load(os.path.join(GOLD, "anchors_overlap.npz"))
    got = K.compute_overlap(_t(g["kat_boxes"], cuda_device), _t(g["kat_query"], cuda_device)).cpu().numpy()
    assert np.array_equal(got, g["kat_iou"])


def test_compute_overlap_batched(cuda_device):
    from oracle import anchors as oa, overlap as oo
    from pyrapose_b200 import _kernels as K
    a = oa.anchors_for_shape((480, 640, 3))[::7].copy()
    qs = np.stack([gt_boxes(1000 + i, k=50) for i in range(4)])
    got = K.compute_overlap(_t(a, cuda_device), _t(qs, cuda_device)).cpu().numpy()
    for i in range(4):
        assert np.array_equal(got[i], oo.compute_overlap(a, qs[i]))


def test_gt_assign_bit_exact(cuda_device):
    from oracle import anchors as oa, overlap as oo
    from pyrapose_b200 import _kernels as K
    import torch
    a = oa.anchors_for_shape((480, 640, 3))
    B = 6
    counts = [100, 1, 37, 100, 64, 3]
    qs = np.zeros((B, 100, 4))
    for i in range(B):
        qs[i, :counts[i]] = gt_boxes(1000 + i, k=counts[i])
    pos, ign, arg, mx = K.gt_assign(_t(a, cuda_device), _t(qs, cuda_device),
                                    torch.tensor(counts, dtype=torch.int32, device=cuda_device), want_max=True)
    for i in range(B):
        p, g, am, m = oo.gt_assign(a, qs[i, :counts[i]])
        p2, g2, am2 = oa.compute_gt_annotations(a, qs[i, :counts[i]])   # numpy path over the materialised matrix
        assert np.array_equal(p, p2) and np.array_equal(g, g2) and np.array_equal(am, am2)
        assert np.array_equal(pos[i].cpu().numpy(), p)
        assert np.array_equal(ign[i].cpu().numpy(), g)
        assert np.array_equal(arg[i].cpu().numpy(), am)
        assert np.array_equal(mx[i].cpu().numpy(), m)


# ------------------------------------------------------------------ anchors
@pytest.mark.parametrize("shape", [(480, 640), (960, 1280), (427, 561)])
def test_anchor_grids_bit_exact(cuda_device, shape):
    from oracle import anchors as oa
    from pyrapose_b200 import _kernels as K
    import torch
    p = oa.AnchorParameters.default
    shapes = oa.guess_shapes(shape, [3, 4, 5])
    lv64, lv32 = [], []
    for i, (fh, fw) in enumerate(shapes):
        base = oa.generate_anchors(p.sizes[i], p.ratios, p.scales)
        lv64.append(K.anchors_shift(_t(base, cuda_device), int(fh), int(fw), p.strides[i]))
        lv32.append(K.anchors_shift(_t(base.astype(np.float32), cuda_device), int(fh), int(fw), p.strides[i]))
    a64 = torch.cat(lv64).cpu().numpy()
    a32 = torch.cat(lv32).cpu().numpy()
    assert np.array_equal(a64, oa.anchors_for_shape(shape + (3,)))
    assert np.array_equal(a32, oa.anchors_layer_f32([(int(h), int(w)) for h, w in shapes]))


# ------------------------------------------------------------------ decode
def test_box3d_transform_inv(cuda_device):
    from oracle import anchors as oa, decode as od
    from pyrapose_b200 import _kernels as K
    rng = np.random.default_rng(5)
    anchors = oa.anchors_layer_f32([(60, 80), (30, 40), (15, 20)])
    B = 3
    d = rng.normal(0, 1, (B, anchors.shape[0], 16)).astype(np.float32)
    ab = np.broadcast_to(anchors[None], (B,) + anchors.shape)
    ref = od.box3D_transform_inv(ab, d)
    got, hull = K.box3d_transform_inv(_t(anchors, cuda_device), _t(d, cuda_device), want_hull=True)
    got = got.cpu().numpy()
    assert np.allclose(got, ref, rtol=1e-4, atol=1e-4)
    assert np.array_equal(got, ref), "expected bit-exact (separately rounded mul/add)"
    assert np.array_equal(hull.cpu().numpy(), od.hull_boxes(ref))
    # batched anchors + custom mean/std + denormalisation
    mean = rng.normal(0, 0.1, 16).astype(np.float32)
    std = rng.uniform(0.1, 0.3, 16).astype(np.float32)
    ref2 = od.box3D_denormalization(ab, d, mean, std)
    got2 = K.box3d_transform_inv(_t(np.ascontiguousarray(ab), cuda_device), _t(d, cuda_device), mean=mean, std=std,
                                 denormalize=True).cpu().numpy()
    assert np.allclose(got2, ref2, rtol=1e-4, atol=1e-5)
    # KAT: anchor [-16,-16,16,16] shifted by (4,4), all deltas 1 -> every corner offset 6.4
    a1 = np.array([[-12, -12, 20, 20]], np.float32)
    out = K.box3d_transform_inv(_t(a1, cuda_device), _t(np.ones((1, 1, 16), np.float32), cuda_device)).cpu().numpy()
    assert np.allclose(out[0, 0], np.tile(a1[0], 4) + np.float32(0.2) * 32, rtol=1e-6)


def test_bbox_transform_inv_and_clip(cuda_device):
    from oracle import anchors as oa, decode as od
    from pyrapose_b200 import _kernels as K
    rng = np.random.default_rng(6)
    anchors = oa.anchors_layer_f32([(15, 20)])
    d = rng.normal(0, 1, (2, anchors.shape[0], 4)).astype(np.float32)
    ab = np.broadcast_to(anchors[None], (2,) + anchors.shape)
    ref = od.bbox_transform_inv(ab, d)
    got = K.bbox_transform_inv(_t(anchors, cuda_device), _t(d, cuda_device)).cpu().numpy()
    assert np.array_equal(got, ref)
    assert np.array_equal(K.clip_boxes(_t(ref, cuda_device), 640, 480).cpu().numpy(), od.clip_boxes((480, 640), ref))


# ------------------------------------------------------------------ filter
def _synthetic_dets(seed, B, N, C, logit_mean, logit_std=1.5):
    """SURVEY §8(d) config-2 generator: i.i.d. logits, tie-free scores, random hull-like boxes."""
    rng = np.random.default_rng(seed)
    cls = (1.0 / (1.0 + np.exp(-rng.normal(logit_mean, logit_std, (B, N, C))))).astype(np.float32)
    # break ties: make every score in an image unique
    for b in range(B):
        flat = cls[b].reshape(-1)
        _, first = np.unique(flat, return_index=True)
        dup = np.ones(flat.size, bool)
        dup[first] = False
        while dup.any():
            flat[dup] = np.nextafter(flat[dup], np.float32(1.0)) + rng.uniform(0, 1e-6, dup.sum()).astype(np.float32)
            _, first = np.unique(flat, return_index=True)
            dup = np.ones(flat.size, bool)
            dup[first] = False
    ctr = rng.uniform(0, 640, (B, N, 2))
    wh = rng.uniform(8, 160, (B, N, 2))
    boxes = np.concatenate([ctr - wh / 2, ctr + wh / 2], axis=2).astype(np.float32)
    b3 = rng.normal(0, 100, (B, N, 16)).astype(np.float32)
    return boxes, b3, cls


def _check_filter(cuda_device, boxes, b3, cls, **kw):
    from oracle import filter as of
    from pyrapose_b200 import _kernels as K
    ref = of.filter_detections_batch(boxes, b3, cls, **kw)
    got = K.filter_detections(_t(boxes, cuda_device), _t(b3, cuda_device), _t(cls, cuda_device), **kw)
    names = ["boxes", "boxes3D", "scores", "labels"]
    for r, g, nm in zip(ref, got, names):
        g = g.cpu().numpy()
        assert g.dtype == r.dtype, nm
        assert np.array_equal(g, r), "%s differ (%s)" % (nm, kw)
    return ref


@pytest.mark.parametrize("kw", [
    dict(),
    dict(class_specific_filter=False),
    dict(nms=False),
    dict(nms=False, class_specific_filter=False),
    dict(score_threshold=0.5, max_detections=100, nms_threshold=0.3),
])
def test_filter_detections_matches_oracle(cuda_device, kw):
    boxes, b3, cls = _synthetic_dets(3, 2, 6000, 7, -4.0)
    ref = _check_filter(cuda_device, boxes, b3, cls, **kw)
    assert (ref[3] >= 0).sum() > 0


def test_filter_detections_full_size_sparse_and_dense(cuda_device):
    boxes, b3, cls = _synthetic_dets(4, 2, 56700, 15, -6.0)        # ~2% above 0.05
    _check_filter(cuda_device, boxes, b3, cls)
    boxes, b3, cls = _synthetic_dets(5, 1, 56700, 15, -4.6)        # dense stress: ~13% -> >4096 per class (global sort)
    _check_filter(cuda_device, boxes, b3, cls)


def _clustered_dets(seed, B, N, C, n_obj=6):
    """Scores and boxes shaped like a TRAINED detector's (i.i.d. logits make an easy NMS load): a few objects
    per image, every candidate near an object scores high for that object's class and regresses to nearly the same
    box (mutual IoU 0.6-0.95), everything else sits at the background prior. Tie-free."""
    rng = np.random.default_rng(seed)
    ctr = rng.uniform(0, 640, (B, N, 2))
    wh = rng.uniform(16, 120, (B, N, 2))
    boxes = np.concatenate([ctr - wh / 2, ctr + wh / 2], axis=2)
    logits = rng.normal(-7.0, 1.0, (B, N, C))
    for b in range(B):
        for _ in range(n_obj):
            oc, ow = rng.uniform(100, 540, 2), rng.uniform(60, 160, 2)
            cls = int(rng.integers(0, C))
            near = np.nonzero(np.abs(ctr[b] - oc).max(axis=1) < ow.max() * 0.4)[0]
            jit = rng.normal(0, 0.04, (near.size, 4)) * np.concatenate([ow, ow])
            boxes[b, near] = np.concatenate([oc - ow / 2, oc + ow / 2]) + jit
            logits[b, near, cls] = rng.normal(1.0, 1.0, near.size)
    cls_p = (1.0 / (1.0 + np.exp(-logits))).astype(np.float32)
    for b in range(B):
        flat = cls_p[b].reshape(-1)
        while np.unique(flat).size != flat.size:
            _, first = np.unique(flat, return_index=True)
            dup = np.ones(flat.size, bool)
            dup[first] = False
            flat[dup] = np.nextafter(flat[dup], np.float32(1.0)) + rng.uniform(0, 1e-6, dup.sum()).astype(np.float32)
    b3 = rng.normal(0, 100, (B, N, 16)).astype(np.float32)
    return boxes.astype(np.float32), b3, cls_p


def test_filter_detections_clustered_candidates(cuda_device):
    """Long suppression chains: hundreds of candidates per object that nearly coincide, so almost every candidate is
    suppressed by an earlier one of its cluster (the greedy resolve has to walk deep into each 256-chunk)."""
    boxes, b3, cls = _clustered_dets(9, 2, 20000, 6)
    ref = _check_filter(cuda_device, boxes, b3, cls)
    kept = int((ref[3] >= 0).sum())
    cand = int((cls > 0.05).sum())
    print("clustered set: %d candidates above 0.05, %d kept after NMS" % (cand, kept))
    assert cand > 20 * kept > 0
    _check_filter(cuda_device, boxes, b3, cls, class_specific_filter=False)
    _check_filter(cuda_device, boxes, b3, cls, nms_threshold=0.3, max_detections=20)


def test_filter_detections_edge_cases(cuda_device):
    from pyrapose_b200 import _kernels as K
    import torch
    # nothing above threshold -> everything -1
    boxes, b3, cls = _synthetic_dets(6, 2, 500, 3, -12.0)
    out = K.filter_detections(_t(boxes, cuda_device), _t(b3, cuda_device), _t(cls, cuda_device))
    assert all(bool((o == -1).all()) for o in out)
    assert out[3].dtype == torch.int32
    # two boxes with IoU exactly 0.5 are NOT suppressed (strict >); a third with IoU > 0.5 is
    bx = np.array([[[0, 0, 10, 10], [0, 0, 10, 5], [0, 0, 10, 9]]], np.float32)
    sc = np.array([[[0.9], [0.8], [0.7]]], np.float32)
    ref = _check_filter(cuda_device, bx, np.zeros((1, 3, 16), np.float32), sc)
    assert (ref[3][0] >= 0).sum() == 2
    # zero-area boxes never suppress and are never suppressed
    bx = np.array([[[5, 5, 5, 9], [5, 5, 5, 9], [0, 0, 10, 10]]], np.float32)
    ref = _check_filter(cuda_device, bx, np.zeros((1, 3, 16), np.float32), sc)
    assert (ref[3][0] >= 0).sum() == 3
    # > max_detections survivors in one class and in total: truncated per class, then top-k across classes
    rng = np.random.default_rng(9)
    N = 2000
    ctr = np.stack([np.arange(N) * 40.0, np.zeros(N)], 1)
    bx = np.concatenate([ctr, ctr + 10], 1).astype(np.float32)[None]          # disjoint boxes: nothing suppressed
    sc = rng.permutation(np.linspace(0.1, 0.9, N * 2)).reshape(1, N, 2).astype(np.float32)
    ref = _check_filter(cuda_device, bx, rng.normal(size=(1, N, 16)).astype(np.float32), sc, max_detections=300)
    assert (ref[3][0] >= 0).sum() == 300
    with pytest.raises(ValueError):
        K.filter_detections(_t(bx, cuda_device), _t(np.zeros((1, N, 16), np.float32), cuda_device), _t(sc, cuda_device),
                            max_detections=5000)


def test_filter_other_gather(cuda_device):
    from pyrapose_b200 import _kernels as K
    boxes, b3, cls = _synthetic_dets(7, 2, 3000, 4, -4.0)
    other = np.random.default_rng(1).normal(size=(2, 3000, 5)).astype(np.float32)
    out = K.filter_detections(_t(boxes, cuda_device), _t(b3, cuda_device), _t(cls, cuda_device), want_indices=True)
    idx = out[4].cpu().numpy()
    g = K.gather_rows(_t(other, cuda_device), out[4]).cpu().numpy()
    for b in range(2):
        for d in range(300):
            exp = other[b, idx[b, d]] if idx[b, d] >= 0 else -np.ones(5, np.float32)
            assert np.array_equal(g[b, d], exp)


def test_vote_filter(cuda_device):
    from oracle import filter as of
    from pyrapose_b200 import _kernels as K
    import torch
    boxes, b3, cls = _synthetic_dets(8, 3, 56700, 15, -3.0)
    mask, counts = K.score_mask(_t(cls, cuda_device), 0.5)
    c = counts.cpu().numpy()
    offsets = np.zeros(c.size + 1, np.int64)
    offsets[1:] = np.cumsum(c.reshape(-1))
    idx, votes, conf = K.vote_gather(mask, _t(cls, cuda_device), _t(b3, cuda_device), _t(offsets, cuda_device),
                                     int(offsets[-1]))
    idx = idx.cpu().numpy()
    votes = votes.cpu().numpy()
    conf = conf.cpu().numpy()
    for b in range(3):
        ref = of.vote_filter(cls[b], b3[b], threshold=0.5, min_votes=0)
        for cl in range(15):
            s, e = offsets[b * 15 + cl], offsets[b * 15 + cl + 1]
            assert np.array_equal(idx[s:e], ref[cl][0])
            assert idx[s:e].dtype == np.int64
            assert np.array_equal(votes[s:e], ref[cl][1])
            assert np.isclose(conf[b, cl], cls[b, ref[cl][0], cl].sum(), rtol=1e-5)


@pytest.mark.parametrize("B,N,C", [(3, 333, 5), (1, 301, 3), (2, 1000, 40), (2, 257, 50), (1, 31, 1), (2, 4097, 15)])
def test_score_mask_ragged_shapes(cuda_device, B, N, C):
    """Every staging path of the score pass: bulk-copy slab (aligned), bulk + scalar tail (slab bytes % 16 != 0),
    misaligned image base (plain loads), C > 32 (two ballot words per row) and C > 44 (one-warp tiles).
    Bit-exact against numpy `scores > thr` packed LSB-first per 32 anchors (linemod_eval.py:317-319 semantics)."""
    from pyrapose_b200 import _kernels as K
    rng = np.random.default_rng(B * 1000 + N + C)
    sc = rng.uniform(0, 1, (B, N, C)).astype(np.float32)
    sc[0, 0, 0] = 0.5  # exactly at the threshold: strict '>' must drop it
    mask, counts = K.score_mask(_t(sc, cuda_device), 0.5)
    W = (N + 31) // 32
    hit = sc > np.float32(0.5)
    pad = np.zeros((B, W * 32, C), bool)
    pad[:, :N] = hit
    ref = (pad.reshape(B, W, 32, C).astype(np.uint64) << np.arange(32, dtype=np.uint64)[None, None, :, None]).sum(2)
    ref = ref.transpose(0, 2, 1).astype(np.uint32)  # [B, C, W]
    got = mask.cpu().numpy().view(np.uint32)
    assert got.shape == ref.shape
    assert np.array_equal(got, ref)
    assert np.array_equal(counts.cpu().numpy(), hit.sum(1).astype(np.int32))
