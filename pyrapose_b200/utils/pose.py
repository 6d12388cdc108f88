"""Pose from corner votes — the evaluators' loop right after predict_on_batch, batched on the GPU.

Reference, per image and per class (PyraPose/utils/linemod_eval.py:307-334, 427-430, 479-485;
PyraPose_ROS_wrapper/scripts/pyrapose_node.py:584-659):
    idx = np.where(scores[0, :, c] > threshold); skip the class when it has fewer than `min_votes` hits
    est_points = boxes3D[0, idx, :] -> (k*8, 1, 2); obj_points = the class's 3D box corners repeated k times
    cv2.solvePnPRansac(obj_points, est_points, K, iterationsCount=300, reprojectionError=5.0, confidence=0.99,
                       flags=cv2.SOLVEPNP_ITERATIVE) -> rvec, tvec; R = cv2.Rodrigues(rvec)
Here: one score-mask launch, one host read of the (B, C) vote counts, one gather, and ONE batched RANSAC + refinement
over all (image, class) groups (csrc/pnp.cu). Only the poses come back to the host.
"""
import numpy as np


def rodrigues(rvec):
    """cv2.Rodrigues for a (3,) rotation vector -> (3, 3) matrix (float64)."""
    r = np.asarray(rvec, dtype=np.float64).reshape(3)
    th = np.linalg.norm(r)
    K = np.array([[0, -r[2], r[1]], [r[2], 0, -r[0]], [-r[1], r[0], 0]])
    if th < 1e-12:
        return np.eye(3) + K
    return np.eye(3) + np.sin(th) / th * K + (1 - np.cos(th)) / (th * th) * K.dot(K)


def estimate_poses(scores, boxes3D, threeD_boxes, camera, threshold=0.5, min_votes=10, iterations=300,
                   reprojection_error=5.0):
    """scores (B, N, C) and boxes3D (B, N, 16): CUDA tensors (the inference model's outputs). threeD_boxes (C, 8, 3):
    box corners per class in model units; camera = (fx, fy, cx, cy). Returns, per image, a list of dicts
    {cls, rvec (3,), R (3,3), t (3,), votes, inliers, confidence, rmse} for every class with >= min_votes votes and a
    RANSAC consensus (cv2's retval)."""
    import torch
    from .. import _kernels
    B, N, C = [int(v) for v in scores.shape]
    dev = scores.device
    mask, counts = _kernels.score_mask(scores, threshold)
    counts_h = counts.cpu().numpy().reshape(B, C)                        # the only synchronising read before the poses
    keep = counts_h >= max(int(min_votes), 1)
    offsets = np.zeros(B * C + 1, dtype=np.int64)
    offsets[1:] = np.cumsum(counts_h.reshape(-1))
    total = int(offsets[-1])
    out = [[] for _ in range(B)]
    if total == 0 or not keep.any():
        return out
    off_d = torch.from_numpy(offsets).to(dev)
    _, votes, conf = _kernels.vote_gather(mask, scores, boxes3D, off_d, total)
    gb, gc = np.nonzero(keep)
    flat = gb * C + gc
    # groups = the kept (image, class) pairs; their vote rows stay where vote_gather put them
    g_off = torch.from_numpy(np.stack([offsets[flat], offsets[flat + 1]], axis=1).astype(np.int64))
    G = len(flat)
    # ppb_pnp_ransac_votes wants offsets[g]..offsets[g+1]: repack the kept groups contiguously
    sizes = (g_off[:, 1] - g_off[:, 0]).numpy()
    new_off = np.zeros(G + 1, dtype=np.int64)
    new_off[1:] = np.cumsum(sizes)
    rows = torch.cat([torch.arange(int(a), int(b)) for a, b in g_off.tolist()]).to(dev)
    gvotes = votes.index_select(0, rows).contiguous()
    obj = torch.from_numpy(np.ascontiguousarray(np.asarray(threeD_boxes, dtype=np.float32)[gc])).to(dev)
    cam = torch.from_numpy(np.tile(np.asarray(camera, dtype=np.float32).reshape(1, 4), (G, 1))).to(dev)
    rt, inl, rmse = _kernels.pnp_ransac_votes(gvotes, torch.from_numpy(new_off).to(dev), obj, cam, iterations,
                                              reprojection_error)
    rt_h, inl_h, rmse_h, conf_h = rt.cpu().numpy(), inl.cpu().numpy(), rmse.cpu().numpy(), conf.cpu().numpy()
    for g in range(G):
        if inl_h[g] <= 0:
            continue
        b, c = int(gb[g]), int(gc[g])
        out[b].append({"cls": c, "rvec": rt_h[g, :3].copy(), "R": rodrigues(rt_h[g, :3]), "t": rt_h[g, 3:].copy(),
                       "votes": int(sizes[g]), "inliers": int(inl_h[g]), "confidence": float(conf_h[b, c]),
                       "rmse": float(rmse_h[g])})
    return out
