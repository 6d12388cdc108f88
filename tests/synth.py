"""Seeded synthetic inputs shared by tests and benches: GT boxes for config 3, pose annotations for
config 4 (random unit quaternions, t_z ~ U(400, 1500) mm, YCB-like box extents, ycbv intrinsics)."""
import numpy as np

YCBV_K = (1066.778, 1067.487, 312.9869, 241.3109)  # PyraPose_ROS_wrapper/scripts/pyrapose_node.py:56-59


def gt_boxes(seed, k=100, width=640, height=480):
    """config 3: x1~U(0,W-16), y1~U(0,H-16), w,h~U(16,256), clipped to the image; float64 (k,4)."""
    rng = np.random.default_rng(seed)
    x1 = rng.uniform(0, width - 16, k)
    y1 = rng.uniform(0, height - 16, k)
    w = rng.uniform(16, 256, k)
    h = rng.uniform(16, 256, k)
    return np.stack([x1, y1, np.minimum(x1 + w, width), np.minimum(y1 + h, height)], axis=1)


def pose_annotations(seed, num_classes, n_obj=None, width=640, height=480):
    """One image's annotation dict in the layout LinemodGenerator.load_annotations produces
    (PyraPose/preprocessing/linemod.py:203-277)."""
    rng = np.random.default_rng(seed)
    n = int(rng.integers(3, 9)) if n_obj is None else n_obj
    fx, fy, cx, cy = YCBV_K
    poses = np.zeros((n, 7))
    segs = np.zeros((n, 8, 3))
    bboxes = np.zeros((n, 4))
    mask = np.zeros((height, width), np.uint8)
    for i in range(n):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        tz = rng.uniform(400, 1500)
        u, v = rng.uniform(0.15 * width, 0.85 * width), rng.uniform(0.15 * height, 0.85 * height)
        poses[i, :3] = ((u - cx) * tz / fx, (v - cy) * tz / fy, tz)
        poses[i, 3:] = q
        ext = rng.uniform(40, 120, 3)
        segs[i] = np.array([[sx * ext[0], sy * ext[1], sz * ext[2]] for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)])
        r = 1.2 * ext.max() * fx / tz
        bboxes[i] = (max(u - r, 0), max(v - r, 0), min(u + r, width - 1), min(v + r, height - 1))
        x0, y0, x1, y1 = [int(t) for t in bboxes[i]]
        mask[y0 + (y1 - y0) // 4:y1 - (y1 - y0) // 4, x0 + (x1 - x0) // 4:x1 - (x1 - x0) // 4] = i + 1
    return {
        'bboxes': bboxes, 'labels': rng.integers(0, num_classes, n).astype(np.float64), 'poses': poses,
        'segmentations': segs, 'cam_params': np.tile(np.array([fx, fy, cx, cy]), (n, 1)),
        'mask': [mask], 'mask_ids': np.arange(1, n + 1),
    }
