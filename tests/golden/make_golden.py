"""Generates the golden vectors under tests/golden/ from the REFERENCE's own code.

Run in the build container (needs /root/reference; it does not travel to the GPU box):
    python tests/golden/make_golden.py

Sources of truth executed here:
  * /root/reference/PyraPose/utils/compute_overlap.pyx   compiled by oracle/build.py into oracle/_ref/
  * /root/reference/PyraPose/utils/anchors.py            numpy functions exec'd from the source text past
                                                          the module-level `import keras`/`transforms3d`
                                                          (stubbed: only keras.backend.floatx() is used)
Nothing from the reference is copied into the repo; only inputs/outputs are stored.
"""
import hashlib
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference/PyraPose"


def load_reference_anchors_module():
    from oracle import build as obuild
    ref_co = obuild.load_ref_compute_overlap()
    assert ref_co is not None, "run oracle/build.py first"
    keras = types.ModuleType("keras")
    keras.backend = types.SimpleNamespace(floatx=lambda: "float32")
    stubs = {"keras": keras, "transforms3d": types.ModuleType("transforms3d"), "cv2": types.ModuleType("cv2"),
             "PIL": types.ModuleType("PIL")}
    stubs["PIL"].Image = None
    src = open(os.path.join(REF, "utils/anchors.py")).read()
    src = src.replace("from ..utils.compute_overlap import compute_overlap", "")
    src = src.replace("from PIL import Image", "Image = None")
    saved = {k: sys.modules.get(k) for k in stubs}
    sys.modules.update(stubs)
    try:
        mod = types.ModuleType("ref_anchors")
        mod.compute_overlap = ref_co
        exec(compile(src, "ref_anchors.py", "exec"), mod.__dict__)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod, ref_co


def gt_boxes(seed, k=100, w=640, h=480):
    """SURVEY.md §8(d) config 3 generator."""
    rng = np.random.default_rng(seed)
    x1 = rng.uniform(0, w - 16, k)
    y1 = rng.uniform(0, h - 16, k)
    bw = rng.uniform(16, 256, k)
    bh = rng.uniform(16, 256, k)
    x2 = np.minimum(x1 + bw, w)
    y2 = np.minimum(y1 + bh, h)
    return np.stack([x1, y1, x2, y2], axis=1).astype(np.float64)


def main():
    ref, ref_co = load_reference_anchors_module()
    out = {}
    # ---- anchors
    for size in (32, 64, 128):
        out["generate_anchors_%d" % size] = ref.generate_anchors(base_size=size)
    a480 = ref.anchors_for_shape((480, 640, 3))
    a960 = ref.anchors_for_shape((960, 1280, 3))
    a_odd = ref.anchors_for_shape((427, 561, 3))
    five = ref.AnchorParameters(sizes=[32, 64, 128, 256, 512], strides=[8, 16, 32, 64, 128],
                                ratios=ref.AnchorParameters.default.ratios, scales=ref.AnchorParameters.default.scales)
    a5 = ref.anchors_for_shape((480, 640, 3), pyramid_levels=[3, 4, 5, 6, 7], anchor_params=five)
    out["anchors_480x640_head"] = a480[:64]
    out["anchors_480x640_tail"] = a480[-64:]
    out["anchors_480x640_sum"] = np.array([a480.sum()])
    out["anchors_960x1280_sum"] = np.array([a960.sum()])
    out["anchors_odd_427x561"] = a_odd[::97]
    out["anchors_odd_shape"] = np.array(a_odd.shape)
    out["anchors_5level_shape"] = np.array(a5.shape)
    out["anchors_5level_sum"] = np.array([a5.sum()])
    sha = {
        "anchors_480x640": hashlib.sha256(a480.tobytes()).hexdigest(),
        "anchors_960x1280": hashlib.sha256(a960.tobytes()).hexdigest(),
        "anchors_5level_480x640": hashlib.sha256(a5.tobytes()).hexdigest(),
        "anchors_odd_427x561": hashlib.sha256(a_odd.tobytes()).hexdigest(),
    }
    # ---- IoU matrix + assignment on the config-3 generator (subsampled anchors keep the file small)
    q = gt_boxes(1000)
    sub = a480[::23].copy()  # 2466 anchors
    ov = ref_co(sub, q)
    pos, ign, arg = ref.compute_gt_annotations(sub, q)
    out["iou_query"] = q
    out["iou_anchor_stride"] = np.array([23])
    out["iou_matrix"] = ov
    out["assign_positive"] = pos
    out["assign_ignore"] = ign
    out["assign_argmax"] = arg
    # full-size assignment, stored as hashes + counts
    for seed in (1000, 1001):
        qf = gt_boxes(seed)
        ovf = ref_co(a480, qf)
        p, i, a = ref.compute_gt_annotations(a480, qf)
        sha["iou_full_seed%d" % seed] = hashlib.sha256(ovf.tobytes()).hexdigest()
        sha["assign_full_seed%d" % seed] = hashlib.sha256(
            p.astype(np.uint8).tobytes() + i.astype(np.uint8).tobytes() + a.astype(np.int64).tobytes()).hexdigest()
        out["assign_full_counts_seed%d" % seed] = np.array([p.sum(), i.sum()])
    # ---- known-answer IoU cases
    kat_b = np.array([[0, 0, 9, 9], [0, 0, 9, 9], [0, 0, 9, 9], [0, 0, 9, 9]], dtype=np.float64)
    kat_q = np.array([[0, 0, 9, 9], [10, 0, 19, 9], [5, 0, 14, 9], [9, 0, 18, 9]], dtype=np.float64)
    out["kat_boxes"] = kat_b
    out["kat_query"] = kat_q
    out["kat_iou"] = ref_co(kat_b, kat_q)
    # ---- regression targets
    rng = np.random.default_rng(7)
    anc = a480[rng.integers(0, a480.shape[0], 200)]
    gt4 = gt_boxes(5, k=200)
    gt16 = rng.uniform(0, 640, (200, 16))
    out["xform_anchors"] = anc
    out["xform_gt4"] = gt4
    out["xform_gt16"] = gt16
    out["bbox_transform"] = ref.bbox_transform(anc, gt4)
    out["box3D_transform"] = ref.box3D_transform(anc, gt16, 15)
    out["toPix_in"] = rng.uniform(100, 1000, (8, 3))
    out["toPix_out"] = ref.toPix_array(out["toPix_in"], fx=572.4114, fy=573.57043, cx=325.2611, cy=242.04899)
    np.savez_compressed(os.path.join(HERE, "anchors_overlap.npz"), **out)
    with open(os.path.join(HERE, "sha256.txt"), "w") as f:
        for k in sorted(sha):
            f.write("%s %s\n" % (k, sha[k]))
    print("wrote", os.path.join(HERE, "anchors_overlap.npz"), os.path.getsize(os.path.join(HERE, "anchors_overlap.npz")))
    for k in sorted(sha):
        print(k, sha[k])


if __name__ == "__main__":
    main()
