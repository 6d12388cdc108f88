"""Golden vectors for the TensorFlow / Keras side of the path, made by IMPORTING AND RUNNING the reference's own modules
over the numpy stand-ins of tests/golden/tf_shim.py (see its header for exactly what is the reference's text and what
is a restated library primitive). Run in the build container (needs /root/reference and oracle/_ref):

    python tests/golden/make_golden_shim.py          -> tests/golden/reference_text.npz (+ sha256 lines printed)

Executed reference code (nothing of it is copied into the repo; only inputs and outputs are stored):
  PyraPose/backend/common.py        bbox_transform_inv :6-22, box3D_transform_inv :25-56, box3D_denormalization :59-90,
                                    shift :93-116
  PyraPose/layers/_misc.py          Anchors.call :58-71, UpsampleLike.call :99-115, RegressBoxes3D.call :195-197,
                                    ClipBoxes.call :263-277
  PyraPose/layers/filter_detections.py   filter_detections :21-118, FilterDetections.call :153-191
  PyraPose/losses.py                focal :22-68, orthogonal_l1 :321-408
  PyraPose/utils/image.py           preprocess_image :35-62, compute_resize_scale :281-305, resize_image :308-323
  PyraPose/preprocessing/generator.py   Generator.compute_inputs :320-336
  PyraPose/utils/anchors.py         anchor_targets_bbox :72-287 (with compute_gt_annotations, box3D_transform,
                                    toPix_array, guess_shapes, anchors_for_shape), compute_overlap = the reference's
                                    Cython file compiled into oracle/_ref
Inputs are seeded; the ones the tests cannot regenerate from a seed are stored next to the outputs."""
import hashlib
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


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def synthetic_dets(seed, N, C, logit_mean, logit_std=1.5):
    """One image of tie-free scores, hull-like boxes, random 16-vectors (the generator of tests/test_kernels_gpu.py)."""
    rng = np.random.default_rng(seed)
    cls = (1.0 / (1.0 + np.exp(-rng.normal(logit_mean, logit_std, (N, C))))).astype(np.float32)
    flat = cls.reshape(-1)
    while np.unique(flat).size != flat.size:
        _, first = np.unique(flat, return_index=True)
        dup = np.ones(flat.size, bool)
        dup[first] = False
        flat[dup] = np.nextafter(flat[dup], np.float32(1.0)) + rng.uniform(0, 1e-6, dup.sum()).astype(np.float32)
    ctr = rng.uniform(0, 640, (N, 2))
    wh = rng.uniform(8, 160, (N, 2))
    boxes = np.concatenate([ctr - wh / 2, ctr + wh / 2], axis=1).astype(np.float32)
    b3 = rng.normal(0, 100, (N, 16)).astype(np.float32)
    return boxes, b3, cls


def main():
    import tf_shim
    from oracle import build as obuild
    from synth import pose_annotations
    co = obuild.load_ref_compute_overlap()
    assert co is not None, "run oracle/build.py first (compiles the reference's compute_overlap.pyx into oracle/_ref)"
    com = types.ModuleType("PyraPose.utils.compute_overlap")
    com.compute_overlap = co
    added = tf_shim.install(compute_overlap_module=com)
    try:
        backend = importlib.import_module("PyraPose.backend")
        losses = importlib.import_module("PyraPose.losses")
        ref_anchors = importlib.import_module("PyraPose.utils.anchors")
        fd = importlib.import_module("PyraPose.layers.filter_detections")
        misc = importlib.import_module("PyraPose.layers._misc")
        out, hashes = {}, {}
        A = lambda x: np.array(np.asarray(x))   # noqa: E731  (drop the shim's ndarray subclass)
        t = tf_shim.t                            # tensors handed to the reference's TF-side code

        # ---------------------------------------------------------------- decode (backend/common.py)
        rng = np.random.default_rng(11)
        a480 = ref_anchors.anchors_for_shape((480, 640, 3)).astype(np.float32)
        pick = rng.integers(0, a480.shape[0], 400)
        anc = np.stack([a480[pick], a480[pick[::-1]]])                      # (2, 400, 4)
        d16 = rng.normal(0, 1, (2, 400, 16)).astype(np.float32)
        mean = rng.normal(0, 0.1, 16).astype(np.float32)
        std = rng.uniform(0.1, 0.3, 16).astype(np.float32)
        out["dec_anchors"], out["dec_deltas"], out["dec_mean"], out["dec_std"] = anc, d16, mean, std
        out["dec_box3D"] = A(backend.box3D_transform_inv(t(anc), t(d16)))
        out["dec_box3D_meanstd"] = A(backend.box3D_transform_inv(t(anc), t(d16), mean=list(mean), std=list(std)))
        out["dec_bbox"] = A(backend.bbox_transform_inv(t(anc), t(d16[:, :, :4])))
        out["dec_denorm"] = A(backend.box3D_denormalization(t(anc), t(d16)))
        # the layers that wrap them
        out["layer_regress3D"] = A(misc.RegressBoxes3D().call([t(anc), t(d16)]))
        img = np.zeros((2, 480, 640, 3), np.float32)
        wide = (anc + rng.normal(0, 300, anc.shape)).astype(np.float32)
        out["clip_in"] = wide
        out["layer_clip"] = A(misc.ClipBoxes().call([t(img), t(wide)]))
        # ---------------------------------------------------------------- Anchors layer (shift) per level + concatenated
        prm = ref_anchors.AnchorParameters.default
        per_level = []
        for size, stride, (fh, fw) in zip(prm.sizes, prm.strides, ((60, 80), (30, 40), (15, 20))):
            layer = misc.Anchors(size=size, stride=stride, ratios=prm.ratios, scales=prm.scales)
            per_level.append(A(layer.call(t(np.zeros((1, fh, fw, 4), np.float32))))[0])
        full = np.concatenate(per_level, axis=0)
        assert full.dtype == np.float32 and full.shape == (56700, 4)
        hashes["anchors_layer_480x640_f32"] = sha(full)
        out["anchors_layer_head"], out["anchors_layer_tail"] = full[:90], full[-90:]
        odd = A(misc.Anchors(size=32, stride=8, ratios=prm.ratios, scales=prm.scales).call(
            t(np.zeros((2, 7, 5, 3), np.float32))))
        out["anchors_layer_odd_7x5"] = odd
        # ---------------------------------------------------------------- UpsampleLike (nearest)
        src = rng.normal(0, 1, (2, 15, 20, 3)).astype(np.float32)
        out["up_src"] = src
        out["up_30x40"] = A(misc.UpsampleLike().call([t(src), t(np.zeros((2, 30, 40, 1), np.float32))]))
        out["up_17x23"] = A(misc.UpsampleLike().call([t(src), t(np.zeros((2, 17, 23, 1), np.float32))]))
        # ---------------------------------------------------------------- filter_detections
        cases = [("default", dict()), ("agnostic", dict(class_specific_filter=False)), ("no_nms", dict(nms=False)),
                 ("no_nms_agnostic", dict(nms=False, class_specific_filter=False)),
                 ("tight", dict(score_threshold=0.5, max_detections=100, nms_threshold=0.3))]
        boxes, b3, cls = synthetic_dets(21, 1500, 4, -3.0)
        out["fd_boxes"], out["fd_boxes3D"], out["fd_cls"] = boxes, b3, cls
        other = rng.normal(0, 1, (1500, 3)).astype(np.float32)
        out["fd_other"] = other
        for name, kw in cases:
            r = fd.filter_detections(t(boxes), t(b3), t(cls), other=[t(other)], **kw)
            for key, val in zip(("boxes", "boxes3D", "scores", "labels", "other"), r):
                out["fd_%s_%s" % (name, key)] = A(val)
        # dense: more candidates than max_detections in every class; empty: nothing above the threshold
        boxes_d, b3_d, cls_d = synthetic_dets(22, 900, 3, 0.5)
        out["fdd_in_boxes"], out["fdd_in_boxes3D"], out["fdd_in_cls"] = boxes_d, b3_d, cls_d
        r = fd.filter_detections(t(boxes_d), t(b3_d), t(cls_d), max_detections=50)
        for key, val in zip(("boxes", "boxes3D", "scores", "labels"), r):
            out["fdd_%s" % key] = A(val)
        r = fd.filter_detections(t(boxes), t(b3), t((cls * np.float32(0.01)).astype(np.float32)))
        out["fd_empty_labels"], out["fd_empty_scores"] = A(r[3]), A(r[2])
        # the layer: map over a batch of two different images
        boxes2, b32, cls2 = synthetic_dets(23, 1500, 4, -3.0)
        layer = fd.FilterDetections()
        r = layer.call([t(np.stack([boxes, boxes2])), t(np.stack([b3, b32])), t(np.stack([cls, cls2]))])
        out["fdl_boxes2"], out["fdl_boxes3D2"], out["fdl_cls2"] = boxes2, b32, cls2
        for key, val in zip(("boxes", "boxes3D", "scores", "labels"), r):
            out["fdl_%s" % key] = A(val)
        # ---------------------------------------------------------------- anchor_targets_bbox (the whole function)
        a64 = ref_anchors.anchors_for_shape((480, 640, 3))
        C = 15
        anns = [pose_annotations(300 + i, C) for i in range(2)] + [pose_annotations(302, C, n_obj=0)]
        if anns[2]['bboxes'].shape[0] != 0:
            anns[2] = {k: (v[:0] if isinstance(v, np.ndarray) else v) for k, v in anns[2].items()}
        images = [np.zeros((480, 640, 3), np.float32)] * 3
        reg, lab, msk = ref_anchors.anchor_targets_bbox(a64, images, anns, C)
        out["tgt_seeds"] = np.array([300, 301, 302])
        for name, arr in (("regression_3D", reg), ("labels", lab), ("mask", msk)):
            assert arr.dtype == np.float32
            hashes["targets_480x640_C15_%s" % name] = sha(arr)
        out["tgt_states"] = reg[:, :, -1].astype(np.int8)
        pos = np.argwhere(reg[:, :, -1] == 1)
        out["tgt_pos_index"] = pos
        out["tgt_pos_regression"] = reg[pos[:, 0], pos[:, 1], :16]
        out["tgt_pos_label"] = lab[pos[:, 0], pos[:, 1], :C].argmax(axis=1)
        out["tgt_mask_nonzero"] = np.argwhere(msk != 0).astype(np.int32)
        out["tgt_shapes"] = np.array([reg.shape, lab.shape, msk.shape])
        # a padded batch: image 1 smaller than the batch canvas (anchors whose centre lies outside it are ignored)
        images_p = [np.zeros((480, 640, 3), np.float32), np.zeros((360, 500, 3), np.float32)]
        reg_p, lab_p, _ = ref_anchors.anchor_targets_bbox(a64, images_p, anns[:2], C)
        out["tgt_pad_states"] = reg_p[:, :, -1].astype(np.int8)
        hashes["targets_padded_regression_3D"] = sha(reg_p)
        hashes["targets_padded_labels"] = sha(lab_p)
        # ---------------------------------------------------------------- losses on those targets
        rngl = np.random.default_rng(31)
        sub = np.sort(np.concatenate([pos[pos[:, 0] == 0][:, 1], rngl.integers(0, a64.shape[0], 3000)]))
        y_reg = reg[:2, sub]                                               # (2, n, 17): positives of image 0 + random
        y_cls = lab[:2, sub]
        p_reg = (y_reg[:, :, :16] + rngl.normal(0, 0.3, y_reg[:, :, :16].shape)).astype(np.float32)
        p_cls = (1.0 / (1.0 + np.exp(-rngl.normal(-3.0, 2.0, y_cls[:, :, :C].shape)))).astype(np.float32)
        out["loss_y_reg"], out["loss_y_cls"], out["loss_p_reg"], out["loss_p_cls"] = y_reg, y_cls, p_reg, p_cls
        out["loss_focal"] = np.array([float(A(losses.focal()(t(y_cls), t(p_cls))))])
        out["loss_focal_a5_g1"] = np.array([float(A(losses.focal(alpha=0.5, gamma=1.0)(t(y_cls), t(p_cls))))])
        out["loss_orth"] = np.array([float(A(losses.orthogonal_l1()(t(y_reg), t(p_reg))))])
        none = y_reg.copy()
        none[:, :, -1] = 0
        out["loss_orth_no_positives"] = np.array([float(A(losses.orthogonal_l1()(t(none), t(p_reg))))])
        nonec = y_cls.copy()
        nonec[:, :, -1] = np.where(nonec[:, :, -1] == 1, 0, nonec[:, :, -1])
        nonec[:, :, :C] = 0
        out["loss_focal_no_positives"] = np.array([float(A(losses.focal()(t(nonec), t(p_cls))))])
        # ---------------------------------------------------------------- input staging (utils/image.py, generator.py)
        ref_image = importlib.import_module("PyraPose.utils.image")
        ref_gen = importlib.import_module("PyraPose.preprocessing.generator")
        shapes = [(375, 1242), (500, 333), (480, 640), (720, 1280)]
        out["img_shapes"] = np.array(shapes)
        out["img_seed"] = np.array([77])
        rngi = np.random.default_rng(77)
        staged, scales = [], []
        for k, (h, w) in enumerate(shapes):
            raw = rngi.integers(0, 256, (h, w, 3)).astype(np.uint8)
            x = ref_image.preprocess_image(raw, mode='caffe')                    # image.py:35-62
            assert x.dtype == np.float32
            hashes["preprocess_%dx%d" % (h, w)] = sha(x)
            y, scale = ref_image.resize_image(x, min_side=480, max_side=640)     # image.py:308-323
            staged.append(y)
            scales.append(scale)
            out["img_resized_shape_%d" % k] = np.array(y.shape)
            out["img_resized_sample_%d" % k] = y[::7, ::11].copy()
        out["img_scales"] = np.array(scales, dtype=np.float64)
        fake_self = types.SimpleNamespace(batch_size=len(staged))
        batch = ref_gen.Generator.compute_inputs(fake_self, staged)              # generator.py:320-336
        out["img_batch_shape"] = np.array(batch.shape)
        out["img_batch_sample"] = batch[:, ::9, ::13].copy()
        out["img_batch_nonzero_rows"] = np.array([[int(np.abs(b).sum(axis=(1, 2)).nonzero()[0].max()) + 1,
                                                   int(np.abs(b).sum(axis=(0, 2)).nonzero()[0].max()) + 1] for b in batch])
    finally:
        tf_shim.uninstall(added)
    path = os.path.join(HERE, "reference_text.npz")
    np.savez_compressed(path, **out)
    with open(os.path.join(HERE, "reference_text_sha256.txt"), "w") as f:
        for k in sorted(hashes):
            f.write("%s %s\n" % (k, hashes[k]))
    print("wrote %s (%d bytes), %d arrays" % (path, os.path.getsize(path), len(out)))
    for k in sorted(hashes):
        print(k, hashes[k])


if __name__ == "__main__":
    main()
