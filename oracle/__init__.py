"""CPU oracle for the PyraPose hot path — TEST INFRASTRUCTURE ONLY.

Nothing in the product package (pyrapose_b200/) imports this. Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may use it, and only as the checker or the
reported CPU baseline.

Pinning status:
  * compute_overlap, compute_gt_annotations, anchors_for_shape, generate_anchors, shift, bbox_transform,
    box3D_transform: PINNED — checked against the reference's own code executed in the build container
    (Cython extension compiled from /root/reference/PyraPose/utils/compute_overlap.pyx into oracle/_ref/,
    numpy functions exec'd from /root/reference/PyraPose/utils/anchors.py); golden vectors committed under
    tests/golden/ with the generating script.
  * PFPN forward (keras-resnet 0.1.0 + Keras 2.3.1 / TensorFlow 2.1.1 graph), box3D_transform_inv,
    filter_detections (tf.image.non_max_suppression, tf.nn.top_k): PARITY UNPINNED — TensorFlow/Keras are not
    installable here and the reference has no tests or golden vectors; these are cited restatements.
"""
