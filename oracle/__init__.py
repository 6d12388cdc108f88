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
  * box3D_transform_inv / bbox_transform_inv / box3D_denormalization, the Anchors / UpsampleLike / ClipBoxes layers,
    filter_detections + FilterDetections.call, focal / orthogonal_l1, anchor_targets_bbox as a whole: PINNED TO THE
    REFERENCE'S TEXT — the reference's own modules (backend/common.py, layers/_misc.py, layers/filter_detections.py,
    losses.py, utils/anchors.py) imported and run in the build container over numpy stand-ins for the tensorflow /
    keras primitives they call (tests/golden/tf_shim.py); vectors in tests/golden/reference_text.npz, generator
    committed, compared in tests/test_oracle_reference_text.py. The primitives in the stand-in (NMS-V3, top_k tie
    rule, Keras' binary_crossentropy, nearest resize, quat2mat) are restatements.
  * PFPN forward (keras-resnet 0.1.0 + Keras 2.3.1 / TensorFlow 2.1.1 graph) and the training step's autograd /
    optimizer: PARITY UNPINNED — TensorFlow/Keras are not installable here and the reference has no tests or golden
    vectors; these are cited restatements with independent cross-checks (HuggingFace ResNet, torchvision).
"""
