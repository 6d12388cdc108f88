"""Host-side mirror of PyraPose/utils for the hot path (anchors, compute_overlap, image preprocessing)."""
