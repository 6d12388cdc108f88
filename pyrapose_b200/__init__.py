"""pyrapose_b200 — B200 (sm_100a) implementation of PyraPose's data-parallel hot path.

Sub-packages mirror the reference's module layout (PyraPose/models, PyraPose/layers,
PyraPose/utils, PyraPose/backend) so that `from pyrapose_b200 import models` is a drop-in for
`from PyraPose import models` on that path. Nothing here initialises CUDA at import time
(forked data-loader workers must stay clean, cf. PyraPose/bin/train.py:372-390).
"""
__version__ = "0.1.0"
