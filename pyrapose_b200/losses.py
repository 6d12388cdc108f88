"""Loss descriptors — mirror of the two losses the reference compiles (PyraPose/bin/train.py:95-102):
focal (PyraPose/losses.py:22-68) and orthogonal_l1 (PyraPose/losses.py:321-408).

The reference returns TF functors; here each factory returns a small descriptor that PFPNModel.compile hands to the
native training step (loss value + gradient are computed by CUDA kernels, csrc/train_kernels.cu). The other losses
of losses.py are not compiled anywhere in the reference and are out of the hot path.
"""


class _Loss(object):
    def __init__(self, kind, **params):
        self.kind = kind
        self.params = params

    def __repr__(self):
        return "%s(%s)" % (self.kind, ", ".join("%s=%r" % kv for kv in sorted(self.params.items())))


def focal(alpha=0.25, gamma=2.0):
    """Focal loss on a sigmoid output with the anchor state in the last target column (-1 ignore, 0 bg, 1 object);
    normalised by max(1, #positive anchors of the batch)."""
    return _Loss("focal", alpha=float(alpha), gamma=float(gamma))


def orthogonal_l1(weight=0.125, sigma=3.0):
    """0.8 * smooth-L1(sigma) on the 16 corner regressions + 0.2 * mean |edge-parallelism difference| (24 terms),
    positives only, each summed and divided by max(1, #positives); times `weight`."""
    return _Loss("orthogonal_l1", weight=float(weight), sigma=float(sigma))
