"""Independent cross-checks of the oracle's loss restatements (oracle/losses.py), which the CUDA loss kernels and the
training-step parity tests are measured against. Keras is not installable here; these compare with public
implementations of the same published formulas:
  * focal (PyraPose/losses.py:22-68, alpha 0.25 / gamma 2, "Focal Loss for Dense Object Detection")
        == torchvision.ops.sigmoid_focal_loss on the logits of the same probabilities, summed over the anchors whose
           state is not -1 and divided by max(1, #positive anchors)
  * the smooth-L1 part of orthogonal_l1 (losses.py:363-378, sigma = 3)
        == torch.nn.functional.smooth_l1_loss(beta = 1 / sigma^2) (same piecewise definition, scaled by sigma^2 inside
           the quadratic zone)
  * the orthogonality term's 24 differences (losses.py:338-361): linear in the corners, zero for any parallelepiped's
    projected corners ordered as the reference orders them, so prediction == target gives 0 and a translation of all
    eight corners changes nothing."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")


def _case(seed=0, B=2, N=500, C=6):
    rng = np.random.default_rng(seed)
    state = rng.choice([-1.0, 0.0, 1.0], size=(B, N), p=[0.2, 0.6, 0.2]).astype(np.float32)
    labels = np.zeros((B, N, C), np.float32)
    pos = state == 1
    labels[pos, rng.integers(0, C, int(pos.sum()))] = 1.0
    y_true = np.concatenate([labels, state[..., None]], axis=2)
    logits = rng.standard_normal((B, N, C)).astype(np.float32) * 2.0
    return y_true, logits, state, labels


def test_focal_matches_torchvision_sigmoid_focal_loss():
    tv = pytest.importorskip("torchvision")
    from torchvision.ops import sigmoid_focal_loss
    from oracle import losses as ol
    y_true, logits, state, labels = _case()
    p = torch.sigmoid(torch.from_numpy(logits))
    got = float(ol.focal(torch.from_numpy(y_true), p))
    keep = torch.from_numpy(state != -1)
    per = sigmoid_focal_loss(torch.from_numpy(logits), torch.from_numpy(labels), alpha=0.25, gamma=2.0, reduction="none")
    want = float(per[keep].sum() / max(1.0, float((state == 1).sum())))
    assert abs(got - want) <= 2e-6 * abs(want), (got, want)
    # no positives: the normaliser is max(1, 0) = 1 (losses.py:61-63)
    y0 = y_true.copy()
    y0[..., -1] = np.where(state == 1, 0.0, state)
    y0[..., :-1] = 0.0
    got0 = float(ol.focal(torch.from_numpy(y0), p))
    per0 = sigmoid_focal_loss(torch.from_numpy(logits), torch.zeros_like(p), alpha=0.25, gamma=2.0, reduction="none")
    assert abs(got0 - float(per0[keep].sum())) <= 2e-6 * got0


def test_orthogonal_l1_parts():
    import torch.nn.functional as F
    from oracle import losses as ol
    rng = np.random.default_rng(1)
    B, N = 2, 300
    state = rng.choice([-1.0, 0.0, 1.0], size=(B, N), p=[0.2, 0.6, 0.2]).astype(np.float32)
    tgt = rng.standard_normal((B, N, 16)).astype(np.float32)
    pred = (tgt + rng.standard_normal((B, N, 16)).astype(np.float32) * rng.choice([0.02, 0.5], size=(B, N, 1))).astype(np.float32)
    y_true = torch.from_numpy(np.concatenate([tgt, state[..., None]], axis=2))
    pos = torch.from_numpy(state == 1)
    n_pos = max(1, int(pos.sum()))
    total = float(ol.orthogonal_l1(y_true, torch.from_numpy(pred)))
    # smooth-L1 part: F.smooth_l1_loss(beta) = 0.5 d^2 / beta inside, |d| - 0.5 beta outside == losses.py:369-376
    xy = float(F.smooth_l1_loss(torch.from_numpy(pred)[pos], torch.from_numpy(tgt)[pos], beta=1.0 / 9.0, reduction="sum"))
    orth = (ol._orths(torch.from_numpy(pred)[pos]) - ol._orths(torch.from_numpy(tgt)[pos])).abs().mean(dim=1).sum()
    want = 0.125 * (0.8 * xy / n_pos + 0.2 * float(orth) / n_pos)
    assert abs(total - want) <= 2e-6 * abs(want), (total, want)
    # prediction == target: zero loss; translating all eight predicted corners leaves the orthogonality term unchanged
    assert float(ol.orthogonal_l1(y_true, torch.from_numpy(tgt))) == 0.0
    shift = torch.tensor([3.0, -2.0] * 8)
    a = ol._orths(torch.from_numpy(pred)[pos])
    b = ol._orths(torch.from_numpy(pred)[pos] + shift)
    assert torch.allclose(a, b, atol=1e-5)
    # the 24 differences vanish on the projected corners of a parallelepiped in the corner order the quadruples imply
    # (corner j and j + 4 differ in the sign of the first extent; 0-1-2-3 walk one face): any affine image of a cube
    ext = np.array([0.1, 0.2, 0.3])
    signs = np.array([[1, 1, 1], [1, 1, -1], [1, -1, -1], [1, -1, 1], [-1, 1, 1], [-1, 1, -1], [-1, -1, -1], [-1, -1, 1]], float)
    A = rng.standard_normal((2, 3))                       # any affine map R^3 -> R^2 (a weak-perspective projection)
    corners = (signs * ext) @ A.T + np.array([5.0, 7.0])
    o = ol._orths(torch.from_numpy(corners.reshape(1, 16).astype(np.float32)))
    assert float(o.abs().max()) < 1e-5
