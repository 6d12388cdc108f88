"""CPU restatement of the training objective of the hot path — TEST INFRASTRUCTURE ONLY.

PINNED TO THE REFERENCE'S TEXT: the reference's own losses.py, imported and run over numpy stand-ins for the
keras.backend calls it makes (tests/golden/tf_shim.py), gives the loss values in tests/golden/reference_text.npz;
tests/test_oracle_reference_text.py compares to 2e-6 relative (float32 reduction order). Keras / TensorFlow themselves
are not installable here, so keras.backend.binary_crossentropy is a restatement (below). These follow
  focal(alpha=0.25, gamma=2.0)          /root/reference/PyraPose/losses.py:22-68
  orthogonal_l1(weight=0.125, sigma=3)  /root/reference/PyraPose/losses.py:321-408
line by line in torch (float32, CPU) so that autograd supplies the reference gradients the CUDA loss kernels
are checked against. keras.backend.binary_crossentropy clips probabilities to [eps, 1 - eps], eps = 1e-7
(Keras 2.3.1 tensorflow_backend.binary_crossentropy), which is restated here.
"""
import numpy as np
import torch

KERAS_EPS = 1e-7

# The 12 (x, y) edge-parallelism differences of losses.py:338-361, as index quadruples (a, b, c, d) of the x
# coordinate: x_k = (r[a] - r[b]) - (r[c] - r[d]); the y twin uses every index + 1.
ORTH_QUADS = [
    (0, 6, 2, 4), (0, 6, 8, 14), (0, 2, 6, 4), (0, 2, 8, 10), (0, 8, 2, 10), (0, 8, 6, 14),
    (12, 10, 14, 8), (12, 10, 4, 2), (12, 4, 10, 2), (12, 4, 14, 6), (12, 14, 4, 6), (12, 14, 10, 8),
]


def focal(y_true, y_pred, alpha=0.25, gamma=2.0):
    """y_true (B, N, C+1) with the anchor state in the last column, y_pred (B, N, C) probabilities -> scalar."""
    labels = y_true[:, :, :-1]
    state = y_true[:, :, -1]
    keep = state != -1                                            # losses.py:48-50
    labels = labels[keep]
    p = y_pred[keep]
    alpha_factor = torch.where(labels == 1, torch.full_like(labels, alpha), torch.full_like(labels, 1 - alpha))
    focal_weight = torch.where(labels == 1, 1 - p, p)
    focal_weight = alpha_factor * focal_weight ** gamma          # losses.py:53-56
    pc = torch.clamp(p, KERAS_EPS, 1 - KERAS_EPS)
    bce = -(labels * torch.log(pc) + (1 - labels) * torch.log(1 - pc))
    cls_loss = focal_weight * bce                                 # losses.py:58
    normalizer = torch.clamp((state == 1).sum().to(y_pred.dtype), min=1.0)   # losses.py:61-63
    return cls_loss.sum() / normalizer


def _orths(r):
    cols = []
    for (a, b, c, d) in ORTH_QUADS:
        cols.append((r[:, a] - r[:, b]) - (r[:, c] - r[:, d]))
        cols.append((r[:, a + 1] - r[:, b + 1]) - (r[:, c + 1] - r[:, d + 1]))
    return torch.stack(cols, dim=1)


def orthogonal_l1(y_true, y_pred, weight=0.125, sigma=3.0):
    """y_true (B, N, 17): 16 regression targets + anchor state; y_pred (B, N, 16) raw regression -> scalar."""
    weight_xy, weight_orth = 0.8, 0.2
    sigma_squared = sigma ** 2
    state = y_true[:, :, -1]
    pos = state == 1                                              # losses.py:334-336
    reg = y_pred[pos]
    tgt = y_true[:, :, :-1][pos]
    diff = (reg - tgt).abs()
    xy = torch.where(diff < 1.0 / sigma_squared, 0.5 * sigma_squared * diff ** 2, diff - 0.5 / sigma_squared)
    orth = (_orths(reg) - _orths(tgt)).abs().mean(dim=1)          # keras.losses.mean_absolute_error over 24 values
    normalizer = max(1, int(pos.sum()))                           # losses.py:401-402
    return weight * (weight_xy * xy.sum() / normalizer + weight_orth * orth.sum() / normalizer)


def losses_and_grads(reg_true, cls_true, mask_true, reg_pred, cls_pred, mask_pred):
    """numpy in / numpy out: (loss_3Dbox, loss_cls, loss_mask) and dLoss/dpred for the three outputs (float32).
    For cls / mask the gradient is with respect to the PROBABILITY (the model output); the chain through the
    sigmoid is the network's job."""
    t = [torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)) for a in
         (reg_true, cls_true, mask_true, reg_pred, cls_pred, mask_pred)]
    preds = [x.clone().requires_grad_(True) for x in t[3:]]
    l_reg = orthogonal_l1(t[0], preds[0])
    l_cls = focal(t[1], preds[1])
    l_mask = focal(t[2], preds[2])
    (l_reg + l_cls + l_mask).backward()
    grads = [p.grad.numpy() if p.grad is not None else np.zeros(tuple(p.shape), np.float32) for p in preds]
    return (float(l_reg), float(l_cls), float(l_mask)), grads


def logit_grad_from_prob_grad(prob, dprob):
    """d/dlogit of a sigmoid output: dprob * p * (1 - p)."""
    return dprob * prob * (1.0 - prob)
