"""CPU restatement of ONE training step of the reference — TEST INFRASTRUCTURE ONLY (PARITY UNPINNED: Keras /
TensorFlow are not installable here; the reference has no tests).

Follows /root/reference/PyraPose/bin/train.py:95-102 (compile: losses + Adam(lr, clipnorm=0.001)),
/root/reference/PyraPose/models/resnet.py:87,100-103 (BatchNorm frozen; non-BN layers with index < 40 frozen = conv1
and stage 2 of keras-resnet 0.1.0's ResNet50), /root/reference/PyraPose/models/retinanet.py:108 (l2(0.001) kernel
regulariser on the five convolutions of the regression sub-model, counted once: the sub-model is one shared Keras
model) and Keras 2.3.1 optimizers.py (Optimizer.get_gradients: clip by the GLOBAL norm when norm >= clipnorm;
Adam.get_updates: lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t), p -= lr_t * m / (sqrt(v) + 1e-7)).
torch autograd (float32, CPU) over oracle/pfpn.py + oracle/losses.py supplies the gradients.
"""
import numpy as np
import torch

from . import losses as olosses
from . import pfpn

L2_REG = 0.001


def _dev(t):
    """pfpn.DEVICE: None = CPU; a GPU test may run the same autograd graph through torch's library kernels."""
    return t.to(pfpn.DEVICE[0]) if pfpn.DEVICE[0] is not None else t


class _RoundBF16(torch.autograd.Function):
    """bf16 round-to-nearest-even of an activation; the gradient map coming back is rounded the same way (the native
    engine stores both as bf16)."""

    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).to(torch.float32)

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).to(torch.float32)


class _PinLinear(torch.autograd.Function):
    """Replaces a stored activation by the tensor the native engine holds for it; the gradient passes through
    (rounded to bf16 like the engine's gradient maps). Used by step(pinned=...)."""

    @staticmethod
    def forward(ctx, x, pinned):
        return pinned

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).to(torch.float32), None


class _PinRelu(torch.autograd.Function):
    """Same for a ReLU output: the backward mask is the ENGINE's (pinned > 0), so both sides differentiate the same
    piecewise-linear function."""

    @staticmethod
    def forward(ctx, z, pinned):
        ctx.save_for_backward(pinned)
        return pinned

    @staticmethod
    def backward(ctx, g):
        (pinned,) = ctx.saved_tensors
        return (g * (pinned > 0)).to(torch.bfloat16).to(torch.float32), None


class _RoundWeightSTE(torch.autograd.Function):
    """bf16 round of a (BatchNorm-folded) kernel, straight-through gradient (fp32 weight gradients)."""

    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).to(torch.float32)

    @staticmethod
    def backward(ctx, g):
        return g


def _emulated_weights(wt):
    """The tensors the native engine actually multiplies with: kernel * frozen-BN scale rounded to bf16, BatchNorm
    replaced by its fp32 shift (engine.cu set_layer)."""
    out = {}
    bn_of = {}
    for k in wt:
        layer, param = k.split("/")
        if param == "gamma":
            conv = "conv1" if layer == "bn_conv1" else "res" + layer[2:]
            bn_of[conv] = layer
    for k, v in wt.items():
        layer, param = k.split("/")
        if param == "kernel":
            if layer in bn_of:
                bn = bn_of[layer]
                scale = wt[bn + "/gamma"] / torch.sqrt(wt[bn + "/moving_variance"] + pfpn.BN_EPS)
                shift = wt[bn + "/beta"] - wt[bn + "/moving_mean"] * scale
                out[k] = _RoundWeightSTE.apply(v * scale.view(1, 1, 1, -1))
                out[bn + "/gamma"] = torch.ones_like(scale)
                out[bn + "/beta"] = shift
                out[bn + "/moving_mean"] = torch.zeros_like(scale)
                out[bn + "/moving_variance"] = torch.full_like(scale, 1.0 - pfpn.BN_EPS)
            else:
                out[k] = _RoundWeightSTE.apply(v)
        elif param == "bias":
            out[k] = v
    return out


def is_trainable(key):
    """key = '<layer>/<param>'."""
    layer, param = key.split("/")
    if param not in ("kernel", "bias"):
        return False  # frozen BatchNorm
    return not (layer == "conv1" or layer.startswith("res2"))


def forward_losses(images_nhwc, weights_t, targets, num_classes, emulate_bf16=False, pinned=None):
    x = _dev(torch.from_numpy(np.ascontiguousarray(images_nhwc, dtype=np.float32))).permute(0, 3, 1, 2).contiguous()
    net_w = weights_t
    saved = list(pfpn._Q)
    if emulate_bf16 or pinned is not None:
        net_w = _emulated_weights(weights_t)
        x = x.to(torch.bfloat16).to(torch.float32)  # the stem reads the image as bf16 (launch_stem_s2d)
        pfpn._Q[0] = lambda t, name: _RoundBF16.apply(t)
    if pinned is not None:
        def pin(name):
            return _dev(torch.from_numpy(np.ascontiguousarray(pinned[name]))).permute(0, 3, 1, 2).contiguous()
        relu_names = set(n for n in pinned if n.endswith("_relu") or "_tower_t" in n)
        # ReLU points go to _PinRelu (pfpn._qrelu hands over the PRE-activation when _Q[1] is None), the rest to
        # _PinLinear
        pfpn._Q[0] = lambda t, name: (_PinRelu if name in relu_names else _PinLinear).apply(t, pin(name))
        pfpn._Q[1] = None
    try:
        _, C3, C4, C5 = pfpn.resnet50(x, net_w)
        feats = pfpn.sparce_fpn(C3, C4, C5, net_w)
        reg = torch.cat([pfpn.regression_head(f, net_w, i) for i, f in enumerate(feats)], dim=1)
        cls = torch.cat([pfpn.classification_head(f, net_w, num_classes, i) for i, f in enumerate(feats)], dim=1)
        msk = pfpn.mask_head(feats[0], net_w, num_classes, 0)
    finally:
        pfpn._Q[0], pfpn._Q[1] = saved
    y_reg, y_cls, y_mask = [_dev(torch.from_numpy(np.ascontiguousarray(t, dtype=np.float32))) for t in targets]
    l_reg = olosses.orthogonal_l1(y_reg, reg)
    l_cls = olosses.focal(y_cls, cls)
    l_mask = olosses.focal(y_mask, msk)
    l2 = sum((weights_t[k] ** 2).sum() for k in weights_t
             if k.startswith("pyramid_regression3D") and k.endswith("/kernel")) * L2_REG
    return l_reg, l_cls, l_mask, l2, (reg, cls, msk)


def step(images_nhwc, weights, targets, num_classes, lr=1e-5, clipnorm=0.001, beta_1=0.9, beta_2=0.999, eps=1e-7,
         num_threads=None, emulate_bf16=False, pinned=None):
    """One Adam step from zero moments. Returns dict with losses, raw gradients {key: HWIO/bias numpy}, the global
    gradient norm and the updated weights {key: numpy}.
    emulate_bf16=True rounds weights (after the BatchNorm fold), stored activations and gradient maps to bf16 at the
    points where the native engine does; everything else stays fp32.
    pinned={engine tensor name: NHWC float32} replaces every stored activation by the engine's own copy (and uses its
    ReLU masks in the backward pass), so that the comparison isolates the backward arithmetic and the wiring of the
    gradient graph from the ~1e-2 forward noise of bf16: see tests/test_train_gpu.py."""
    if num_threads:
        torch.set_num_threads(num_threads)
    wt = {}
    for k, v in weights.items():
        t = _dev(torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32)).clone())
        if is_trainable(k):
            t.requires_grad_(True)
        wt[k] = t
    l_reg, l_cls, l_mask, l2, outs = forward_losses(images_nhwc, wt, targets, num_classes, emulate_bf16, pinned)
    total = l_reg + l_cls + l_mask + l2
    total.backward()
    keys = [k for k in wt if wt[k].requires_grad]
    grads = {k: (wt[k].grad if wt[k].grad is not None else torch.zeros_like(wt[k])) for k in keys}
    norm = torch.sqrt(sum((g.double() ** 2).sum() for g in grads.values())).item()
    scale = clipnorm / norm if (clipnorm > 0 and norm >= clipnorm) else 1.0
    lr_t = lr * np.sqrt(1.0 - beta_2) / (1.0 - beta_1)
    new_w = {}
    for k in keys:
        g = grads[k] * scale
        m = (1.0 - beta_1) * g
        v = (1.0 - beta_2) * g * g
        new_w[k] = (wt[k].detach() - lr_t * m / (torch.sqrt(v) + eps)).cpu().numpy()
    return {
        "losses": {"3Dbox": l_reg.item(), "cls": l_cls.item(), "mask": l_mask.item(), "l2": l2.item(),
                   "total": total.item()},
        "grads": {k: g.cpu().numpy() for k, g in grads.items()},
        "grad_norm": norm,
        "weights": new_w,
        "outputs": [o.detach().cpu().numpy() for o in outs],
    }
