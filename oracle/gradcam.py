"""Grad-CAM oracle (TEST INFRASTRUCTURE ONLY).

Restates the algorithm the reference runs at /root/reference/web_ui.py:275-279
(``GradCAM(model, target_layers=[t])(input_tensor, targets=[ClassifierOutputTarget(c)])``)
with the target layer chosen by /root/reference/web_ui.py:96-114 (``_conv_head``
or the last ``nn.Conv2d``; for timm EfficientNet-B0 that is ``conv_head``).
The implementation lives in the third-party ``grad-cam`` package pinned at git
781dbc0d (/root/reference/requirements.txt:21), which is not on disk, so its
published algorithm is restated:

    A = target-layer output, G = d(sum_n logits[n, c_n]) / dA
    w[n,k] = mean_hw G[n,k];  cam[n] = relu(sum_k w[n,k] A[n,k])
    per image: cam = (cam - min) / (max + 1e-7); bilinear resize to the input HxW
    (cv2.resize semantics: half-pixel centres, edge clamp); relu; min-max again.

Two independent routes are provided and tests require that they agree:
``gradcam_autograd`` (hooks + backward, exactly the package's mechanism) and
``gradcam_closed_form`` (analytic gradient through BN -> SiLU -> mean -> Linear).
Parity vs the real package is *unpinned* (not installable here).
"""

from __future__ import annotations

import torch
from torch.nn import functional as F


def _minmax(cam: torch.Tensor) -> torch.Tensor:
    flat = cam.flatten(1)
    lo = flat.min(1).values.view(-1, 1, 1)
    cam = cam - lo
    hi = cam.flatten(1).max(1).values.view(-1, 1, 1)
    return cam / (hi + 1e-7)


def postprocess_cam(cam_lowres: torch.Tensor, out_hw: tuple[int, int]) -> torch.Tensor:
    """relu'd low-res cam [N,h,w] -> float32 [N,H,W] in [0,1] (scale -> resize -> relu -> scale)."""
    cam = _minmax(cam_lowres)
    cam = F.interpolate(cam[:, None], size=out_hw, mode="bilinear", align_corners=False)[:, 0]
    cam = cam.clamp_min(0)
    return _minmax(cam)


def gradcam_on_layer(model, target: torch.nn.Module, images: torch.Tensor, targets: torch.Tensor):
    """Hook-and-backward Grad-CAM on an explicit target layer (pytorch-grad-cam's mechanism): the route for the networks
    whose target -- the last ``nn.Conv2d`` in ``modules()`` order, /root/reference/web_ui.py:94-112 -- is not the conv head.
    Returns (logits [N,C], cam_lowres [N,h,w] after ReLU, cam [N,H,W] in [0,1], A [N,K,h,w], G [N,K,h,w])."""
    store = {}

    def hook(_m, _inp, out):
        store["A"] = out
        out.register_hook(lambda g: store.__setitem__("G", g))

    handle = target.register_forward_hook(hook)
    try:
        with torch.enable_grad():
            model.zero_grad(set_to_none=True)
            logits = model(images)
            logits[torch.arange(images.shape[0]), targets].sum().backward()
    finally:
        handle.remove()
    a, g = store["A"].detach(), store["G"].detach()
    cam_low = (g.mean((2, 3), keepdim=True) * a).sum(1).clamp_min(0)
    return logits.detach(), cam_low, postprocess_cam(cam_low, tuple(images.shape[-2:])), a, g


def gradcam_autograd(model, images: torch.Tensor, targets: torch.Tensor):
    """Hook-and-backward Grad-CAM on the layer the reference's target rule picks (/root/reference/web_ui.py:108-110):
    ``model._conv_head`` when it exists (efficientnet_pytorch layout), else ``model.conv_head`` (timm layout).

    Returns (logits [N,C], cam_lowres [N,h,w] after ReLU, cam [N,H,W] in [0,1]).
    """
    store = {}

    def hook(_m, _inp, out):
        store["A"] = out
        out.register_hook(lambda g: store.__setitem__("G", g))

    target = getattr(model, "_conv_head", None)
    if target is None:
        target = model.conv_head
    handle = target.register_forward_hook(hook)
    try:
        with torch.enable_grad():
            model.zero_grad(set_to_none=True)
            logits = model(images)
            loss = logits[torch.arange(images.shape[0]), targets].sum()
            loss.backward()
    finally:
        handle.remove()
    a, g = store["A"].detach(), store["G"].detach()
    w = g.mean((2, 3), keepdim=True)
    cam_low = (w * a).sum(1).clamp_min(0)
    return logits.detach(), cam_low, postprocess_cam(cam_low, tuple(images.shape[-2:]))


def gradcam_closed_form(model, images: torch.Tensor, targets: torch.Tensor):
    """Analytic route: everything after A is BN(eval) -> SiLU -> mean_hw -> Linear, so
    G[n,k,h,w] = W_fc[c_n,k] / (h*w) * s_k * silu'(s_k A + t_k)."""
    with torch.no_grad():
        a = model.forward_features_raw(images)
        bn = model.bn2
        s = bn.weight / torch.sqrt(bn.running_var + bn.eps)
        t = bn.bias - bn.running_mean * s
        z = a * s.view(1, -1, 1, 1) + t.view(1, -1, 1, 1)
        sig = torch.sigmoid(z)
        dsilu = sig * (1 + z * (1 - sig))
        hw = a.shape[2] * a.shape[3]
        wfc = model.classifier.weight[targets]  # [N,K]
        g = wfc.view(*wfc.shape, 1, 1) / hw * s.view(1, -1, 1, 1) * dsilu
        w = g.mean((2, 3), keepdim=True)
        cam_low = (w * a).sum(1).clamp_min(0)
        logits = model.forward_head(a)
    return logits, cam_low, postprocess_cam(cam_low, tuple(images.shape[-2:]))
