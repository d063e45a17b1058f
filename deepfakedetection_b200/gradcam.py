"""pytorch-grad-cam compatible front (``GradCAM``, ``ClassifierOutputTarget``, ``show_cam_on_image``).

The reference drives Grad-CAM as (/root/reference/web_ui.py:275-282)::

    with GradCAM(model=model, target_layers=[target_layer]) as cam:
        grayscale = cam(input_tensor=batch, targets=[ClassifierOutputTarget(cls_idx)])[0]
    overlay = show_cam_on_image(rgb, grayscale, use_rgb=True)

For a B200 engine module whose target layer is its ``_conv_head`` this class calls the fused,
batched, device-resident kernel (``b200df_forward_cam``): closed-form channel weights, weighted
sum, ReLU, min-max, bilinear upsample, min-max -- one pass, no autograd graph, no host numpy.
``use_hooks=True`` (or any other module / layer) follows the package's own mechanism instead:
forward hook on the target layer, ``register_hook`` on its output, backward from the summed
target logits -- which, on the B200 module, runs the trunk/head custom ops and the
``b200df_head_backward`` kernel.
"""

from __future__ import annotations

import numpy as np
import torch
from torch.nn import functional as F


class ClassifierOutputTarget:
    def __init__(self, category: int) -> None:
        self.category = int(category)

    def __call__(self, model_output: torch.Tensor) -> torch.Tensor:
        if model_output.ndim == 1:
            return model_output[self.category]
        return model_output[:, self.category]


def _scale(cam: torch.Tensor) -> torch.Tensor:
    flat = cam.flatten(1)
    cam = cam - flat.min(1).values.view(-1, 1, 1)
    return cam / (cam.flatten(1).max(1).values.view(-1, 1, 1) + 1e-7)


class GradCAM:
    def __init__(self, model, target_layers, reshape_transform=None, use_hooks: bool | None = None) -> None:
        self.model = model.eval()
        self.target_layers = list(target_layers)
        self.reshape_transform = reshape_transform
        if use_hooks is None:   # B200DF_GRADCAM_HOOKS=1: pytorch-grad-cam's own mechanism (hooks + backward) instead of the fused call
            import os

            use_hooks = os.environ.get("B200DF_GRADCAM_HOOKS", "0") not in ("", "0")
        self.use_hooks = use_hooks
        self.outputs: torch.Tensor | None = None

    def __enter__(self):
        return self

    def __exit__(self, *exc) -> bool:
        return False

    def _fused_ok(self) -> bool:
        from .module import B200Module

        return (
            not self.use_hooks
            and isinstance(self.model, B200Module)
            and self.model.supports_cam
            and len(self.target_layers) == 1
            and self.target_layers[0] is self.model._cam_point()
            and self.reshape_transform is None
        )

    def __call__(self, input_tensor: torch.Tensor, targets=None, aug_smooth: bool = False, eigen_smooth: bool = False) -> np.ndarray:
        if aug_smooth or eigen_smooth:
            raise NotImplementedError("aug_smooth / eigen_smooth are not used by the reference (web_ui.py:275-279)")
        dev = next(self.model.parameters()).device
        x = input_tensor.to(dev)
        if self._fused_ok():
            tgt = None
            if targets is not None:
                if not all(isinstance(t, ClassifierOutputTarget) for t in targets):
                    raise TypeError("the fused path supports ClassifierOutputTarget only; pass use_hooks=True")
                if len(targets) != x.shape[0]:
                    raise ValueError("need one target per image")
                tgt = torch.tensor([t.category for t in targets], dtype=torch.int32, device=dev)
            logits, _low, cam = self.model.forward_cam(x, tgt)
            self.outputs = logits
            return cam.cpu().numpy()
        return self._hooked(x, targets)

    # pytorch-grad-cam's mechanism, restated
    def _hooked(self, x: torch.Tensor, targets) -> np.ndarray:
        acts: list[torch.Tensor] = []
        grads: list[torch.Tensor] = []

        def fwd_hook(_m, _inp, out):
            acts.append(out)
            if out.requires_grad:
                out.register_hook(lambda g: grads.insert(0, g))

        handles = [layer.register_forward_hook(fwd_hook) for layer in self.target_layers]
        try:
            with torch.enable_grad():
                x = x.detach().requires_grad_(True)
                out = self.model(x)
                self.outputs = out.detach()
                if targets is None:
                    targets = [ClassifierOutputTarget(int(c)) for c in out.argmax(1)]
                loss = sum(t(o) for t, o in zip(targets, out))
                self.model.zero_grad(set_to_none=True)
                loss.backward()
        finally:
            for h in handles:
                h.remove()
        if len(grads) != len(acts):
            raise RuntimeError("target layer output did not receive a gradient")
        per_layer = []
        for a, g in zip(acts, grads):
            a, g = a.detach().float(), g.detach().float()
            w = g.mean((2, 3), keepdim=True)
            cam = (w * a).sum(1).clamp_min(0)
            cam = _scale(cam)
            cam = F.interpolate(cam[:, None], size=tuple(x.shape[-2:]), mode="bilinear", align_corners=False)[:, 0]
            per_layer.append(cam)
        cam = torch.stack(per_layer, 1).clamp_min(0).mean(1)
        return _scale(cam).cpu().numpy().astype(np.float32)


def _jet_lut() -> np.ndarray:
    try:
        import cv2

        lut = cv2.applyColorMap(np.arange(256, dtype=np.uint8).reshape(-1, 1), cv2.COLORMAP_JET)
        return lut.reshape(256, 3)  # BGR
    except Exception:  # pragma: no cover - cv2 is present in this image
        v = np.linspace(0, 1, 256)
        r = np.clip(1.5 - np.abs(4 * v - 3), 0, 1)
        g = np.clip(1.5 - np.abs(4 * v - 2), 0, 1)
        b = np.clip(1.5 - np.abs(4 * v - 1), 0, 1)
        return (np.stack([b, g, r], 1) * 255).astype(np.uint8)


def show_cam_on_image(img: np.ndarray, mask: np.ndarray, use_rgb: bool = False, colormap=None, image_weight: float = 0.5) -> np.ndarray:
    """JET heat-map blended over a float RGB image in [0,1] -> uint8 (web_ui.py:282)."""
    if np.max(img) > 1:
        raise ValueError("The input image should np.float32 in the range [0, 1]")
    if not 0 <= image_weight <= 1:
        raise ValueError("image_weight should be in the range [0, 1]")
    heat = _jet_lut()[np.uint8(255 * np.clip(mask, 0, 1))]
    if use_rgb:
        heat = heat[..., ::-1]
    heat = np.float32(heat) / 255
    cam = (1 - image_weight) * heat + image_weight * img
    cam = cam / np.max(cam)
    return np.uint8(255 * cam)


def show_cam_on_image_device(images: torch.Tensor, cams: torch.Tensor, normalize: bool = True,
                             mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225),
                             image_weight: float = 0.5) -> torch.Tensor:
    """``_tensor_to_rgb(tensor, normalize=...)`` + ``show_cam_on_image(rgb, cam, use_rgb=True)`` for a whole batch
    on the GPU (/root/reference/web_ui.py:72-93, 281-282): ``images`` float32 [N,3,H,W] (what the model saw),
    ``cams`` float32 [N,H,W] in [0,1] -> uint8 [N,H,W,3] RGB overlays, byte-identical to the host functions."""
    import ctypes as C

    from . import _lib

    if images.dim() != 4 or images.shape[1] != 3 or cams.dim() != 3 or cams.shape[0] != images.shape[0] \
            or tuple(cams.shape[1:]) != tuple(images.shape[2:]):
        raise ValueError("expected images [N,3,H,W] and cams [N,H,W]")
    if not images.is_cuda or not cams.is_cuda:
        raise RuntimeError("show_cam_on_image_device needs CUDA tensors (the host path is show_cam_on_image)")
    if not 0 <= image_weight <= 1:
        raise ValueError("image_weight should be in the range [0, 1]")
    x = images.detach().float().contiguous()
    m = cams.detach().float().contiguous()
    n, _, h, w = x.shape
    out = torch.empty((n, h, w, 3), dtype=torch.uint8, device=x.device)
    mean_c = (C.c_float * 3)(*[float(v) for v in mean])
    std_c = (C.c_float * 3)(*[float(v) for v in std])
    with torch.cuda.device(x.device):
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        _lib.check(_lib.lib().b200df_cam_overlay(C.c_void_p(x.data_ptr()), C.c_void_p(m.data_ptr()), n, h, w,
                                                  1 if normalize else 0, C.cast(mean_c, C.c_void_p),
                                                  C.cast(std_c, C.c_void_p), C.c_float(image_weight),
                                                  C.c_void_p(out.data_ptr()), stream))
    return out
