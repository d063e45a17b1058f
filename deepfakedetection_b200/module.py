"""``nn.Module`` front of the B200 engine: what ``ModelSpec.builder`` returns.

Mirrors the lifecycle the reference drives (/root/reference/orchestration/orchestrator.py:350-377):
``builder(name, num_classes)`` -> ``.to(device)`` -> ``.eval()`` ->
``load_state_dict(state, strict=False)`` -> ``model(images)`` under ``inference_mode``
(orchestrator.py:587-592).  Parameters and buffers carry *timm's* ``efficientnet_b0`` names
and shapes so the caller's ``load_state_dict`` fills them; they are only
containers -- no forward math ever runs in PyTorch.  On first use (and after every
``load_state_dict`` / ``.to()``) the weights are pushed through the C ABI, BN-folded in fp32
and packed for the kernels.

Grad-CAM protocol (/root/reference/web_ui.py:96-114, 275-279): ``_conv_head`` is an
``nn.Module``; when autograd is enabled ``forward`` runs trunk -> ``_conv_head`` (hook point,
sees the raw conv_head output A) -> head, with the head's backward supplied by
``b200df_head_backward``.
"""

from __future__ import annotations

import math
import warnings

import torch
from torch import nn

from . import _lib

# (depthwise-separable?, kernel, stride, expand, out channels, repeats)
_B0_STAGES = (
    (True, 3, 1, 1, 16, 1),
    (False, 3, 2, 6, 24, 2),
    (False, 5, 2, 6, 40, 2),
    (False, 3, 2, 6, 80, 3),
    (False, 5, 1, 6, 112, 3),
    (False, 5, 2, 6, 192, 4),
    (False, 3, 1, 6, 320, 1),
)

PRECISIONS = {"bf16": _lib.PREC_BF16, "fp32": _lib.PREC_FP32}


class _NoForward(nn.Module):
    """Parameter container: the arithmetic lives in libb200df, never here."""

    def forward(self, *args, **kwargs):  # noqa: D102
        raise RuntimeError(
            f"{type(self).__name__} is a parameter container of the b200 backend; call the top-level model "
            "(there is no per-layer PyTorch/CPU path)."
        )


class _Conv(nn.Conv2d):
    def forward(self, x):  # noqa: D102
        return _NoForward.forward(self)


class _BN(nn.BatchNorm2d):
    def forward(self, x):  # noqa: D102
        return _NoForward.forward(self)


class _SE(_NoForward):
    def __init__(self, chs: int, rd: int) -> None:
        super().__init__()
        self.conv_reduce = _Conv(chs, rd, 1, bias=True)
        self.conv_expand = _Conv(rd, chs, 1, bias=True)


class _DSBlock(_NoForward):
    def __init__(self, cin: int, cout: int, k: int, stride: int) -> None:
        super().__init__()
        self.conv_dw = _Conv(cin, cin, k, stride, (k - 1) // 2, groups=cin, bias=False)
        self.bn1 = _BN(cin)
        self.se = _SE(cin, max(1, cin // 4))
        self.conv_pw = _Conv(cin, cout, 1, bias=False)
        self.bn2 = _BN(cout)


class _IRBlock(_NoForward):
    def __init__(self, cin: int, cout: int, k: int, stride: int, exp: int) -> None:
        super().__init__()
        mid = cin * exp
        self.conv_pw = _Conv(cin, mid, 1, bias=False)
        self.bn1 = _BN(mid)
        self.conv_dw = _Conv(mid, mid, k, stride, (k - 1) // 2, groups=mid, bias=False)
        self.bn2 = _BN(mid)
        self.se = _SE(mid, max(1, cin // 4))
        self.conv_pwl = _Conv(mid, cout, 1, bias=False)
        self.bn3 = _BN(cout)


class _Seq(nn.Sequential):
    def forward(self, x):  # noqa: D102
        return _NoForward.forward(self)


class CamHookPoint(nn.Conv2d):
    """``conv_head``: holds the 1x1 conv weight and acts as the Grad-CAM hook point.

    The engine computes the conv itself; ``forward`` receives the already computed raw output A
    and hands it on (making it a grad-requiring leaf when autograd is on), so forward hooks
    registered on this module -- what pytorch-grad-cam does with the layer returned by
    ``_resolve_cam_target`` (/root/reference/web_ui.py:108-114) -- observe A and its gradient.
    """

    def forward(self, a: torch.Tensor) -> torch.Tensor:  # noqa: D102
        if torch.is_grad_enabled() and not a.requires_grad:
            a = a.detach().requires_grad_(True)
        return a


class B200Module(nn.Module):
    """Lifecycle shared by every backbone served by the engine: timm/efficientnet_pytorch/fastervit-named parameter
    containers + lazy (re)packing through the C ABI + the hot ``forward`` (orchestrator.py:350-377, 587-592)."""

    arch_id = -1
    arch_name = ""
    default_image_size = 224
    supports_cam = True          # fused Grad-CAM (forward_cam / the gradcam.GradCAM front)
    supports_cam_hooks = True    # pytorch-grad-cam's hook protocol (forward with grad enabled: trunk -> A -> head)

    def __init__(self, num_classes: int = 2, precision: str = "bf16") -> None:
        super().__init__()
        if precision not in PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(PRECISIONS)}, got {precision!r}")
        self.num_classes = num_classes
        self.precision = precision
        self._build_containers(num_classes)
        self._init_weights()
        self._engine = None
        self._engine_id = 0
        self._dirty = True
        self.register_load_state_dict_post_hook(B200Module._post_load)

    def _build_containers(self, num_classes: int) -> None:
        raise NotImplementedError

    def _cam_point(self) -> nn.Module:
        raise NotImplementedError

    def _init_weights(self) -> None:
        pass

    # ---- weight lifecycle ----------------------------------------------------------------------------
    @staticmethod
    def _post_load(module: "B200Module", incompatible) -> None:
        module._dirty = True
        # the reference loads with strict=False, which hides mistakes; say what did not line up
        if incompatible.missing_keys:
            warnings.warn(f"b200 backend: checkpoint is missing {len(incompatible.missing_keys)} tensors "
                          f"(kept at their current values), e.g. {incompatible.missing_keys[:3]}", stacklevel=2)
        if incompatible.unexpected_keys:
            warnings.warn(f"b200 backend: checkpoint has {len(incompatible.unexpected_keys)} unexpected tensors "
                          f"(ignored), e.g. {incompatible.unexpected_keys[:3]}", stacklevel=2)

    def _apply(self, fn, *args, **kwargs):  # .to()/.cuda()/.float() move the containers -> repack
        self._dirty = True
        return super()._apply(fn, *args, **kwargs)

    def train(self, mode: bool = True):  # noqa: D102
        if mode:
            warnings.warn("b200 backend is inference-only: BatchNorm always uses running statistics", stacklevel=2)
        return super().train(mode)

    def refresh(self) -> None:
        """Re-fold and re-upload the weights (call after editing parameters in place)."""
        self._dirty = True

    def set_precision(self, precision: str) -> None:
        if precision not in PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(PRECISIONS)}")
        if precision != self.precision:
            self.precision = precision
            self._dirty = True

    def _ready(self, device: torch.device):
        from . import ops
        from .engine import Engine

        index = device.index if device.index is not None else torch.cuda.current_device()
        if self._engine is None or self._engine.device.index != index:
            self._engine = Engine(self.arch_id, self.num_classes, torch.device("cuda", index))
            self._engine_id = ops.register_engine(self._engine)
            self._dirty = True
        if self._dirty:
            self._engine.load_state(self.state_dict(), PRECISIONS[self.precision])
            self._dirty = False
        return self._engine

    # ---- the hot call -----------------------------------------------------------------------------------
    def _check_input(self, images: torch.Tensor) -> None:
        if images.ndim != 4 or images.shape[1] != 3 or images.shape[2] != 224 or images.shape[3] != 224:
            raise ValueError(f"b200 {self.arch_name} expects float32 [N,3,224,224] images, got {tuple(images.shape)}")
        if not images.is_cuda:
            raise RuntimeError(
                f"images are on '{images.device}': the b200 backend has no CPU fallback. The reference silently "
                "falls back to CPU when CUDA is unavailable (orchestrator.py:463-469); this backend refuses instead."
            )

    def forward(self, images: torch.Tensor) -> torch.Tensor:  # noqa: D102
        from . import ops

        self._check_input(images)
        if images.dtype != torch.float32:
            images = images.float()
        self._ready(images.device)
        if torch.is_grad_enabled():
            # Grad-CAM hook protocol (web_ui.py:275-279): pytorch-grad-cam hooks `_conv_head` and calls
            # `.register_hook` on its output, so A must be a graph leaf that requires grad.  The trunk has no
            # backward (inference engine): gradients stop at A, which is all Grad-CAM needs.
            if not self.supports_cam_hooks:
                raise NotImplementedError(
                    f"b200 {self.arch_name}: no autograd path to the Grad-CAM target of this backbone.  "
                    "Call the model under torch.inference_mode()/no_grad(), or use forward_cam().")
            a, aux = ops.trunk(images.detach(), self._engine_id)
            a = self._cam_point()(a.requires_grad_(True))
            return ops.head(a, aux, self._engine_id)
        return ops.forward(images, self._engine_id)

    def forward_u8(self, images_hwc: torch.Tensor, mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225)) -> torch.Tensor:
        """uint8 [N,224,224,3] crops -> logits, with ToTensor + Normalize fused into the stem
        (orchestrator.py:340-345).  Pass ``mean=std=None`` for ``val_normalize: false``."""
        if images_hwc.ndim != 4 or tuple(images_hwc.shape[1:]) != (224, 224, 3) or images_hwc.dtype != torch.uint8:
            raise ValueError(f"b200 {self.arch_name} forward_u8 expects uint8 [N,224,224,3] (HWC) crops, got "
                             f"{images_hwc.dtype} {tuple(images_hwc.shape)}")
        for name, v in (("mean", mean), ("std", std)):
            if v is not None and len(v) != 3:
                raise ValueError(f"{name} must have 3 entries, got {len(v)}")
        if (mean is None) != (std is None):
            raise ValueError("pass mean and std together (or both None for `val_normalize: false`)")
        if not images_hwc.is_cuda:
            raise RuntimeError(f"images are on '{images_hwc.device}': the b200 backend has no CPU fallback")
        eng = self._ready(images_hwc.device)
        return eng.forward_u8(images_hwc, mean, std)

    @torch.no_grad()
    def forward_cam(self, images: torch.Tensor, targets: torch.Tensor | None = None):
        """Batched fused Grad-CAM: (logits [N,C], cam_lowres [N,7,7], cam [N,224,224] in [0,1])."""
        from . import ops

        self._check_input(images)
        self._ready(images.device)
        if targets is not None:
            targets = targets.to(device=images.device, dtype=torch.int32)
        return ops.forward_cam(images.float(), targets, self._engine_id)


class B200EfficientNetB0(B200Module):
    """Drop-in for ``timm.create_model("efficientnet_b0", num_classes=...)`` in eval mode on a B200."""

    arch_id = _lib.ARCH_EFFICIENTNET_B0
    arch_name = "efficientnet_b0"

    def _build_containers(self, num_classes: int) -> None:
        self.conv_stem = _Conv(3, 32, 3, 2, 1, bias=False)
        self.bn1 = _BN(32)
        stages = []
        cin = 32
        for ds, k, s, e, cout, reps in _B0_STAGES:
            blocks = []
            for r in range(reps):
                stride = s if r == 0 else 1
                blocks.append(_DSBlock(cin, cout, k, stride) if ds else _IRBlock(cin, cout, k, stride, e))
                cin = cout
            stages.append(_Seq(*blocks))
        self.blocks = _Seq(*stages)
        self.conv_head = CamHookPoint(cin, 1280, 1, bias=False)
        self.bn2 = _BN(1280)
        self.classifier = nn.Linear(1280, num_classes)

    def _cam_point(self) -> nn.Module:
        return self.conv_head

    # timm's EfficientNet initialisation, restated: conv N(0, sqrt(2/fan_out)); BN 1/0; linear U(+-1/sqrt(out)).
    def _init_weights(self) -> None:
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                fan_out = m.kernel_size[0] * m.kernel_size[1] * m.out_channels // m.groups
                nn.init.normal_(m.weight, 0.0, math.sqrt(2.0 / fan_out))
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
            elif isinstance(m, nn.Linear):
                bound = 1.0 / math.sqrt(m.weight.size(0))
                nn.init.uniform_(m.weight, -bound, bound)
                nn.init.zeros_(m.bias)

    # ---- Grad-CAM target (web_ui.py:108-110 looks for `_conv_head` first) --------------------------
    @property
    def _conv_head(self) -> nn.Module:
        return self.conv_head


# ---- the reference's registry default: efficientnet_pytorch EfficientNet-B3 ---------------------------------------

# (kernel, stride, expand, out channels, repeats): the base table scaled by width 1.2 / depth 1.4
_B3_STAGES = ((3, 1, 1, 24, 2), (3, 2, 6, 32, 3), (5, 2, 6, 48, 3), (3, 2, 6, 96, 5), (5, 1, 6, 136, 5),
              (5, 2, 6, 232, 6), (3, 1, 6, 384, 2))


class _EPBlock(_NoForward):
    """efficientnet_pytorch ``MBConvBlock`` parameter names (no ``_expand_conv``/``_bn0`` at expand ratio 1)."""

    def __init__(self, cin: int, cout: int, k: int, stride: int, exp: int) -> None:
        super().__init__()
        mid = cin * exp
        if exp != 1:
            self._expand_conv = _Conv(cin, mid, 1, bias=False)
            self._bn0 = _BN(mid, eps=1e-3)
        self._depthwise_conv = _Conv(mid, mid, k, stride, 0, groups=mid, bias=False)
        self._bn1 = _BN(mid, eps=1e-3)
        rd = max(1, int(cin * 0.25))
        self._se_reduce = _Conv(mid, rd, 1, bias=True)
        self._se_expand = _Conv(rd, mid, 1, bias=True)
        self._project_conv = _Conv(mid, cout, 1, bias=False)
        self._bn2 = _BN(cout, eps=1e-3)


class _List(nn.ModuleList):
    def forward(self, x):  # noqa: D102
        return _NoForward.forward(self)


class B200EfficientNetB3(B200EfficientNetB0):
    """Drop-in for ``EfficientNet.from_name("efficientnet-b3")`` with a fresh ``_fc``
    (/root/reference/orchestration/model_registry.py:32-36): the registry's ``efficientnet_b3`` entry (:50-58), the
    model ``config/inference.yaml`` and the published release weights use.  Parameters carry efficientnet_pytorch's
    names (``_conv_stem``, ``_bn0``, ``_blocks.N.*``, ``_conv_head``, ``_bn1``, ``_fc``); the engine applies that
    package's static TF-"same" padding for its 300-pixel native size to the 224-pixel crops and BN eps 1e-3."""

    arch_id = _lib.ARCH_EFFICIENTNET_B3
    arch_name = "efficientnet_b3"

    def _build_containers(self, num_classes: int) -> None:
        self._conv_stem = _Conv(3, 40, 3, 2, 0, bias=False)
        self._bn0 = _BN(40, eps=1e-3)
        blocks = []
        cin = 40
        for k, s, e, cout, reps in _B3_STAGES:
            for r in range(reps):
                blocks.append(_EPBlock(cin, cout, k, s if r == 0 else 1, e))
                cin = cout
        self._blocks = _List(blocks)
        self._conv_head = CamHookPoint(cin, 1536, 1, bias=False)
        self._bn1 = _BN(1536, eps=1e-3)
        self._fc = nn.Linear(1536, num_classes)

    # B0 exposes `_conv_head` as an alias property; here it is the real submodule of that name
    @property
    def _conv_head(self) -> nn.Module:
        return self._modules["_conv_head"]

    def _cam_point(self) -> nn.Module:
        return self._modules["_conv_head"]
