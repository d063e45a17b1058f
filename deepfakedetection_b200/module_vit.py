"""``nn.Module`` fronts for the two transformer-style backbones the reference registry serves through its prefix table
(/root/reference/orchestration/model_registry.py:39-47, 60-75):

* ``B200EfficientFormerV2S1`` -- drop-in for ``timm.create_model("efficientformerv2_s1", pretrained=False,
  num_classes=...)`` (model_registry.py:39-40; ``config/inference.yaml:25-37``);
* ``B200FasterViT`` -- drop-in for ``fastervit.create_model("faster_vit_{0,2}_224", pretrained=False)`` with the fresh
  ``head`` the registry attaches (model_registry.py:43-47; ``config/inference.yaml:38-50``).

As in ``module.py`` the sub-modules are parameter CONTAINERS that carry the third-party packages' parameter / buffer
names and shapes so that the reference's ``load_state_dict(state, strict=False)``
(orchestrator.py:370-375) fills them; no arithmetic ever runs in PyTorch.  Module registration ORDER follows the
packages too, because the reference picks its Grad-CAM target as "the last nn.Conv2d in modules() order"
(/root/reference/web_ui.py:96-105): that is ``stages.3.blocks.5.mlp.fc2.conv`` for EfficientFormerV2-S1 (served: the
map after it is affine, closed form) and ``levels.2.global_tokenizer.pos_embed`` for FasterViT (served by the engine's
backward pass through the transformer blocks behind it, csrc/cam_bwd.cu).
"""

from __future__ import annotations

import math

import torch
from torch import nn

from . import _lib
from .module import _BN, B200Module, CamHookPoint, _Conv, _List, _NoForward, _Seq


class _Linear(nn.Linear):
    def forward(self, x):  # noqa: D102
        return _NoForward.forward(self)


class _LN(nn.LayerNorm):
    def forward(self, x):  # noqa: D102
        return _NoForward.forward(self)


# ---- EfficientFormerV2 (timm models/efficientformer_v2.py layout) -----------------------------------------------------

class _ConvNorm(_NoForward):
    """timm ``ConvNorm`` / ``ConvNormAct``: ``conv`` (bias=True) + ``bn``."""

    def __init__(self, cin: int, cout: int, k: int = 1, stride: int = 1, groups: int = 1, hook: bool = False) -> None:
        super().__init__()
        cls = CamHookPoint if hook else _Conv
        self.conv = cls(cin, cout, k, stride, (k - 1) // 2, groups=groups, bias=True)
        self.bn = _BN(cout)


class _Attention2d(_NoForward):
    def __init__(self, dim: int, resolution: int, stride: int | None, key_dim: int = 32, heads: int = 8, ratio: int = 4) -> None:
        super().__init__()
        if stride is not None:
            resolution = math.ceil(resolution / stride)
            self.stride_conv = _ConvNorm(dim, dim, 3, stride, groups=dim)
        else:
            self.stride_conv = None
        dh = ratio * key_dim * heads
        self.q = _ConvNorm(dim, key_dim * heads)
        self.k = _ConvNorm(dim, key_dim * heads)
        self.v = _ConvNorm(dim, dh)
        self.v_local = _ConvNorm(dh, dh, 3, groups=dh)
        self.talking_head1 = _Conv(heads, heads, 1)
        self.talking_head2 = _Conv(heads, heads, 1)
        self.proj = _ConvNorm(dh, dim, 1)
        self.attention_biases = nn.Parameter(torch.zeros(heads, resolution * resolution))


class _LocalGlobalQuery(_NoForward):
    def __init__(self, cin: int, cout: int) -> None:
        super().__init__()
        self.local = _Conv(cin, cin, 3, 2, 1, groups=cin)
        self.proj = _ConvNorm(cin, cout, 1)


class _Attention2dDownsample(_NoForward):
    def __init__(self, dim: int, out_dim: int, resolution: int, key_dim: int = 16, heads: int = 8, ratio: int = 4) -> None:
        super().__init__()
        dh = ratio * key_dim * heads
        self.q = _LocalGlobalQuery(dim, key_dim * heads)
        self.k = _ConvNorm(dim, key_dim * heads, 1)
        self.v = _ConvNorm(dim, dh, 1)
        self.v_local = _ConvNorm(dh, dh, 3, 2, groups=dh)
        self.proj = _ConvNorm(dh, out_dim, 1)
        self.attention_biases = nn.Parameter(torch.zeros(heads, resolution * resolution))


class _Downsample(_NoForward):
    def __init__(self, cin: int, cout: int, resolution: int, use_attn: bool) -> None:
        super().__init__()
        self.conv = _ConvNorm(cin, cout, 3, 2)
        self.attn = _Attention2dDownsample(cin, cout, resolution) if use_attn else None


class _ConvMlp(_NoForward):
    def __init__(self, dim: int, hidden: int, hook: bool) -> None:
        super().__init__()
        self.fc1 = _ConvNorm(dim, hidden, 1)
        self.mid = _ConvNorm(hidden, hidden, 3, groups=hidden)
        self.fc2 = _ConvNorm(hidden, dim, 1, hook=hook)


class _LayerScale(_NoForward):
    def __init__(self, dim: int, init: float = 1e-5) -> None:
        super().__init__()
        self.gamma = nn.Parameter(init * torch.ones(dim))


class _EfBlock(_NoForward):
    def __init__(self, dim: int, ratio: int, resolution: int, stride: int | None, use_attn: bool, hook: bool) -> None:
        super().__init__()
        if use_attn:
            self.token_mixer = _Attention2d(dim, resolution, stride)
            self.ls1 = _LayerScale(dim)
        else:
            self.token_mixer = None
            self.ls1 = None
        self.mlp = _ConvMlp(dim, dim * ratio, hook)
        self.ls2 = _LayerScale(dim)


class _EfStage(_NoForward):
    def __init__(self, dim: int, dim_out: int, depth: int, resolution: int, downsample: bool, block_stride: int | None,
                 ds_attn: bool, block_attn: bool, num_vit: int, ratios, last: bool) -> None:
        super().__init__()
        if downsample:
            self.downsample = _Downsample(dim, dim_out, resolution, ds_attn)
            dim = dim_out
            resolution = math.ceil(resolution / 2)
        else:
            self.downsample = nn.Identity()
        remain = depth - num_vit - 1
        self.blocks = _Seq(*[
            _EfBlock(dim, ratios[i], resolution, block_stride, block_attn and i > remain, hook=last and i == depth - 1)
            for i in range(depth)
        ])


class _Stem4(_Seq):
    def __init__(self, cin: int, cout: int) -> None:
        super().__init__()
        self.conv1 = _ConvNorm(cin, cout // 2, 3, 2)
        self.conv2 = _ConvNorm(cout // 2, cout, 3, 2)


_EFV2_S1 = dict(dims=(32, 48, 120, 224), depths=(3, 3, 9, 6),
                ratios=((4, 4, 4), (4, 4, 4), (4, 4, 3, 3, 3, 3, 4, 4, 4), (4, 4, 3, 3, 4, 4)), num_vit=2)


class B200EfficientFormerV2S1(B200Module):
    """Drop-in for timm ``efficientformerv2_s1`` (eval mode: logits = (head(x) + head_dist(x)) / 2)."""

    arch_id = _lib.ARCH_EFFICIENTFORMERV2_S1
    arch_name = "efficientformerv2_s1"

    def _build_containers(self, num_classes: int) -> None:
        cfg = _EFV2_S1
        dims, depths = cfg["dims"], cfg["depths"]
        self.stem = _Stem4(3, dims[0])
        prev, stride, stages = dims[0], 4, []
        for i in range(4):
            stages.append(_EfStage(prev, dims[i], depths[i], math.ceil(224 / stride), i > 0, 2 if i == 2 else None, i >= 3,
                                   i >= 2, cfg["num_vit"], cfg["ratios"][i], last=i == 3))
            if i > 0:
                stride *= 2
            prev = dims[i]
        self.stages = _Seq(*stages)
        self.norm = _BN(dims[-1])
        self.head = _Linear(dims[-1], num_classes)
        self.head_dist = _Linear(dims[-1], num_classes)

    def _init_weights(self) -> None:  # timm init_weights: Linear trunc_normal(.02) / zero bias; convs keep torch defaults
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=0.02)
                nn.init.zeros_(m.bias)

    def _cam_point(self) -> nn.Module:
        return self.stages[3].blocks[-1].mlp.fc2.conv


# ---- FasterViT (fastervit/models/faster_vit.py layout) -------------------------------------------------------------------

class _PosEmb1D(_NoForward):
    def __init__(self, dim: int, seq_length: int) -> None:
        super().__init__()
        self.cpb_mlp = _Seq(_Linear(2, 512, bias=True), nn.ReLU(), _Linear(512, dim, bias=False))
        self.register_buffer("relative_bias", torch.zeros(1, seq_length, dim))


class _PosEmb2D(_NoForward):
    def __init__(self, window: int, heads: int, seq_length: int) -> None:
        super().__init__()
        self.cpb_mlp = _Seq(_Linear(2, 512, bias=True), nn.ReLU(inplace=True), _Linear(512, heads, bias=False))
        rel = torch.arange(-(window - 1), window, dtype=torch.float32)
        table = torch.stack(torch.meshgrid([rel, rel], indexing="ij")).permute(1, 2, 0).contiguous().unsqueeze(0)
        table = table / max(window - 1, 1) * 8
        table = torch.sign(table) * torch.log2(torch.abs(table) + 1.0) / math.log2(8)
        self.register_buffer("relative_coords_table", table)
        co = torch.arange(window)
        coords = torch.flatten(torch.stack(torch.meshgrid([co, co], indexing="ij")), 1)
        rc = (coords[:, :, None] - coords[:, None, :]).permute(1, 2, 0).contiguous()
        rc[:, :, 0] += window - 1
        rc[:, :, 1] += window - 1
        rc[:, :, 0] *= 2 * window - 1
        self.register_buffer("relative_position_index", rc.sum(-1))
        self.register_buffer("relative_bias", torch.zeros(1, heads, seq_length, seq_length))


class _WindowAttention(_NoForward):
    def __init__(self, dim: int, heads: int, resolution: int, seq_length: int) -> None:
        super().__init__()
        self.qkv = _Linear(dim, dim * 3, bias=True)
        self.proj = _Linear(dim, dim)
        self.pos_emb_funct = _PosEmb2D(resolution, heads, seq_length)


class _FvMlp(_NoForward):
    def __init__(self, dim: int, hidden: int) -> None:
        super().__init__()
        self.fc1 = _Linear(dim, hidden)
        self.fc2 = _Linear(hidden, dim)


class _HAT(_NoForward):
    def __init__(self, dim: int, heads: int, sr_ratio: int, window: int = 7, ct_size: int = 2) -> None:
        super().__init__()
        self.pos_embed = _PosEmb1D(dim, window ** 2)
        self.norm1 = _LN(dim)
        ct_per_window = ct_size ** 2 if sr_ratio > 1 else 0
        ct_total = ct_per_window * sr_ratio * sr_ratio
        self.attn = _WindowAttention(dim, heads, window, window ** 2 + ct_per_window)
        self.norm2 = _LN(dim)
        self.mlp = _FvMlp(dim, dim * 4)
        if sr_ratio > 1:
            self.hat_norm1 = _LN(dim)
            self.hat_norm2 = _LN(dim)
            self.hat_attn = _WindowAttention(dim, heads, int(ct_total ** 0.5), ct_total)
            self.hat_mlp = _FvMlp(dim, dim * 4)
            self.hat_pos_embed = _PosEmb1D(dim, ct_total)


class _ConvBlock(_NoForward):
    def __init__(self, dim: int) -> None:
        super().__init__()
        self.conv1 = _Conv(dim, dim, 3, 1, 1)
        self.norm1 = _BN(dim, eps=1e-5)
        self.conv2 = _Conv(dim, dim, 3, 1, 1)
        self.norm2 = _BN(dim, eps=1e-5)


class _FvDownsample(_NoForward):
    def __init__(self, dim: int) -> None:
        super().__init__()
        self.norm = _LN(dim, eps=1e-6)
        self.reduction = _Seq(_Conv(dim, 2 * dim, 3, 2, 1, bias=False))


class _TokenInitializer(_NoForward):
    def __init__(self, dim: int, input_resolution: int, window: int = 7, ct_size: int = 2) -> None:
        super().__init__()
        output_size = int(ct_size * input_resolution / window)
        stride = int(input_resolution / output_size)
        kernel = input_resolution - (output_size - 1) * stride
        self.pos_embed = CamHookPoint(dim, dim, 3, padding=1, groups=dim)
        self.to_global_feature = _Seq()
        self.to_global_feature.add_module("pos", self.pos_embed)
        self.to_global_feature.add_module("pool", nn.AvgPool2d(kernel_size=kernel, stride=stride))


class _FvLevel(_NoForward):
    def __init__(self, dim: int, depth: int, input_resolution: int, heads: int, conv: bool, downsample: bool,
                 only_local: bool) -> None:
        super().__init__()
        if conv:
            self.blocks = _List([_ConvBlock(dim) for _ in range(depth)])
        else:
            sr = input_resolution // 7 if not only_local else 1
            self.blocks = _List([_HAT(dim, heads, sr) for _ in range(depth)])
        self.downsample = _FvDownsample(dim) if downsample else None   # registered BEFORE global_tokenizer, as in the package
        if (not conv) and (not only_local) and input_resolution // 7 > 1:
            self.global_tokenizer = _TokenInitializer(dim, input_resolution)


class _PatchEmbed(_NoForward):
    def __init__(self, in_dim: int, dim: int) -> None:
        super().__init__()
        self.conv_down = _Seq(_Conv(3, in_dim, 3, 2, 1, bias=False), _BN(in_dim, eps=1e-4), nn.ReLU(),
                              _Conv(in_dim, dim, 3, 2, 1, bias=False), _BN(dim, eps=1e-4), nn.ReLU())


_FV_CFG = {
    "faster_vit_0_224": dict(arch=_lib.ARCH_FASTERVIT_0, dim=64, in_dim=64, depths=(2, 3, 6, 5), heads=(2, 4, 8, 16)),
    "faster_vit_2_224": dict(arch=_lib.ARCH_FASTERVIT_2, dim=96, in_dim=64, depths=(3, 3, 8, 5), heads=(2, 4, 8, 16)),
}


class B200FasterViT(B200Module):
    """Drop-in for ``fastervit.create_model("faster_vit_0_224" | "faster_vit_2_224")`` + ``head = nn.Linear(...)``.

    Grad-CAM: the reference's target rule (web_ui.py:94-112) picks ``levels.2.global_tokenizer.pos_embed``; the engine's
    backward program runs through every block behind it (csrc/cam_bwd.cu), in the fused ``forward_cam`` and -- for the
    real pytorch-grad-cam package -- behind the hook protocol: a forward with grad enabled returns logits whose autograd
    edge to the (14 x 14) target activation is that backward program."""

    supports_cam = True
    supports_cam_hooks = True

    def __init__(self, num_classes: int = 2, precision: str = "bf16", variant: str = "faster_vit_0_224") -> None:
        if variant not in _FV_CFG:
            raise KeyError(f"b200 backend serves {sorted(_FV_CFG)}, not '{variant}'")
        self.variant = variant
        self.arch_id = _FV_CFG[variant]["arch"]
        self.arch_name = variant
        super().__init__(num_classes=num_classes, precision=precision)

    def _build_containers(self, num_classes: int) -> None:
        cfg = _FV_CFG[self.variant]
        dim, depths, heads = cfg["dim"], cfg["depths"], cfg["heads"]
        self.patch_embed = _PatchEmbed(cfg["in_dim"], dim)
        hat = (False, False, True, False)
        self.levels = _List([
            _FvLevel(dim * 2 ** i, depths[i], int(2 ** (-2 - i) * 224), heads[i], i < 2, i < 3, not hat[i]) for i in range(4)
        ])
        self.num_features = dim * 8
        self.norm = _BN(self.num_features)
        self.head = _Linear(self.num_features, num_classes)

    def _init_weights(self) -> None:  # fastervit _init_weights: Linear trunc_normal(.02) / zero bias, norms 1 / 0
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=0.02)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)

    def _cam_point(self) -> nn.Module:
        return self.levels[2].global_tokenizer.pos_embed
