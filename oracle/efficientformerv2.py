"""fp32 eager restatement of timm ``efficientformerv2_s1`` (TEST INFRASTRUCTURE ONLY).

CPU oracle for the hot call ``logits = model(images)``
(/root/reference/orchestration/orchestrator.py:529, :590) when the model is the one the reference registry
builds through ``timm.create_model(model_name, pretrained=False, num_classes=...)``
(/root/reference/orchestration/model_registry.py:39-40, prefix entry ``efficientformer`` :61-67; trainer call
/root/reference/trainers/efficientformer_v2.py:327; ``config/inference.yaml:25-37``).

timm 1.0.20 (/root/reference/requirements.txt:95) is a third-party dependency that is NOT on disk and cannot be
installed here, and the reference holds no golden vectors for this path, so the network is restated from the
published algorithm (timm ``models/efficientformer_v2.py``; Li et al., "Rethinking Vision Transformers for
MobileNet Size and Speed", 2022) in timm's state_dict layout:

    stem.conv{1,2}.{conv,bn}                                  dense 3x3 s2 (bias) + BN + GELU, 3 -> C0/2 -> C0
    stages.S.downsample.conv.{conv,bn}                        dense 3x3 s2 (bias) + BN            (S >= 1)
    stages.3.downsample.attn.{q.local,q.proj.{conv,bn},k.{conv,bn},v.{conv,bn},v_local.{conv,bn},
                              proj.{conv,bn},attention_biases}  Attention2dDownsample (LocalGlobalQuery)
    stages.S.blocks.B.mlp.{fc1,mid,fc2}.{conv,bn}             1x1 (bias)+BN+GELU -> dw3x3 (bias)+BN+GELU -> 1x1 (bias)+BN
    stages.S.blocks.B.ls2.gamma                               LayerScale on the MLP branch
    stages.S.blocks.B.token_mixer.{stride_conv.{conv,bn},q,k,v,v_local,talking_head1,talking_head2,proj,
                                   attention_biases}, ls1.gamma   Attention2d (last `num_vit` blocks of stages 2, 3)
    norm (BN), head, head_dist                                 logits = (head(x) + head_dist(x)) / 2 in eval mode

**PARITY UNPINNED**: there is no second implementation of this network on disk.  What IS pinned
(``tests/test_oracle_efv2.py``): parameter counts of all four published variants against the figures timm /
the paper publish (S0 3.60 M, S1 6.19 M, S2 12.71 M, L 26.32 M at 1000 classes), the S1 MAC count (~0.65 G),
the state_dict key list (``STATE_DICT_KEYS_S1_SHA``; a real timm install can diff it), and the Grad-CAM closed
form against autograd.  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this file.
"""

from __future__ import annotations

import hashlib
import math

import torch
from torch import nn
from torch.nn import functional as F

WIDTHS = {"L": (40, 80, 192, 384), "S2": (32, 64, 144, 288), "S1": (32, 48, 120, 224), "S0": (32, 48, 96, 176)}
DEPTHS = {"L": (5, 5, 15, 10), "S2": (4, 4, 12, 8), "S1": (3, 3, 9, 6), "S0": (2, 2, 6, 4)}
EXPANSION = {
    "L": (4, 4, (4, 4, 4, 4, 3, 3, 3, 3, 3, 3, 3, 4, 4, 4, 4), (4, 4, 4, 3, 3, 3, 3, 4, 4, 4)),
    "S2": (4, 4, (4, 4, 3, 3, 3, 3, 3, 3, 4, 4, 4, 4), (4, 4, 3, 3, 3, 3, 4, 4)),
    "S1": (4, 4, (4, 4, 3, 3, 3, 3, 4, 4, 4), (4, 4, 3, 3, 4, 4)),
    "S0": (4, 4, (4, 3, 3, 3, 4, 4), (4, 3, 3, 4)),
}
NUM_VIT = {"L": 6, "S2": 4, "S1": 2, "S0": 2}
BN_EPS = 1e-5


class ConvNorm(nn.Module):
    """conv (bias=True unless told otherwise, symmetric padding (k-1)//2) + BatchNorm2d [+ activation]:
    timm ``ConvNorm`` / ``ConvNormAct``; keys ``conv.{weight,bias}``, ``bn.*``."""

    def __init__(self, cin: int, cout: int, k: int = 1, stride: int = 1, groups: int = 1, act: bool = False) -> None:
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, k, stride, (k - 1) // 2, groups=groups, bias=True)
        self.bn = nn.BatchNorm2d(cout, eps=BN_EPS)
        self.act = act

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        x = self.bn(self.conv(x))
        return F.gelu(x) if self.act else x


def _bias_idxs(q_side: int, k_side: int, q_step: int) -> torch.Tensor:
    """attention_bias_idxs: |dy| * k_side + |dx| between query position (step ``q_step`` on the key grid) and key."""
    kp = torch.stack(torch.meshgrid(torch.arange(k_side), torch.arange(k_side), indexing="ij")).flatten(1)
    qp = torch.stack(torch.meshgrid(torch.arange(0, q_side * q_step, q_step), torch.arange(0, q_side * q_step, q_step),
                                    indexing="ij")).flatten(1)
    rel = (qp[..., :, None] - kp[..., None, :]).abs()
    return rel[0] * k_side + rel[1]


class Attention2d(nn.Module):
    def __init__(self, dim: int, resolution: int, stride: int | None, key_dim: int = 32, num_heads: int = 8,
                 attn_ratio: int = 4) -> None:
        super().__init__()
        self.num_heads, self.key_dim, self.scale = num_heads, key_dim, key_dim ** -0.5
        if stride is not None:
            resolution = math.ceil(resolution / stride)
            self.stride_conv = ConvNorm(dim, dim, 3, stride, groups=dim)
            self.upsample = nn.Upsample(scale_factor=stride, mode="bilinear")
        else:
            self.stride_conv = None
            self.upsample = None
        self.resolution = resolution
        self.N = resolution * resolution
        self.d = attn_ratio * key_dim
        self.dh = self.d * num_heads
        kh = key_dim * num_heads
        self.q = ConvNorm(dim, kh)
        self.k = ConvNorm(dim, kh)
        self.v = ConvNorm(dim, self.dh)
        self.v_local = ConvNorm(self.dh, self.dh, 3, groups=self.dh)
        self.talking_head1 = nn.Conv2d(num_heads, num_heads, 1)
        self.talking_head2 = nn.Conv2d(num_heads, num_heads, 1)
        self.proj = ConvNorm(self.dh, dim, 1)
        self.attention_biases = nn.Parameter(torch.zeros(num_heads, self.N))
        self.register_buffer("attention_bias_idxs", _bias_idxs(resolution, resolution, 1), persistent=False)

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        b = x.shape[0]
        if self.stride_conv is not None:
            x = self.stride_conv(x)
        q = self.q(x).reshape(b, self.num_heads, -1, self.N).permute(0, 1, 3, 2)
        k = self.k(x).reshape(b, self.num_heads, -1, self.N)
        v = self.v(x)
        v_local = self.v_local(v)
        v = v.reshape(b, self.num_heads, -1, self.N).permute(0, 1, 3, 2)
        attn = (q @ k) * self.scale + self.attention_biases[:, self.attention_bias_idxs]
        attn = self.talking_head1(attn).softmax(dim=-1)
        attn = self.talking_head2(attn)
        x = (attn @ v).transpose(2, 3).reshape(b, self.dh, self.resolution, self.resolution) + v_local
        if self.upsample is not None:
            x = self.upsample(x)
        return self.proj(F.gelu(x))


class LocalGlobalQuery(nn.Module):
    def __init__(self, cin: int, cout: int) -> None:
        super().__init__()
        self.pool = nn.AvgPool2d(1, 2, 0)
        self.local = nn.Conv2d(cin, cin, 3, 2, 1, groups=cin)
        self.proj = ConvNorm(cin, cout, 1)

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return self.proj(self.local(x) + self.pool(x))


class Attention2dDownsample(nn.Module):
    def __init__(self, dim: int, out_dim: int, resolution: int, key_dim: int = 16, num_heads: int = 8,
                 attn_ratio: int = 4) -> None:
        super().__init__()
        self.num_heads, self.key_dim, self.scale = num_heads, key_dim, key_dim ** -0.5
        self.resolution, self.resolution2 = resolution, math.ceil(resolution / 2)
        self.N, self.N2 = resolution ** 2, self.resolution2 ** 2
        self.d = attn_ratio * key_dim
        self.dh = self.d * num_heads
        kh = key_dim * num_heads
        self.q = LocalGlobalQuery(dim, kh)
        self.k = ConvNorm(dim, kh, 1)
        self.v = ConvNorm(dim, self.dh, 1)
        self.v_local = ConvNorm(self.dh, self.dh, 3, 2, groups=self.dh)
        self.proj = ConvNorm(self.dh, out_dim, 1)
        self.attention_biases = nn.Parameter(torch.zeros(num_heads, self.N))
        self.register_buffer("attention_bias_idxs", _bias_idxs(self.resolution2, resolution, 2), persistent=False)

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        b = x.shape[0]
        q = self.q(x).reshape(b, self.num_heads, -1, self.N2).permute(0, 1, 3, 2)
        k = self.k(x).reshape(b, self.num_heads, -1, self.N)
        v = self.v(x)
        v_local = self.v_local(v)
        v = v.reshape(b, self.num_heads, -1, self.N).permute(0, 1, 3, 2)
        attn = ((q @ k) * self.scale + self.attention_biases[:, self.attention_bias_idxs]).softmax(dim=-1)
        x = (attn @ v).transpose(2, 3).reshape(b, self.dh, self.resolution2, self.resolution2) + v_local
        return self.proj(F.gelu(x))


class Downsample(nn.Module):
    def __init__(self, cin: int, cout: int, resolution: int, use_attn: bool) -> None:
        super().__init__()
        self.conv = ConvNorm(cin, cout, 3, 2)
        self.attn = Attention2dDownsample(cin, cout, resolution) if use_attn else None

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        out = self.conv(x)
        return self.attn(x) + out if self.attn is not None else out


class ConvMlpWithNorm(nn.Module):
    def __init__(self, dim: int, hidden: int) -> None:
        super().__init__()
        self.fc1 = ConvNorm(dim, hidden, 1, act=True)
        self.mid = ConvNorm(hidden, hidden, 3, groups=hidden, act=True)
        self.fc2 = ConvNorm(hidden, dim, 1)

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return self.fc2(self.mid(self.fc1(x)))


class LayerScale2d(nn.Module):
    def __init__(self, dim: int, init: float = 1e-5) -> None:
        super().__init__()
        self.gamma = nn.Parameter(init * torch.ones(dim))

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return x * self.gamma.view(1, -1, 1, 1)


class Block(nn.Module):
    def __init__(self, dim: int, mlp_ratio: float, resolution: int, stride: int | None, use_attn: bool) -> None:
        super().__init__()
        if use_attn:  # registration order matters for the reference's "last nn.Conv2d" Grad-CAM rule (web_ui.py:96-105)
            self.token_mixer = Attention2d(dim, resolution, stride)
            self.ls1 = LayerScale2d(dim)
        else:
            self.token_mixer = None
            self.ls1 = None
        self.mlp = ConvMlpWithNorm(dim, int(dim * mlp_ratio))
        self.ls2 = LayerScale2d(dim)

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        if self.token_mixer is not None:
            x = x + self.ls1(self.token_mixer(x))
        return x + self.ls2(self.mlp(x))


class Stage(nn.Module):
    def __init__(self, dim: int, dim_out: int, depth: int, resolution: int, downsample: bool, block_stride: int | None,
                 downsample_use_attn: bool, block_use_attn: bool, num_vit: int, mlp_ratios) -> None:
        super().__init__()
        if downsample:
            self.downsample = Downsample(dim, dim_out, resolution, downsample_use_attn)
            dim = dim_out
            resolution = math.ceil(resolution / 2)
        else:
            self.downsample = nn.Identity()
        if not isinstance(mlp_ratios, (tuple, list)):
            mlp_ratios = (mlp_ratios,) * depth
        remain = depth - num_vit - 1
        self.blocks = nn.Sequential(*[
            Block(dim, mlp_ratios[i], resolution, block_stride, block_use_attn and i > remain) for i in range(depth)
        ])

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return self.blocks(self.downsample(x))


class Stem4(nn.Sequential):
    def __init__(self, cin: int, cout: int) -> None:
        super().__init__()
        self.conv1 = ConvNorm(cin, cout // 2, 3, 2, act=True)
        self.conv2 = ConvNorm(cout // 2, cout, 3, 2, act=True)


class EfficientFormerV2Oracle(nn.Module):
    def __init__(self, num_classes: int = 2, variant: str = "S1", img_size: int = 224) -> None:
        super().__init__()
        dims, depths, ratios = WIDTHS[variant], DEPTHS[variant], EXPANSION[variant]
        self.stem = Stem4(3, dims[0])
        prev, stride, stages = dims[0], 4, []
        for i in range(4):
            stages.append(Stage(prev, dims[i], depths[i], math.ceil(img_size / stride), i > 0, 2 if i == 2 else None,
                                i >= 3, i >= 2, NUM_VIT[variant], ratios[i]))
            if i > 0:
                stride *= 2
            prev = dims[i]
        self.stages = nn.Sequential(*stages)
        self.norm = nn.BatchNorm2d(dims[-1], eps=BN_EPS)
        self.head = nn.Linear(dims[-1], num_classes)
        self.head_dist = nn.Linear(dims[-1], num_classes)
        for m in self.modules():  # timm init_weights: Linear trunc_normal(.02) / zero bias; convs keep torch defaults
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=0.02)
                nn.init.zeros_(m.bias)

    # ---- Grad-CAM split: A = raw output of the LAST nn.Conv2d in modules() order (web_ui.py:96-105) -----------
    def cam_target(self) -> nn.Module:
        last = None
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                last = m
        return last

    def forward_features_raw(self, x: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
        """-> (A [N,C,7,7] = raw ``stages.3.blocks.-1.mlp.fc2.conv`` output, the block's residual input)."""
        x = self.stem(x)
        for s in self.stages[:-1]:
            x = s(x)
        last = self.stages[-1]
        x = last.downsample(x)
        for b in last.blocks[:-1]:
            x = b(x)
        blk = last.blocks[-1]
        if blk.token_mixer is not None:
            x = x + blk.ls1(blk.token_mixer(x))
        a = blk.mlp.fc2.conv(blk.mlp.mid(blk.mlp.fc1(x)))
        return a, x

    def forward_head(self, a: torch.Tensor, resid: torch.Tensor) -> torch.Tensor:
        blk = self.stages[-1].blocks[-1]
        x = resid + blk.ls2(blk.mlp.fc2.bn(a))
        x = self.norm(x).mean((2, 3))
        return (self.head(x) + self.head_dist(x)) / 2

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        x = self.stages(self.stem(x))
        x = self.norm(x).mean((2, 3))
        return (self.head(x) + self.head_dist(x)) / 2


def cam_channel_weights(model: EfficientFormerV2Oracle, targets: torch.Tensor) -> torch.Tensor:
    """Closed-form Grad-CAM weights: everything after A is linear, so dlogit_c/dA is constant over
    (h, w):  w[n,k] = (W_head + W_dist)[c_n,k]/2 * s_norm[k] * ls2[k] * s_fc2bn[k] / (h*w)."""
    blk = model.stages[-1].blocks[-1]
    bn, nm = blk.mlp.fc2.bn, model.norm
    s_bn = bn.weight / torch.sqrt(bn.running_var + bn.eps)
    s_nm = nm.weight / torch.sqrt(nm.running_var + nm.eps)
    wfc = (model.head.weight + model.head_dist.weight)[targets] / 2
    side = 7
    return wfc * (s_nm * blk.ls2.gamma * s_bn).view(1, -1) / (side * side)


def randomize_(model: nn.Module, g: torch.Generator) -> None:
    """Synthetic weights that exercise every term: BN stats/affine (SURVEY 8d recipe), layer scales of O(0.1-1)
    (the 1e-5 default would hide the blocks), attention biases, talking-head mixes near identity, conv biases, heads."""
    for name, m in model.named_modules():
        with torch.no_grad():
            if isinstance(m, nn.BatchNorm2d):
                c = m.num_features
                m.running_mean.copy_(torch.randn(c, generator=g) * 0.1)
                m.running_var.copy_(torch.rand(c, generator=g) + 0.5)
                m.weight.copy_(torch.rand(c, generator=g) + 0.5)
                m.bias.copy_(torch.randn(c, generator=g) * 0.1)
            elif isinstance(m, LayerScale2d):
                m.gamma.copy_(torch.rand(m.gamma.shape, generator=g) * 0.4 + 0.1)
            elif isinstance(m, (Attention2d, Attention2dDownsample)):
                m.attention_biases.copy_(torch.randn(m.attention_biases.shape, generator=g))
            elif isinstance(m, nn.Conv2d) and name.split(".")[-1].startswith("talking_head"):
                h = m.weight.shape[0]
                m.weight.copy_((torch.eye(h) + torch.randn(h, h, generator=g) * 0.2).view(h, h, 1, 1))
                m.bias.copy_(torch.randn(h, generator=g) * 0.1)
            elif isinstance(m, nn.Conv2d):
                m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
            elif isinstance(m, nn.Linear):
                m.weight.copy_(torch.randn(m.weight.shape, generator=g) / math.sqrt(m.weight.size(1)))
                m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)


def make_seeded_oracle(seed: int = 0, num_classes: int = 2, variant: str = "S1") -> EfficientFormerV2Oracle:
    g = torch.Generator().manual_seed(seed)
    prev = torch.random.get_rng_state()
    torch.manual_seed(seed)
    try:
        model = EfficientFormerV2Oracle(num_classes, variant)
    finally:
        torch.random.set_rng_state(prev)
    randomize_(model, g)
    return model.eval()


def make_calibrated_oracle(seed: int = 0, num_classes: int = 2, calib_n: int = 16) -> EfficientFormerV2Oracle:
    """BN running stats calibrated on a synthetic batch so activations stay O(1) through the whole net and the logits
    depend on the image (std ~2): the network the fp32 |dlogit| <= 1e-3 bound is checked on."""
    from .effnet_b0 import synth_images

    model = make_seeded_oracle(seed, num_classes)
    bns = [m for m in model.modules() if isinstance(m, nn.BatchNorm2d)]
    for m in bns:
        m.momentum = None
        m.reset_running_stats()
    model.train()
    with torch.no_grad():
        model(synth_images(calib_n, seed + 2000))
    model.eval()
    for m in bns:
        m.momentum = 0.1
    g = torch.Generator().manual_seed(seed + 1000)
    with torch.no_grad():
        x = synth_images(calib_n, seed + 2000)
        pooled = model.norm(model.stages(model.stem(x))).mean((2, 3))
        w = torch.randn(num_classes, pooled.shape[1], generator=g)
        raw = (pooled - pooled.mean(0)) @ w.t()
        w = w * (2.0 / raw.std().clamp_min(1e-6))
        bias = -(pooled.mean(0) @ w.t())
        model.head.weight.copy_(w * 1.1)
        model.head_dist.weight.copy_(w * 0.9)
        model.head.bias.copy_(bias + 0.1)
        model.head_dist.bias.copy_(bias - 0.1)
    return model


def state_dict_keys_sha(model: nn.Module) -> str:
    """sha256 over 'key shape' lines: lets a real timm install diff the layout without the weights."""
    lines = [f"{k} {tuple(v.shape)}" for k, v in model.state_dict().items()]
    return hashlib.sha256("\n".join(lines).encode()).hexdigest()
