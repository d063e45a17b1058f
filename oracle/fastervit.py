"""fp32 eager restatement of ``fastervit`` ``faster_vit_0_224`` / ``faster_vit_2_224`` (TEST INFRASTRUCTURE ONLY).

CPU oracle for the hot call ``logits = model(images)`` (/root/reference/orchestration/orchestrator.py:529, :590)
when the model is the one the reference registry builds through ``fastervit.create_model(model_name,
pretrained=False)`` followed by a fresh ``head = nn.Linear(in_features, num_classes)``
(/root/reference/orchestration/model_registry.py:43-47, prefix entry ``faster_vit`` :68-74; trainer call
/root/reference/trainers/fastervit.py:371-373; ``config/inference.yaml:38-50``).

``fastervit==1.0.0`` (/root/reference/requirements.txt:16) is a third-party dependency that is NOT on disk and
cannot be installed here, and the reference holds no golden vectors for this path, so the network is restated from
the published algorithm (NVlabs/FasterViT ``fastervit/models/faster_vit.py``; Hatamizadeh et al., "FasterViT: Fast
Vision Transformers with Hierarchical Attention", 2023) in that package's state_dict layout:

    patch_embed.conv_down.{0,1,3,4}                dense 3x3 s2 (no bias) + BN(eps 1e-4) + ReLU, twice
    levels.{0,1}.blocks.B.{conv1,norm1,conv2,norm2}  ConvBlock: x + BN(conv3x3(GELU(BN(conv3x3(x)))))  (conv bias)
    levels.L.downsample.{norm,reduction.0}         LayerNorm2d(eps 1e-6) -> dense 3x3 s2 (no bias), C -> 2C
    levels.2.global_tokenizer.{pos_embed / to_global_feature.pos}   dw3x3 (bias) -> AvgPool(5, 3): 4x4 carrier tokens
    levels.{2,3}.blocks.B.{pos_embed.cpb_mlp.{0,2}, norm1, attn.{qkv,proj,pos_emb_funct.cpb_mlp.{0,2}}, norm2,
                           mlp.{fc1,fc2}}          HAT block (window 7, level 3 has one window and no carrier tokens)
    levels.2.blocks.B.{hat_norm1,hat_attn.*,hat_norm2,hat_mlp.*,hat_pos_embed.*}   carrier-token attention
    norm (BN), head

**PARITY UNPINNED**: no second implementation of this network is on disk.  What IS pinned
(``tests/test_oracle_fastervit.py``): the parameter counts of the published variants (FasterViT-0 31.4 M,
-1 53.4 M, -2 75.9 M, -3 159.5 M at 1000 classes), window partition/reverse round trips, and the state_dict key
hash for a real install to diff.  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this.
"""

from __future__ import annotations

import math

import torch
from torch import nn
from torch.nn import functional as F

VARIANTS = {
    # name: dim, in_dim, depths, heads, layer_scale
    "faster_vit_0_224": dict(dim=64, in_dim=64, depths=(2, 3, 6, 5), heads=(2, 4, 8, 16), layer_scale=None),
    "faster_vit_1_224": dict(dim=80, in_dim=32, depths=(1, 3, 8, 5), heads=(2, 4, 8, 16), layer_scale=None),
    "faster_vit_2_224": dict(dim=96, in_dim=64, depths=(3, 3, 8, 5), heads=(2, 4, 8, 16), layer_scale=None),
    "faster_vit_3_224": dict(dim=128, in_dim=64, depths=(3, 3, 12, 5), heads=(2, 4, 8, 16), layer_scale=1e-5),
}
WINDOW = 7
CT_SIZE = 2
MLP_RATIO = 4


def window_partition(x: torch.Tensor, ws: int) -> torch.Tensor:
    b, c, h, w = x.shape
    x = x.view(b, c, h // ws, ws, w // ws, ws)
    return x.permute(0, 2, 4, 3, 5, 1).reshape(-1, ws * ws, c)


def window_reverse(windows: torch.Tensor, ws: int, h: int, w: int) -> torch.Tensor:
    b = int(windows.shape[0] / (h * w / ws / ws))
    x = windows.reshape(b, h // ws, w // ws, ws, ws, -1)
    return x.permute(0, 5, 1, 3, 2, 4).reshape(b, windows.shape[2], h, w)


def ct_dewindow(ct: torch.Tensor, w: int, h: int, ws: int) -> torch.Tensor:
    bs, n = ct.shape[0], ct.shape[2]
    ct2 = ct.view(-1, w // ws, h // ws, ws, ws, n).permute(0, 5, 1, 3, 2, 4)
    return ct2.reshape(bs, n, w * h).transpose(1, 2)


def ct_window(ct: torch.Tensor, w: int, h: int, ws: int) -> torch.Tensor:
    bs, n = ct.shape[0], ct.shape[2]
    ct = ct.view(bs, h // ws, ws, w // ws, ws, n)
    return ct.permute(0, 1, 3, 2, 4, 5)


class LayerNorm2d(nn.LayerNorm):
    """timm ``LayerNorm2d``: LayerNorm over the channel dim of an NCHW tensor, eps 1e-6."""

    def __init__(self, c: int) -> None:
        super().__init__(c, eps=1e-6)

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return F.layer_norm(x.permute(0, 2, 3, 1), self.normalized_shape, self.weight, self.bias, self.eps).permute(0, 3, 1, 2)


class PosEmbMLPSwinv2D(nn.Module):
    """Relative-position bias from a 2 -> 512 -> heads MLP over log-spaced offsets, 16*sigmoid, zero rows/columns for
    the carrier tokens in front.  Depends only on weights: the engine precomputes it at load time."""

    def __init__(self, window: int, num_heads: int, seq_length: int) -> None:
        super().__init__()
        self.window, self.num_heads = window, num_heads
        self.cpb_mlp = nn.Sequential(nn.Linear(2, 512, bias=True), nn.ReLU(inplace=True), nn.Linear(512, num_heads, bias=False))
        rel = torch.arange(-(window - 1), window, dtype=torch.float32)
        table = torch.stack(torch.meshgrid([rel, rel], indexing="ij")).permute(1, 2, 0).contiguous().unsqueeze(0)
        table = table / (window - 1) * 8
        table = torch.sign(table) * torch.log2(torch.abs(table) + 1.0) / math.log2(8)
        self.register_buffer("relative_coords_table", table)
        co = torch.arange(window)
        coords = torch.flatten(torch.stack(torch.meshgrid([co, co], indexing="ij")), 1)
        rc = (coords[:, :, None] - coords[:, None, :]).permute(1, 2, 0).contiguous()
        rc[:, :, 0] += window - 1
        rc[:, :, 1] += window - 1
        rc[:, :, 0] *= 2 * window - 1
        self.register_buffer("relative_position_index", rc.sum(-1))
        self.register_buffer("relative_bias", torch.zeros(1, num_heads, seq_length, seq_length))

    def bias(self, n_tokens: int) -> torch.Tensor:
        w2 = self.window * self.window
        tbl = self.cpb_mlp(self.relative_coords_table).view(-1, self.num_heads)
        b = tbl[self.relative_position_index.view(-1)].view(w2, w2, -1).permute(2, 0, 1).contiguous()
        b = 16 * torch.sigmoid(b)
        n_global = n_tokens - w2
        return F.pad(b, (n_global, 0, n_global, 0)).unsqueeze(0)

    def forward(self, attn: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return attn + self.bias(attn.shape[2])


class PosEmbMLPSwinv1D(nn.Module):
    """Absolute position embedding from a 2 -> 512 -> dim MLP over a normalised grid (rank 2)."""

    def __init__(self, dim: int, seq_length: int) -> None:
        super().__init__()
        self.cpb_mlp = nn.Sequential(nn.Linear(2, 512, bias=True), nn.ReLU(), nn.Linear(512, dim, bias=False))
        self.register_buffer("relative_bias", torch.zeros(1, seq_length, dim))

    def embedding(self, seq_length: int) -> torch.Tensor:
        side = int(seq_length ** 0.5)
        co = torch.arange(0, side, dtype=torch.float32)
        table = torch.stack(torch.meshgrid([co, co], indexing="ij")).contiguous().unsqueeze(0)
        table = (table - side // 2) / (side // 2)
        return self.cpb_mlp(table.flatten(2).transpose(1, 2))

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return x + self.embedding(x.shape[1])


class Mlp(nn.Module):
    def __init__(self, dim: int, hidden: int) -> None:
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden)
        self.fc2 = nn.Linear(hidden, dim)

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return self.fc2(F.gelu(self.fc1(x)))


class WindowAttention(nn.Module):
    def __init__(self, dim: int, num_heads: int, resolution: int, seq_length: int) -> None:
        super().__init__()
        self.num_heads = num_heads
        self.scale = (dim // num_heads) ** -0.5
        self.qkv = nn.Linear(dim, dim * 3, bias=True)
        self.proj = nn.Linear(dim, dim)
        self.pos_emb_funct = PosEmbMLPSwinv2D(resolution, num_heads, seq_length)

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        b, n, c = x.shape
        qkv = self.qkv(x).reshape(b, -1, 3, self.num_heads, c // self.num_heads).permute(2, 0, 3, 1, 4)
        q, k, v = qkv[0], qkv[1], qkv[2]
        attn = self.pos_emb_funct((q @ k.transpose(-2, -1)) * self.scale).softmax(dim=-1)
        return self.proj((attn @ v).transpose(1, 2).reshape(b, -1, c))


class HAT(nn.Module):
    def __init__(self, dim: int, num_heads: int, sr_ratio: int, last: bool, layer_scale) -> None:
        super().__init__()
        self.pos_embed = PosEmbMLPSwinv1D(dim, WINDOW ** 2)
        self.norm1 = nn.LayerNorm(dim)
        ct_per_window = CT_SIZE ** 2 if sr_ratio > 1 else 0
        ct_total = ct_per_window * sr_ratio * sr_ratio
        self.attn = WindowAttention(dim, num_heads, WINDOW, WINDOW ** 2 + ct_per_window)
        self.norm2 = nn.LayerNorm(dim)
        self.mlp = Mlp(dim, dim * MLP_RATIO)
        use_ls = layer_scale is not None
        self.gamma3 = nn.Parameter(layer_scale * torch.ones(dim)) if use_ls else 1
        self.gamma4 = nn.Parameter(layer_scale * torch.ones(dim)) if use_ls else 1
        self.sr_ratio = sr_ratio
        if sr_ratio > 1:
            self.hat_norm1 = nn.LayerNorm(dim)
            self.hat_norm2 = nn.LayerNorm(dim)
            self.hat_attn = WindowAttention(dim, num_heads, int(ct_total ** 0.5), ct_total)
            self.hat_mlp = Mlp(dim, dim * MLP_RATIO)
            self.hat_pos_embed = PosEmbMLPSwinv1D(dim, ct_total)
            self.gamma1 = nn.Parameter(layer_scale * torch.ones(dim)) if use_ls else 1
            self.gamma2 = nn.Parameter(layer_scale * torch.ones(dim)) if use_ls else 1
        self.last = last

    def forward(self, x: torch.Tensor, ct):  # noqa: D102
        x = self.pos_embed(x)
        if self.sr_ratio > 1:
            bg, ng, hg = ct.shape
            side = CT_SIZE * self.sr_ratio
            ct = ct_dewindow(ct, side, side, CT_SIZE)
            ct = self.hat_pos_embed(ct)
            ct = ct + self.gamma1 * self.hat_attn(self.hat_norm1(ct))
            ct = ct + self.gamma2 * self.hat_mlp(self.hat_norm2(ct))
            ct = ct_window(ct, side, side, CT_SIZE).reshape(x.shape[0], -1, x.shape[2])
            x = torch.cat((ct, x), dim=1)
        x = x + self.gamma3 * self.attn(self.norm1(x))
        x = x + self.gamma4 * self.mlp(self.norm2(x))
        if self.sr_ratio > 1:
            ctr, x = x.split([x.shape[1] - WINDOW * WINDOW, WINDOW * WINDOW], dim=1)
            ct = ctr.reshape(bg, ng, hg)
        return x, ct


class TokenInitializer(nn.Module):
    def __init__(self, dim: int, input_resolution: int) -> None:
        super().__init__()
        output_size = int(CT_SIZE * input_resolution / WINDOW)
        stride = int(input_resolution / output_size)
        kernel = input_resolution - (output_size - 1) * stride
        self.pos_embed = nn.Conv2d(dim, dim, 3, padding=1, groups=dim)
        self.to_global_feature = nn.Sequential()
        self.to_global_feature.add_module("pos", self.pos_embed)
        self.to_global_feature.add_module("pool", nn.AvgPool2d(kernel_size=kernel, stride=stride))

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        x = self.to_global_feature(x)
        b, c, h, w = x.shape
        ct = x.view(b, c, h // CT_SIZE, CT_SIZE, w // CT_SIZE, CT_SIZE)
        return ct.permute(0, 2, 4, 3, 5, 1).reshape(-1, h * w, c)


class ConvBlock(nn.Module):
    def __init__(self, dim: int, layer_scale=None) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(dim, dim, 3, 1, 1)
        self.norm1 = nn.BatchNorm2d(dim, eps=1e-5)
        self.conv2 = nn.Conv2d(dim, dim, 3, 1, 1)
        self.norm2 = nn.BatchNorm2d(dim, eps=1e-5)
        self.layer_scale = layer_scale is not None
        if self.layer_scale:
            self.gamma = nn.Parameter(layer_scale * torch.ones(dim))

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        y = self.norm2(self.conv2(F.gelu(self.norm1(self.conv1(x)))))
        if self.layer_scale:
            y = y * self.gamma.view(1, -1, 1, 1)
        return x + y


class Downsample(nn.Module):
    def __init__(self, dim: int) -> None:
        super().__init__()
        self.norm = LayerNorm2d(dim)
        self.reduction = nn.Sequential(nn.Conv2d(dim, 2 * dim, 3, 2, 1, bias=False))

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return self.reduction(self.norm(x))


class PatchEmbed(nn.Module):
    def __init__(self, in_dim: int, dim: int) -> None:
        super().__init__()
        self.conv_down = nn.Sequential(
            nn.Conv2d(3, in_dim, 3, 2, 1, bias=False), nn.BatchNorm2d(in_dim, eps=1e-4), nn.ReLU(),
            nn.Conv2d(in_dim, dim, 3, 2, 1, bias=False), nn.BatchNorm2d(dim, eps=1e-4), nn.ReLU())

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return self.conv_down(x)


class FasterViTLayer(nn.Module):
    def __init__(self, dim: int, depth: int, input_resolution: int, num_heads: int, conv: bool, downsample: bool,
                 only_local: bool, layer_scale) -> None:
        super().__init__()
        self.conv = conv
        if conv:
            self.blocks = nn.ModuleList([ConvBlock(dim) for _ in range(depth)])
        else:
            sr = input_resolution // WINDOW if not only_local else 1
            self.blocks = nn.ModuleList([HAT(dim, num_heads, sr, i == depth - 1, layer_scale) for i in range(depth)])
        self.downsample = Downsample(dim) if downsample else None   # registered BEFORE global_tokenizer
        self.do_gt = (not conv) and (not only_local) and input_resolution // WINDOW > 1
        if self.do_gt:
            self.global_tokenizer = TokenInitializer(dim, input_resolution)

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        ct = self.global_tokenizer(x) if self.do_gt else None
        _, _, h, w = x.shape
        if not self.conv:
            x = window_partition(x, WINDOW)
        for blk in self.blocks:
            if self.conv:
                x = blk(x)
            else:
                x, ct = blk(x, ct)
        if not self.conv:
            x = window_reverse(x, WINDOW, h, w)
        return x if self.downsample is None else self.downsample(x)


class FasterViTOracle(nn.Module):
    def __init__(self, num_classes: int = 2, variant: str = "faster_vit_0_224", resolution: int = 224) -> None:
        super().__init__()
        cfg = VARIANTS[variant]
        dim, depths, heads = cfg["dim"], cfg["depths"], cfg["heads"]
        self.patch_embed = PatchEmbed(cfg["in_dim"], dim)
        hat = (False, False, True, False)
        self.levels = nn.ModuleList([
            FasterViTLayer(dim * 2 ** i, depths[i], int(2 ** (-2 - i) * resolution), heads[i], i < 2, i < 3, not hat[i],
                           cfg["layer_scale"]) for i in range(4)
        ])
        self.num_features = dim * 8
        self.norm = nn.BatchNorm2d(self.num_features)
        self.head = nn.Linear(self.num_features, num_classes)
        for m in self.modules():  # fastervit _init_weights: Linear trunc_normal(.02)/0, LN 1/0, BN 1/0
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=0.02)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)

    def forward_features(self, x: torch.Tensor) -> torch.Tensor:
        x = self.patch_embed(x)
        for level in self.levels:
            x = level(x)
        return self.norm(x).mean((2, 3))

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return self.head(self.forward_features(x))

    def cam_target(self) -> tuple[str, nn.Module]:
        """The reference's rule (web_ui.py:96-105): the last nn.Conv2d in modules() order."""
        last = None
        for name, m in self.named_modules():
            if isinstance(m, nn.Conv2d):
                last = (name, m)
        return last


def randomize_(model: nn.Module, g: torch.Generator) -> None:
    """Synthetic weights that exercise every term (the package's default init has zero biases, unit norms, and tiny
    trunc_normal(0.02) Linear weights under which attention is uniform)."""
    for name, m in model.named_modules():
        with torch.no_grad():
            if isinstance(m, nn.BatchNorm2d):
                c = m.num_features
                m.running_mean.copy_(torch.randn(c, generator=g) * 0.1)
                m.running_var.copy_(torch.rand(c, generator=g) + 0.5)
                m.weight.copy_(torch.rand(c, generator=g) + 0.5)
                m.bias.copy_(torch.randn(c, generator=g) * 0.1)
            elif isinstance(m, nn.LayerNorm):
                m.weight.copy_(torch.rand(m.weight.shape, generator=g) * 0.5 + 0.75)
                m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
            elif isinstance(m, nn.Conv2d) and m.bias is not None:
                m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
            elif isinstance(m, nn.Linear):
                fan_in = m.weight.size(1)
                m.weight.copy_(torch.randn(m.weight.shape, generator=g) * (1.0 / math.sqrt(fan_in)))
                if m.bias is not None:
                    m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
            if isinstance(m, (HAT, ConvBlock)):
                for pname in ("gamma", "gamma1", "gamma2", "gamma3", "gamma4"):
                    p = getattr(m, pname, None)
                    if isinstance(p, nn.Parameter):
                        p.copy_(torch.rand(p.shape, generator=g) * 0.4 + 0.1)


def make_seeded_oracle(seed: int = 0, num_classes: int = 2, variant: str = "faster_vit_0_224") -> FasterViTOracle:
    g = torch.Generator().manual_seed(seed)
    prev = torch.random.get_rng_state()
    torch.manual_seed(seed)
    try:
        model = FasterViTOracle(num_classes, variant)
    finally:
        torch.random.set_rng_state(prev)
    randomize_(model, g)
    with torch.no_grad():  # damp the residual branches so 16 blocks of random weights keep activations O(1)
        for m in model.modules():
            if isinstance(m, (WindowAttention,)):
                m.proj.weight.mul_(0.3)
            if isinstance(m, Mlp):
                m.fc2.weight.mul_(0.3)
            if isinstance(m, ConvBlock):
                m.norm2.weight.mul_(0.3)
    return model.eval()


def make_calibrated_oracle(seed: int = 0, num_classes: int = 2, calib_n: int = 8,
                           variant: str = "faster_vit_0_224") -> FasterViTOracle:
    """BN running stats calibrated on a synthetic batch; head scaled to give logits of std ~2."""
    from .effnet_b0 import synth_images

    model = make_seeded_oracle(seed, num_classes, variant)
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
        pooled = model.forward_features(synth_images(calib_n, seed + 2000))
        w = torch.randn(num_classes, pooled.shape[1], generator=g)
        raw = (pooled - pooled.mean(0)) @ w.t()
        w = w * (2.0 / raw.std().clamp_min(1e-6))
        model.head.weight.copy_(w)
        model.head.bias.copy_(-(pooled.mean(0) @ w.t()) + torch.randn(num_classes, generator=g) * 0.1)
    return model
