"""fp32 eager restatement of ``efficientnet_pytorch`` EfficientNet-B3 (TEST INFRASTRUCTURE ONLY).

This is the CPU oracle for the reference's *registry default* backbone: ``efficientnet_b3`` is built by
``EfficientNet.from_name("efficientnet-b3")`` with a fresh ``_fc``
(/root/reference/orchestration/model_registry.py:32-36, 50-58) and called as ``model(images)`` on
224 x 224 crops (orchestrator.py:587-592; ``default_image_size=224``).  ``efficientnet_pytorch==0.7.1``
(/root/reference/requirements.txt:14) is a third-party dependency absent from /root/reference and from this
image, so the network is restated here from its published algorithm, in its state_dict layout
(SURVEY.md 8f row 2):

    _conv_stem, _bn0, _blocks.N.{_expand_conv,_bn0,_depthwise_conv,_bn1,_se_reduce,_se_expand,
    _project_conv,_bn2}, _conv_head, _bn1, _fc

What differs from the timm B0 layout (oracle/effnet_b0.py):
  * compound scaling: width x1.2 (divisor 8, never below 90 %), depth x1.4 (ceil) -> stem 40, stages
    24/32/48/96/136/232/384 with 2/3/3/5/5/6/2 blocks, head 1536;
  * the first stage's blocks (expand ratio 1) have no ``_expand_conv`` but are otherwise the same block;
  * SE squeeze width = ``max(1, int(block_input_filters * 0.25))``;
  * BatchNorm eps 1e-3;
  * TF "same" padding made STATIC for ``image_size=300`` (the B3 default resolution) by
    ``Conv2dStaticSamePadding``: total pad = max((ceil(i/s) - 1) * s + k - i, 0) with i the layer's input size in
    the 300-pixel network (300, 150, 75, 38, 19, 10), split ``left = total // 2``, ``right = total - left``.
    The reference feeds 224-pixel crops through those fixed pads.  For this network the only asymmetric layers are
    the 3x3 stride-2 convs (stem, stage 2, stage 4): pad (0, 1); every other conv is symmetric (k-1)/2.

Pinning: ``tests/test_oracle_b3.py`` (a) maps this state_dict, built with ``padding="symmetric"`` and BN eps
1e-5, onto the independently written ``torchvision.models.efficientnet_b3`` and requires identical logits -- that
pins widths, depths, SE widths, block wiring and the head; (b) checks the static-same pads against the formula and
against an explicit ``F.pad`` + valid conv.  Parity against efficientnet_pytorch itself is *unpinned* (not
installable here).

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this.
"""

from __future__ import annotations

import math

import torch
from torch import nn
from torch.nn import functional as F

# (kernel, stride, expand ratio, B0 out channels, B0 repeats) -- the EfficientNet base stage table
BASE_STAGES = [(3, 1, 1, 16, 1), (3, 2, 6, 24, 2), (5, 2, 6, 40, 2), (3, 2, 6, 80, 3), (5, 1, 6, 112, 3),
               (5, 2, 6, 192, 4), (3, 1, 6, 320, 1)]
WIDTH, DEPTH, NATIVE_SIZE = 1.2, 1.4, 300
BN_EPS = 1e-3


def round_filters(filters: int, width: float = WIDTH, divisor: int = 8) -> int:
    f = filters * width
    new = max(divisor, int(f + divisor / 2) // divisor * divisor)
    if new < 0.9 * f:
        new += divisor
    return int(new)


def round_repeats(repeats: int, depth: float = DEPTH) -> int:
    return int(math.ceil(depth * repeats))


def static_same_pad(size: int, k: int, s: int) -> tuple[int, int, int]:
    """-> (pad_before, pad_after, output size of the NATIVE network) for one spatial axis."""
    out = -(-size // s)
    total = max((out - 1) * s + k - size, 0)
    return total // 2, total - total // 2, out


def b3_blocks() -> list[dict]:
    """Block table: cin, cout, mid, k, stride, rd, pad (before, after), has_expand, skip."""
    blocks = []
    size = -(-NATIVE_SIZE // 2)  # after the stem
    cin = round_filters(32)
    for k, s, e, c, r in BASE_STAGES:
        cout = round_filters(c)
        for i in range(round_repeats(r)):
            stride = s if i == 0 else 1
            lo, hi, out = static_same_pad(size, k, stride)
            blocks.append(dict(cin=cin, cout=cout, mid=cin * e, k=k, stride=stride, rd=max(1, int(cin * 0.25)),
                               pad=(lo, hi), has_expand=e != 1, skip=stride == 1 and cin == cout))
            size = out
            cin = cout
    return blocks


STEM_CHS = round_filters(32)
HEAD_CHS = round_filters(1280)
STEM_PAD = static_same_pad(NATIVE_SIZE, 3, 2)[:2]


class StaticSamePadConv2d(nn.Conv2d):
    """Conv2d with a fixed ZeroPad2d in front (efficientnet_pytorch ``Conv2dStaticSamePadding``)."""

    def __init__(self, cin: int, cout: int, k: int, stride: int = 1, groups: int = 1, bias: bool = False,
                 pad: tuple[int, int] = (0, 0)) -> None:
        super().__init__(cin, cout, k, stride, 0, groups=groups, bias=bias)
        self.static_pad = pad

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        lo, hi = self.static_pad
        if lo or hi:
            x = F.pad(x, (lo, hi, lo, hi))
        return F.conv2d(x, self.weight, self.bias, self.stride, 0, 1, self.groups)


class MBConvBlock(nn.Module):
    def __init__(self, cfg: dict, symmetric: bool, eps: float) -> None:
        super().__init__()
        self.cfg = cfg
        mid, k = cfg["mid"], cfg["k"]
        pad = ((k - 1) // 2, (k - 1) // 2) if symmetric else cfg["pad"]
        if cfg["has_expand"]:
            self._expand_conv = StaticSamePadConv2d(cfg["cin"], mid, 1)
            self._bn0 = nn.BatchNorm2d(mid, eps=eps)
        self._depthwise_conv = StaticSamePadConv2d(mid, mid, k, cfg["stride"], groups=mid, pad=pad)
        self._bn1 = nn.BatchNorm2d(mid, eps=eps)
        self._se_reduce = StaticSamePadConv2d(mid, cfg["rd"], 1, bias=True)
        self._se_expand = StaticSamePadConv2d(cfg["rd"], mid, 1, bias=True)
        self._project_conv = StaticSamePadConv2d(mid, cfg["cout"], 1)
        self._bn2 = nn.BatchNorm2d(cfg["cout"], eps=eps)

    def forward(self, inputs: torch.Tensor) -> torch.Tensor:  # noqa: D102
        x = inputs
        if self.cfg["has_expand"]:
            x = F.silu(self._bn0(self._expand_conv(x)))
        x = F.silu(self._bn1(self._depthwise_conv(x)))
        s = F.adaptive_avg_pool2d(x, 1)
        s = self._se_expand(F.silu(self._se_reduce(s)))
        x = torch.sigmoid(s) * x
        x = self._bn2(self._project_conv(x))
        return x + inputs if self.cfg["skip"] else x  # drop_connect is training-only


class EfficientNetB3Oracle(nn.Module):
    """efficientnet_pytorch-keyed EfficientNet-B3, fp32 eager; ``forward`` == the reference's ``model(images)``."""

    def __init__(self, num_classes: int = 2, padding: str = "static_same", eps: float = BN_EPS) -> None:
        super().__init__()
        if padding not in ("static_same", "symmetric"):
            raise ValueError(padding)
        sym = padding == "symmetric"
        self.num_classes = num_classes
        self._conv_stem = StaticSamePadConv2d(3, STEM_CHS, 3, 2, pad=(1, 1) if sym else STEM_PAD)
        self._bn0 = nn.BatchNorm2d(STEM_CHS, eps=eps)
        self._blocks = nn.ModuleList(MBConvBlock(c, sym, eps) for c in b3_blocks())
        self._conv_head = StaticSamePadConv2d(b3_blocks()[-1]["cout"], HEAD_CHS, 1)
        self._bn1 = nn.BatchNorm2d(HEAD_CHS, eps=eps)
        self._fc = nn.Linear(HEAD_CHS, num_classes)
        for m in self.modules():  # synthetic init only (the reference always loads a checkpoint)
            if isinstance(m, nn.Conv2d):
                fan_out = m.kernel_size[0] * m.kernel_size[1] * m.out_channels // m.groups
                nn.init.normal_(m.weight, 0.0, math.sqrt(2.0 / fan_out))
                if m.bias is not None:
                    nn.init.zeros_(m.bias)

    def forward_features_raw(self, x: torch.Tensor) -> torch.Tensor:
        """Raw ``_conv_head`` output A (pre-BN): the Grad-CAM activation (/root/reference/web_ui.py:108-110)."""
        x = F.silu(self._bn0(self._conv_stem(x)))
        for b in self._blocks:
            x = b(x)
        return self._conv_head(x)

    def forward_head(self, a: torch.Tensor) -> torch.Tensor:
        x = F.silu(self._bn1(a))
        return self._fc(x.mean((2, 3)))  # dropout is training-only

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return self.forward_head(self.forward_features_raw(x))


def make_calibrated_oracle(seed: int = 0, num_classes: int = 2, calib_n: int = 16, padding: str = "static_same",
                           eps: float = BN_EPS) -> EfficientNetB3Oracle:
    """Synthetic weights that behave like a trained net (same recipe as oracle/effnet_b0.py): default conv init,
    BN running stats calibrated on a synthetic batch, mildly random BN affine, SE biases, classifier scaled to
    logits of std ~2.  Deterministic given ``seed``."""
    from .effnet_b0 import synth_images

    prev = torch.random.get_rng_state()
    torch.manual_seed(seed)
    try:
        model = EfficientNetB3Oracle(num_classes, padding, eps)
    finally:
        torch.random.set_rng_state(prev)
    g = torch.Generator().manual_seed(seed + 1000)
    bns = [m for m in model.modules() if isinstance(m, nn.BatchNorm2d)]
    for m in bns:
        with torch.no_grad():
            m.weight.copy_(torch.rand(m.num_features, generator=g) * 0.5 + 0.75)
            m.bias.copy_(torch.randn(m.num_features, generator=g) * 0.1)
        m.momentum = None
        m.reset_running_stats()
    for m in model.modules():
        if isinstance(m, nn.Conv2d) and m.bias is not None:
            with torch.no_grad():
                m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.2)
    model.train()
    with torch.no_grad():
        model(synth_images(calib_n, seed + 2000))
    model.eval()
    for m in bns:
        m.momentum = 0.01
    with torch.no_grad():
        x = synth_images(calib_n, seed + 2000)
        pooled = F.silu(model._bn1(model.forward_features_raw(x))).mean((2, 3))
        w = torch.randn(num_classes, HEAD_CHS, generator=g)
        raw = (pooled - pooled.mean(0)) @ w.t()
        model._fc.weight.copy_(w * (2.0 / raw.std().clamp_min(1e-6)))
        model._fc.bias.copy_(-(pooled.mean(0) @ model._fc.weight.t()) + torch.randn(num_classes, generator=g) * 0.1)
    return model


def make_seeded_oracle(seed: int = 0, num_classes: int = 2) -> EfficientNetB3Oracle:
    """BASELINE.json's parity recipe on the B3 layout (same as oracle/effnet_b0.make_seeded_oracle): default init,
    then randomised BN statistics/affine, SE biases and classifier.  Deterministic given ``seed``."""
    from .effnet_b0 import randomize_bn_, randomize_head_

    g = torch.Generator().manual_seed(seed)
    prev = torch.random.get_rng_state()
    torch.manual_seed(seed)
    try:
        model = EfficientNetB3Oracle(num_classes)
    finally:
        torch.random.set_rng_state(prev)
    randomize_bn_(model, g)
    randomize_head_(model, g)
    return model.eval()


def to_torchvision_state_dict(sd: dict[str, torch.Tensor]) -> dict[str, torch.Tensor]:
    """Map efficientnet_pytorch keys onto torchvision.models.efficientnet_b3 keys."""
    blocks = b3_blocks()
    # torchvision: features.0 stem, features.1..7 stages (Sequential of MBConv), features.8 head
    where, stage, idx, prev = [], 0, 0, None
    for i, b in enumerate(blocks):
        if i and (b["stride"] != 1 or b["cin"] != b["cout"]):
            stage, idx = stage + 1, 0
        where.append((stage + 1, idx))
        idx += 1
    out: dict[str, torch.Tensor] = {}
    for k, v in sd.items():
        p = k.split(".")
        if p[0] == "_conv_stem":
            out["features.0.0." + p[1]] = v
        elif p[0] == "_bn0":
            out["features.0.1." + p[1]] = v
        elif p[0] == "_conv_head":
            out["features.8.0." + p[1]] = v
        elif p[0] == "_bn1":
            out["features.8.1." + p[1]] = v
        elif p[0] == "_fc":
            out["classifier.1." + p[1]] = v
        elif p[0] == "_blocks":
            n = int(p[1])
            s, j = where[n]
            base = f"features.{s}.{j}.block."
            if blocks[n]["has_expand"]:
                m = {"_expand_conv": "0.0", "_bn0": "0.1", "_depthwise_conv": "1.0", "_bn1": "1.1",
                     "_se_reduce": "2.fc1", "_se_expand": "2.fc2", "_project_conv": "3.0", "_bn2": "3.1"}
            else:
                m = {"_depthwise_conv": "0.0", "_bn1": "0.1", "_se_reduce": "1.fc1", "_se_expand": "1.fc2",
                     "_project_conv": "2.0", "_bn2": "2.1"}
            out[base + m[p[2]] + "." + p[3]] = v
        else:
            raise KeyError(k)
    return out
