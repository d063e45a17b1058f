"""fp32 eager restatement of timm ``efficientnet_b0`` (TEST INFRASTRUCTURE ONLY).

This is the CPU oracle for the hot call ``logits = model(images)`` that the
reference makes at /root/reference/orchestration/orchestrator.py:529 and :590
(fp32, NCHW, eval mode, no autocast -- orchestrator.py:587-592).  The reference
builds its backbones through third-party packages that are not on disk
(``timm==1.0.20``, /root/reference/requirements.txt:95; builder call site
/root/reference/orchestration/model_registry.py:39-40), so the network is
restated here from the published EfficientNet-B0 architecture in *timm's
state_dict layout* (SURVEY.md section 8a row W):

    conv_stem, bn1, blocks.S.B.{conv_pw,bn1,conv_dw,bn2,se.conv_reduce,
    se.conv_expand,conv_pwl,bn3}, conv_head, bn2, classifier

Pinning: ``tests/test_oracle.py`` maps this state_dict onto the independently
written ``torchvision.models.efficientnet_b0`` and requires identical logits
(<= 1e-5), and checks committed golden vectors under ``tests/golden/``.
Parity against timm itself is *unpinned* (timm is not installable here).

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this.
"""

from __future__ import annotations

import math

import torch
from torch import nn
from torch.nn import functional as F

# (block type, kernel, stride, expand ratio, out channels, repeats) -- the B0 stage table.
B0_STAGES: list[tuple[str, int, int, int, int, int]] = [
    ("ds", 3, 1, 1, 16, 1),
    ("ir", 3, 2, 6, 24, 2),
    ("ir", 5, 2, 6, 40, 2),
    ("ir", 3, 2, 6, 80, 3),
    ("ir", 5, 1, 6, 112, 3),
    ("ir", 5, 2, 6, 192, 4),
    ("ir", 3, 1, 6, 320, 1),
]
STEM_CHS = 32
HEAD_CHS = 1280
BN_EPS = 1e-5


class BatchNormAct2d(nn.BatchNorm2d):
    """BatchNorm2d with the activation folded into the module (timm convention),
    so the state_dict keys are ``bnX.{weight,bias,running_mean,running_var}``."""

    def __init__(self, chs: int, act: bool = True) -> None:
        super().__init__(chs, eps=BN_EPS)
        self.apply_act = act

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        x = super().forward(x)
        return F.silu(x) if self.apply_act else x


class SqueezeExcite(nn.Module):
    def __init__(self, chs: int, rd_chs: int) -> None:
        super().__init__()
        self.conv_reduce = nn.Conv2d(chs, rd_chs, 1, bias=True)
        self.conv_expand = nn.Conv2d(rd_chs, chs, 1, bias=True)

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        s = x.mean((2, 3), keepdim=True)
        s = F.silu(self.conv_reduce(s))
        return x * torch.sigmoid(self.conv_expand(s))


class DepthwiseSeparableConv(nn.Module):
    def __init__(self, cin: int, cout: int, k: int, stride: int) -> None:
        super().__init__()
        self.has_skip = stride == 1 and cin == cout
        self.conv_dw = nn.Conv2d(cin, cin, k, stride, (k - 1) // 2, groups=cin, bias=False)
        self.bn1 = BatchNormAct2d(cin)
        self.se = SqueezeExcite(cin, max(1, cin // 4))
        self.conv_pw = nn.Conv2d(cin, cout, 1, bias=False)
        self.bn2 = BatchNormAct2d(cout, act=False)

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        y = self.bn1(self.conv_dw(x))
        y = self.se(y)
        y = self.bn2(self.conv_pw(y))
        return x + y if self.has_skip else y


class InvertedResidual(nn.Module):
    def __init__(self, cin: int, cout: int, k: int, stride: int, exp: int) -> None:
        super().__init__()
        mid = cin * exp
        self.has_skip = stride == 1 and cin == cout
        self.conv_pw = nn.Conv2d(cin, mid, 1, bias=False)
        self.bn1 = BatchNormAct2d(mid)
        self.conv_dw = nn.Conv2d(mid, mid, k, stride, (k - 1) // 2, groups=mid, bias=False)
        self.bn2 = BatchNormAct2d(mid)
        self.se = SqueezeExcite(mid, max(1, cin // 4))
        self.conv_pwl = nn.Conv2d(mid, cout, 1, bias=False)
        self.bn3 = BatchNormAct2d(cout, act=False)

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        y = self.bn1(self.conv_pw(x))
        y = self.bn2(self.conv_dw(y))
        y = self.se(y)
        y = self.bn3(self.conv_pwl(y))
        return x + y if self.has_skip else y


class EfficientNetB0Oracle(nn.Module):
    """timm-keyed EfficientNet-B0, fp32 eager.  ``forward`` == the reference's ``model(images)``."""

    def __init__(self, num_classes: int = 2) -> None:
        super().__init__()
        self.num_classes = num_classes
        self.conv_stem = nn.Conv2d(3, STEM_CHS, 3, 2, 1, bias=False)
        self.bn1 = BatchNormAct2d(STEM_CHS)
        stages = []
        cin = STEM_CHS
        for kind, k, s, e, cout, reps in B0_STAGES:
            blocks = []
            for r in range(reps):
                stride = s if r == 0 else 1
                if kind == "ds":
                    blocks.append(DepthwiseSeparableConv(cin, cout, k, stride))
                else:
                    blocks.append(InvertedResidual(cin, cout, k, stride, e))
                cin = cout
            stages.append(nn.Sequential(*blocks))
        self.blocks = nn.Sequential(*stages)
        self.conv_head = nn.Conv2d(cin, HEAD_CHS, 1, bias=False)
        self.bn2 = BatchNormAct2d(HEAD_CHS)
        self.classifier = nn.Linear(HEAD_CHS, num_classes)
        self._init_weights()

    # timm's "goog" initialisation for EfficientNets (restated): conv ~ N(0, sqrt(2/fan_out)),
    # BN gamma=1/beta=0, linear ~ U(+-1/sqrt(fan_out)), biases 0.
    def _init_weights(self) -> None:
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                fan_out = m.kernel_size[0] * m.kernel_size[1] * m.out_channels // m.groups
                nn.init.normal_(m.weight, 0.0, math.sqrt(2.0 / fan_out))
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)
            elif isinstance(m, nn.Linear):
                r = 1.0 / math.sqrt(m.weight.size(0))
                nn.init.uniform_(m.weight, -r, r)
                nn.init.zeros_(m.bias)

    def forward_features_raw(self, x: torch.Tensor) -> torch.Tensor:
        """Raw ``conv_head`` output A (pre-BN): the Grad-CAM activation
        (/root/reference/web_ui.py:96-114 picks the last nn.Conv2d)."""
        x = self.bn1(self.conv_stem(x))
        x = self.blocks(x)
        return self.conv_head(x)

    def forward_head(self, a: torch.Tensor) -> torch.Tensor:
        x = self.bn2(a)
        x = x.mean((2, 3))
        return self.classifier(x)

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return self.forward_head(self.forward_features_raw(x))


def randomize_bn_(model: nn.Module, generator: torch.Generator) -> None:
    """Randomise BN affine + running stats so BN folding is exercised
    (BASELINE.md section 2 / SURVEY.md section 8d synthetic-weight recipe)."""
    for m in model.modules():
        if isinstance(m, nn.BatchNorm2d):
            c = m.num_features
            with torch.no_grad():
                m.running_mean.copy_(torch.randn(c, generator=generator) * 0.1)
                m.running_var.copy_(torch.rand(c, generator=generator) + 0.5)
                m.weight.copy_(torch.rand(c, generator=generator) + 0.5)
                m.bias.copy_(torch.randn(c, generator=generator) * 0.1)


def randomize_head_(model: nn.Module, generator: torch.Generator, scale: float = 1.0) -> None:
    """Give SE biases and the classifier non-trivial values (defaults are zero bias),
    so every term of the forward is exercised by the parity tests."""
    for m in model.modules():
        if isinstance(m, nn.Conv2d) and m.bias is not None:
            with torch.no_grad():
                m.bias.copy_(torch.randn(m.bias.shape, generator=generator) * 0.2)
        if isinstance(m, nn.Linear):
            with torch.no_grad():
                m.weight.copy_(torch.randn(m.weight.shape, generator=generator) * (scale / math.sqrt(m.weight.size(1))))
                m.bias.copy_(torch.randn(m.bias.shape, generator=generator) * 0.1)


def make_seeded_oracle(seed: int = 0, num_classes: int = 2) -> EfficientNetB0Oracle:
    """Deterministic synthetic weights: default init, then randomised BN/SE-bias/classifier."""
    g = torch.Generator().manual_seed(seed)
    prev = torch.random.get_rng_state()
    torch.manual_seed(seed)
    try:
        model = EfficientNetB0Oracle(num_classes)
    finally:
        torch.random.set_rng_state(prev)
    randomize_bn_(model, g)
    randomize_head_(model, g)
    return model.eval()


def to_torchvision_state_dict(sd: dict[str, torch.Tensor]) -> dict[str, torch.Tensor]:
    """Map timm-layout keys onto torchvision.models.efficientnet_b0 keys
    (torchvision/models/efficientnet.py:105-169 MBConv, :233-344 network)."""
    out: dict[str, torch.Tensor] = {}
    for k, v in sd.items():
        p = k.split(".")
        if p[0] == "conv_stem":
            out["features.0.0." + p[1]] = v
        elif p[0] == "bn1":
            out["features.0.1." + p[1]] = v
        elif p[0] == "conv_head":
            out["features.8.0." + p[1]] = v
        elif p[0] == "bn2":
            out["features.8.1." + p[1]] = v
        elif p[0] == "classifier":
            out["classifier.1." + p[1]] = v
        elif p[0] == "blocks":
            s, b, name = int(p[1]), int(p[2]), p[3]
            rest = ".".join(p[4:])
            base = f"features.{s + 1}.{b}.block."
            if s == 0:  # DepthwiseSeparableConv: dw(0) se(1) project(2)
                m = {"conv_dw": "0.0", "bn1": "0.1", "conv_pw": "2.0", "bn2": "2.1"}
                se = "1"
            else:  # InvertedResidual: expand(0) dw(1) se(2) project(3)
                m = {"conv_pw": "0.0", "bn1": "0.1", "conv_dw": "1.0", "bn2": "1.1", "conv_pwl": "3.0", "bn3": "3.1"}
                se = "2"
            if name == "se":
                fc = {"conv_reduce": "fc1", "conv_expand": "fc2"}[p[4]]
                out[base + f"{se}.{fc}." + p[5]] = v
            else:
                out[base + m[name] + "." + rest] = v
        else:
            raise KeyError(k)
    return out


def synth_images(n: int, seed: int, size: int = 224) -> torch.Tensor:
    """Deterministic synthetic 'ImageNet-normalised' batch with per-image variety
    (noise + per-channel offset + a low-frequency pattern), so logits differ between images."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, 3, size, size, generator=g) * 0.6
    off = torch.randn(n, 3, 1, 1, generator=g) * 0.3
    lin = torch.linspace(-1, 1, size)
    yy, xx = torch.meshgrid(lin, lin, indexing="ij")
    fx = torch.randn(n, 3, 1, 1, generator=g) * 3
    fy = torch.randn(n, 3, 1, 1, generator=g) * 3
    ph = torch.rand(n, 3, 1, 1, generator=g) * 6.2831853
    return (x + off + 0.7 * torch.sin(fx * xx + fy * yy + ph)).contiguous()


def make_calibrated_oracle(seed: int = 0, num_classes: int = 2, calib_n: int = 32) -> EfficientNetB0Oracle:
    """Synthetic weights that behave like a trained net: default conv init, BN running stats
    *calibrated* on a synthetic batch (so activations stay O(1) through all 81 layers and logits
    depend on the input), mildly randomised BN affine, SE biases, and a classifier scaled to
    give logits of std ~2.  Deterministic given ``seed`` (CPU generator)."""
    model = make_seeded_oracle(seed, num_classes)
    g = torch.Generator().manual_seed(seed + 1000)
    bns = [m for m in model.modules() if isinstance(m, nn.BatchNorm2d)]
    for m in bns:
        with torch.no_grad():
            m.weight.copy_(torch.rand(m.num_features, generator=g) * 0.5 + 0.75)
            m.bias.copy_(torch.randn(m.num_features, generator=g) * 0.1)
        m.momentum = None
        m.reset_running_stats()
    model.train()
    with torch.no_grad():
        model(synth_images(calib_n, seed + 2000))
    model.eval()
    for m in bns:
        m.momentum = 0.1
    with torch.no_grad():
        x = synth_images(calib_n, seed + 2000)
        pooled = model.bn2(model.forward_features_raw(x)).mean((2, 3))
        w = torch.randn(num_classes, HEAD_CHS, generator=g)
        raw = (pooled - pooled.mean(0)) @ w.t()
        model.classifier.weight.copy_(w * (2.0 / raw.std().clamp_min(1e-6)))
        model.classifier.bias.copy_(-(pooled.mean(0) @ model.classifier.weight.t()) + torch.randn(num_classes, generator=g) * 0.1)
    return model
