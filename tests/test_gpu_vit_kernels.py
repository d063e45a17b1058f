"""Per-kernel parity on a B200 for the kernels EfficientFormerV2-S1 / FasterViT add (through the C ABI) vs plain
PyTorch fp32 math of the same op on the same seeded inputs.  Shapes are those nets' layer shapes."""

import math

import pytest
import torch
from torch.nn import functional as F

from deepfakedetection_b200 import _lib

pytestmark = pytest.mark.gpu

BF, FP = _lib.DT_BF16, _lib.DT_FP32
ACT_NONE, ACT_SILU, ACT_GELU, ACT_RELU = 0, 1, 2, 3


@pytest.fixture(autouse=True)
def _need_gpu():
    assert torch.cuda.is_available() and _lib.lib().b200df_device_count() > 0, "these tests need a B200"
    torch.cuda.set_device(0)
    # the PyTorch side is the fp32 REFERENCE of each op: no TF32 in its convolutions / matmuls
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False


def gen(seed):
    return torch.Generator(device="cuda").manual_seed(seed)


def tol(dt):
    return dict(atol=2e-2, rtol=2e-2) if dt == BF else dict(atol=2e-4, rtol=2e-4)


def act_ref(x, act):
    return {ACT_NONE: lambda t: t, ACT_SILU: F.silu, ACT_GELU: F.gelu, ACT_RELU: F.relu}[act](x)


@pytest.mark.parametrize("dt", [FP, BF])
@pytest.mark.parametrize("oc,act", [(16, ACT_GELU), (64, ACT_RELU)])
def test_stem_variants(oc, act, dt):
    """timm efficientformerv2 `stem.conv1` (3 -> 16, GELU) and fastervit `patch_embed.conv_down.0` (3 -> 64, ReLU)."""
    from gpu_util import k_stem_act

    g = gen(oc)
    x = torch.randn(3, 3, 224, 224, device="cuda", generator=g)
    w = torch.randn(oc, 3, 3, 3, device="cuda", generator=g) * 0.2
    b = torch.randn(oc, device="cuda", generator=g) * 0.1
    xr, wr = (x.bfloat16().float(), w.bfloat16().float()) if dt == BF else (x, w)
    ref = act_ref(F.conv2d(xr, wr, b, stride=2, padding=1), act).permute(0, 2, 3, 1)
    out = k_stem_act(x, w.reshape(oc, 27).contiguous(), b, dt, oc, 1, act).float()
    torch.testing.assert_close(out, ref, **tol(dt))


# (c, h, stride, act): EfficientFormerV2 mlp.mid (GELU), stride_conv / v_local / q.local (none), FasterViT pos_embed (none)
DW_ACT_CASES = [(128, 56, 1, ACT_GELU), (192, 28, 1, ACT_GELU), (480, 14, 1, ACT_GELU), (360, 14, 1, ACT_GELU),
                (896, 7, 1, ACT_GELU), (120, 14, 2, ACT_NONE), (1024, 7, 1, ACT_NONE), (512, 14, 2, ACT_NONE),
                (256, 14, 1, ACT_NONE), (24, 9, 1, ACT_GELU)]


@pytest.mark.parametrize("dt", [FP, BF])
@pytest.mark.parametrize("c,h,stride,act", DW_ACT_CASES)
def test_dwconv_activations(c, h, stride, act, dt):
    from gpu_util import k_dwconv_act, tdtype

    g = gen(c + h)
    n = 3 if h > 30 else 5
    x = torch.randn(n, h, h, c, device="cuda", generator=g).to(tdtype(dt))
    w = torch.randn(c, 1, 3, 3, device="cuda", generator=g) * 0.3
    b = torch.randn(c, device="cuda", generator=g) * 0.1
    ref = act_ref(F.conv2d(x.float().permute(0, 3, 1, 2), w, b, stride=stride, padding=1, groups=c), act).permute(0, 2, 3, 1)
    out = k_dwconv_act(x, w.reshape(c, 9).t().contiguous(), b, 3, stride, act, dt).float()
    torch.testing.assert_close(out, ref, **tol(dt))


# (cin, cout, h, stride, act, residual): EfficientFormerV2 stem.conv2 / downsamples, FasterViT patch_embed / ConvBlocks /
# downsamples (+ FasterViT-2 widths), plus ragged sizes
CONV_CASES = [
    (16, 32, 112, 2, ACT_GELU, False), (32, 48, 56, 2, ACT_NONE, False), (48, 120, 28, 2, ACT_NONE, False),
    (120, 224, 14, 2, ACT_NONE, False), (64, 64, 112, 2, ACT_RELU, False), (64, 64, 56, 1, ACT_GELU, False),
    (64, 64, 56, 1, ACT_NONE, True), (128, 128, 28, 1, ACT_GELU, False), (128, 128, 28, 1, ACT_NONE, True),
    (64, 128, 56, 2, ACT_NONE, False), (128, 256, 28, 2, ACT_NONE, False), (256, 512, 14, 2, ACT_NONE, False),
    (96, 96, 56, 1, ACT_GELU, False), (192, 384, 28, 2, ACT_NONE, False), (24, 40, 13, 1, ACT_GELU, True),
    (8, 16, 9, 2, ACT_RELU, False), (192, 192, 28, 1, ACT_NONE, True), (64, 64, 20, 1, ACT_GELU, True),
]


def _conv_inputs(cin, cout, h, dt, n):
    from gpu_util import tdtype

    g = gen(cin * 3 + cout + h)
    x = torch.randn(n, h, h, cin, device="cuda", generator=g).to(tdtype(dt))
    w = torch.randn(cout, cin, 3, 3, device="cuda", generator=g) * (1.0 / math.sqrt(9 * cin))
    b = torch.randn(cout, device="cuda", generator=g) * 0.1
    wk = w.permute(0, 2, 3, 1).reshape(cout, 9 * cin).contiguous().to(tdtype(dt))   # k = (ky*3+kx)*cin + c
    return x, w, b, wk


@pytest.mark.parametrize("impl", ["simt", "tc"])
@pytest.mark.parametrize("dt", [FP, BF])
@pytest.mark.parametrize("cin,cout,h,stride,act,res", CONV_CASES)
def test_conv3x3(cin, cout, h, stride, act, res, dt, impl):
    from gpu_util import k_conv3x3, tdtype

    if impl == "tc" and dt == FP:
        pytest.skip("the tcgen05 conv is bf16-only")
    n = 3 if h > 30 else 5   # odd batch: ragged image groups of the tcgen05 tiles
    x, w, b, wk = _conv_inputs(cin, cout, h, dt, n)
    ho = (h - 1) // stride + 1
    r = torch.randn(n, ho, ho, cout, device="cuda", generator=gen(5)).to(tdtype(dt)) if res else None
    wr = w.to(tdtype(dt)).float()
    ref = act_ref(F.conv2d(x.float().permute(0, 3, 1, 2), wr, b, stride=stride, padding=1), act).permute(0, 2, 3, 1)
    if res:
        ref = ref + r.float()
    out = k_conv3x3(x, wk, b, stride, act, dt, 1 if impl == "simt" else 0, r).float()
    torch.testing.assert_close(out, ref, **tol(dt))


def _attention_ref(q, k, v, scale, bias, th, vlocal, act):
    """q [S,H,nq,dk], k [S,H,nk,dk], v [S,H,nk,dv] fp32 -> [S, nq, H*dv]"""
    attn = (q @ k.transpose(-2, -1)) * scale
    if bias is not None:
        attn = attn + bias[None]
    if th is not None:
        attn = torch.einsum("oh,shij->soij", th[0], attn) + th[1].view(1, -1, 1, 1)
    attn = attn.softmax(-1)
    if th is not None:
        attn = torch.einsum("oh,shij->soij", th[2], attn) + th[3].view(1, -1, 1, 1)
    o = (attn @ v).permute(0, 2, 1, 3).reshape(q.shape[0], q.shape[2], -1)
    if vlocal is not None:
        o = o + vlocal
    return act_ref(o, act)


# (nseq, nq, nk, dk, dv, heads, talking heads, v_local, act): EfficientFormerV2 Attention2d / Attention2dDownsample,
# FasterViT window (53), carrier-token (16), level-3 (49) attention, FasterViT-2 head dim 48
ATTN_CASES = [(5, 49, 49, 32, 128, 8, True, True, ACT_GELU), (5, 49, 49, 32, 128, 8, True, True, ACT_NONE),
              (3, 49, 196, 16, 64, 8, False, True, ACT_GELU), (12, 53, 53, 32, 32, 8, False, False, ACT_NONE),
              (3, 16, 16, 32, 32, 8, False, False, ACT_NONE), (3, 49, 49, 32, 32, 16, False, False, ACT_NONE),
              (4, 53, 53, 48, 48, 8, False, False, ACT_NONE)]


@pytest.mark.parametrize("dt", [FP, BF])
@pytest.mark.parametrize("nseq,nq,nk,dk,dv,heads,th,vl,act", ATTN_CASES)
def test_attention(nseq, nq, nk, dk, dv, heads, th, vl, act, dt):
    from gpu_util import k_attention, tdtype

    g = gen(nq * 7 + nk + dk)
    T = tdtype(dt)
    q = torch.randn(nseq, heads, nq, dk, device="cuda", generator=g).to(T)
    k = torch.randn(nseq, heads, nk, dk, device="cuda", generator=g).to(T)
    v = torch.randn(nseq, heads, nk, dv, device="cuda", generator=g).to(T)
    bias = torch.randn(heads, nq, nk, device="cuda", generator=g)
    thp = None
    if th:
        eye = torch.eye(heads, device="cuda")
        thp = ((eye + 0.2 * torch.randn(heads, heads, device="cuda", generator=g)).contiguous(),
               0.1 * torch.randn(heads, device="cuda", generator=g),
               (eye + 0.2 * torch.randn(heads, heads, device="cuda", generator=g)).contiguous(),
               0.1 * torch.randn(heads, device="cuda", generator=g))
    vlocal = torch.randn(nseq, nq, heads * dv, device="cuda", generator=g).to(T) if vl else None
    scale = dk ** -0.5
    ref = _attention_ref(q.float(), k.float(), v.float(), scale, bias, thp, vlocal.float() if vl else None, act)
    if nq == nk and dk == dv:
        # fused qkv buffer [tokens, 3*H*d] as the FasterViT qkv Linear writes it: q | k | v column blocks
        qkv = torch.cat([t.permute(0, 2, 1, 3).reshape(nseq * nq, heads * dk) for t in (q, k, v)], dim=1).contiguous()
        out = k_attention(qkv, qkv, qkv, nseq, nq, nk, dk, dv, heads, scale, bias, thp, None if not vl else vlocal.reshape(nseq * nq, -1),
                          act, dt, 0, heads * dk, 2 * heads * dk)
    else:
        qf = q.permute(0, 2, 1, 3).reshape(nseq * nq, heads * dk).contiguous()
        kf = k.permute(0, 2, 1, 3).reshape(nseq * nk, heads * dk).contiguous()
        vf = v.permute(0, 2, 1, 3).reshape(nseq * nk, heads * dv).contiguous()
        out = k_attention(qf, kf, vf, nseq, nq, nk, dk, dv, heads, scale, bias, thp, None if not vl else vlocal.reshape(nseq * nq, -1),
                          act, dt)
    t = dict(atol=4e-2, rtol=2e-2) if dt == BF else dict(atol=2e-4, rtol=2e-4)
    torch.testing.assert_close(out.float().view(nseq, nq, -1), ref, **t)


@pytest.mark.parametrize("dt", [FP, BF])
@pytest.mark.parametrize("rows,c,eps", [(1000, 256, 1e-5), (777, 512, 1e-5), (3136, 64, 1e-6), (50, 768, 1e-5), (33, 128, 1e-6), (1001, 96, 1e-6), (67, 48, 1e-5), (5, 32, 1e-5)])
def test_layernorm(rows, c, eps, dt):
    from gpu_util import k_layernorm, tdtype

    g = gen(rows + c)
    x = (torch.randn(rows, c, device="cuda", generator=g) * 2 + 0.5).to(tdtype(dt))
    gamma = torch.rand(c, device="cuda", generator=g) + 0.5
    beta = torch.randn(c, device="cuda", generator=g) * 0.1
    ref = F.layer_norm(x.float(), (c,), gamma, beta, eps)
    out = k_layernorm(x, gamma, beta, eps, dt).float()
    torch.testing.assert_close(out, ref, **tol(dt))


@pytest.mark.parametrize("dt", [FP, BF])
@pytest.mark.parametrize("side,c,act", [(7, 1024, ACT_GELU), (5, 64, ACT_NONE)])
def test_upsample2x(side, c, act, dt):
    from gpu_util import k_upsample2x_act, tdtype

    g = gen(side + c)
    x = torch.randn(3, side, side, c, device="cuda", generator=g).to(tdtype(dt))
    ref = act_ref(torch.nn.Upsample(scale_factor=2, mode="bilinear")(x.float().permute(0, 3, 1, 2)), act).permute(0, 2, 3, 1)
    out = k_upsample2x_act(x, act, dt).float()
    torch.testing.assert_close(out, ref, **tol(dt))


@pytest.mark.parametrize("impl", [0, 1])
@pytest.mark.parametrize("m,cin,cout,act,res", [(49 * 7, 120, 480, ACT_GELU, False), (212 * 3, 256, 1024, ACT_GELU, False),
                                                (3136 * 2, 32, 128, ACT_GELU, False), (212 * 3, 1024, 256, ACT_NONE, True),
                                                (49 * 5, 512, 2048, ACT_GELU, False), (49 * 5, 768, 3072, ACT_GELU, False)])
def test_pwconv_gelu_epilogue(m, cin, cout, act, res, impl):
    """EfficientFormerV2 `mlp.fc1` / FasterViT `mlp.fc1` (GELU epilogue) and the residual projections, tcgen05 and SIMT."""
    from gpu_util import k_pwconv

    g = gen(m + cin)
    x = torch.randn(m, cin, device="cuda", generator=g).bfloat16()
    w = (torch.randn(cout, cin, device="cuda", generator=g) / math.sqrt(cin)).bfloat16()
    b = torch.randn(cout, device="cuda", generator=g) * 0.1
    r = torch.randn(m, cout, device="cuda", generator=g).bfloat16() if res else None
    ref = act_ref(x.float() @ w.float().t() + b, act)
    if res:
        ref = ref + r.float()
    out = k_pwconv(x, w, b, residual=r, act=act, dt=BF, impl=impl).float()
    torch.testing.assert_close(out, ref, atol=2e-2, rtol=2e-2)
