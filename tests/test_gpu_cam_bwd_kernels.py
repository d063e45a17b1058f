"""Backward-DATA kernels of the FasterViT Grad-CAM program (csrc/cam_bwd.cu) against torch autograd on the same inputs:
what pytorch-grad-cam's ``loss.backward()`` (/root/reference/web_ui.py:275-279) computes through fastervit's
WindowAttention / LayerNorm / GELU / downsample conv / TokenInitializer pool.  fp32: 1e-4 relative to the largest
gradient; bf16 tensors (fp32 math inside): 2e-2."""

import ctypes as C
import math

import pytest
import torch
from torch.nn import functional as F

from deepfakedetection_b200 import _lib

pytestmark = pytest.mark.gpu
BF, FP = _lib.DT_BF16, _lib.DT_FP32


@pytest.fixture(autouse=True)
def _need_gpu():
    assert torch.cuda.is_available() and _lib.lib().b200df_device_count() > 0, "these tests need a B200"
    torch.cuda.set_device(0)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False


def P(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def S():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def tdt(dt):
    return torch.bfloat16 if dt == BF else torch.float32


def close(got, want, dt):
    tol = (2e-2 if dt == BF else 1e-4) * max(want.abs().max().item(), 1e-30)
    err = (got.float() - want).abs().max().item()
    assert err <= tol, f"max error {err} > {tol} (max |ref| {want.abs().max().item()})"


def gen(seed):
    return torch.Generator(device="cuda").manual_seed(seed)


# (sequences, tokens, head dim, heads): FasterViT-0 level 2 windows / carrier tokens / level 3, FasterViT-2 head dim 48
@pytest.mark.parametrize("dt", [FP, BF])
@pytest.mark.parametrize("nseq,ntok,d,heads", [(12, 53, 32, 8), (3, 16, 32, 8), (5, 49, 32, 16), (4, 53, 48, 8), (2, 7, 16, 2)])
def test_attention_bwd(nseq, ntok, d, heads, dt):
    g = gen(nseq * 100 + ntok + d)
    c = heads * d
    qkv = torch.randn(nseq * ntok, 3 * c, device="cuda", generator=g).to(tdt(dt))
    bias = torch.randn(heads, ntok, ntok, device="cuda", generator=g)
    dout = torch.randn(nseq * ntok, c, device="cuda", generator=g).to(tdt(dt))
    scale = 1.0 / math.sqrt(d)
    x = qkv.float().requires_grad_(True)
    q, k, v = (x.view(nseq, ntok, 3, heads, d).permute(2, 0, 3, 1, 4)[i] for i in range(3))
    out = (((q @ k.transpose(-2, -1)) * scale + bias).softmax(-1) @ v).transpose(1, 2).reshape(nseq * ntok, c)
    out.backward(dout.float())
    got = torch.empty_like(qkv)
    _lib.check(_lib.lib().b200df_k_attention_bwd(P(qkv), 3 * c, 0, c, 2 * c, nseq, ntok, d, heads, C.c_float(scale), P(bias),
                                                 P(dout), P(got), dt, S()))
    close(got, x.grad, dt)


@pytest.mark.parametrize("dt", [FP, BF])
@pytest.mark.parametrize("rows,c,res", [(500, 256, True), (77, 512, False), (196, 64, False), (33, 384, True)])
def test_layernorm_bwd(rows, c, res, dt):
    g = gen(rows + c)
    x = (torch.randn(rows, c, device="cuda", generator=g) * 2 + 0.3).to(tdt(dt))
    gamma = torch.rand(c, device="cuda", generator=g) + 0.5
    dy = torch.randn(rows, c, device="cuda", generator=g).to(tdt(dt))
    r = torch.randn(rows, c, device="cuda", generator=g).to(tdt(dt)) if res else None
    xf = x.float().requires_grad_(True)
    F.layer_norm(xf, (c,), gamma, torch.zeros_like(gamma), 1e-5).backward(dy.float())
    want = xf.grad + (r.float() if res else 0)
    got = torch.empty_like(x)
    _lib.check(_lib.lib().b200df_k_layernorm_bwd(P(x), P(gamma), P(dy), P(r), P(got), rows, c, C.c_float(1e-5), dt, S()))
    close(got, want, dt)


@pytest.mark.parametrize("dt", [FP, BF])
def test_gelu_bwd(dt):
    g = gen(3)
    pre = (torch.randn(5000 * 8, device="cuda", generator=g) * 2).to(tdt(dt))
    dh = torch.randn(5000 * 8, device="cuda", generator=g).to(tdt(dt))
    xf = pre.float().requires_grad_(True)
    F.gelu(xf).backward(dh.float())
    got = dh.clone()
    _lib.check(_lib.lib().b200df_k_gelu_bwd(P(pre), P(got), pre.numel(), dt, S()))
    close(got, xf.grad, dt)


@pytest.mark.parametrize("dt", [FP, BF])
@pytest.mark.parametrize("n,hin,cin,cout", [(3, 14, 256, 512), (9, 14, 384, 768), (2, 6, 24, 40)])
def test_conv3x3_s2_bwd(n, hin, cin, cout, dt):
    g = gen(n + hin + cin)
    w = (torch.randn(cout, cin, 3, 3, device="cuda", generator=g) / math.sqrt(9 * cin))
    wk = w.permute(0, 2, 3, 1).reshape(cout, 9, cin).contiguous().to(tdt(dt))
    dy = torch.randn(n, hin // 2, hin // 2, cout, device="cuda", generator=g).to(tdt(dt))
    x = torch.zeros(n, cin, hin, hin, device="cuda", requires_grad=True)
    F.conv2d(x, wk.float().view(cout, 3, 3, cin).permute(0, 3, 1, 2), None, stride=2, padding=1).backward(dy.float().permute(0, 3, 1, 2))
    got = torch.empty(n, hin, hin, cin, device="cuda", dtype=tdt(dt))
    _lib.check(_lib.lib().b200df_k_conv3x3_s2_bwd(P(dy), P(wk), P(got), n, hin, cin, cout, dt, S()))
    close(got, x.grad.permute(0, 2, 3, 1), dt)


@pytest.mark.parametrize("dt", [FP, BF])
def test_ct_pool_bwd(dt):
    """fastervit TokenInitializer at level 2: AvgPool2d(5, 3) of a 14 x 14 map -> 4 x 4 tokens in windowed (2 x 2 per window)
    order; backward = fp32 NCHW gradient of the map."""
    g = gen(9)
    n, side, c, kernel, stride, cs = 3, 14, 64, 5, 3, 2
    gct = torch.randn(n, 16, c, device="cuda", generator=g).to(tdt(dt))
    x = torch.zeros(n, c, side, side, device="cuda", requires_grad=True)
    pooled = F.avg_pool2d(x, kernel, stride)                                     # [n, c, 4, 4]
    ct = pooled.view(n, c, 2, cs, 2, cs).permute(0, 2, 4, 3, 5, 1).reshape(n, 16, c)
    ct.backward(gct.float())
    got = torch.empty(n, c, side, side, device="cuda")
    _lib.check(_lib.lib().b200df_k_ct_pool_bwd(P(gct), P(got), n, side, c, kernel, stride, cs, dt, S()))
    close(got, x.grad, FP)
