"""Per-kernel parity on a B200: every kernel (through the C ABI) vs plain PyTorch fp32 math of the
same op on the same seeded inputs.  Shapes are the EfficientNet-B0 layer shapes."""

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


def gen(seed):
    return torch.Generator(device="cuda").manual_seed(seed)


def tol(dt):
    return dict(atol=2e-2, rtol=2e-2) if dt == BF else dict(atol=1e-4, rtol=1e-4)


@pytest.mark.parametrize("dt", [FP, BF])
def test_stem(dt):
    from gpu_util import k_stem

    g = gen(0)
    x = torch.randn(3, 3, 224, 224, device="cuda", generator=g)
    w = torch.randn(32, 3, 3, 3, device="cuda", generator=g) * 0.2
    b = torch.randn(32, device="cuda", generator=g) * 0.1
    if dt == BF:  # the bf16 stem is an implicit GEMM on tcgen05: inputs and weights are rounded to bf16
        ref = F.silu(F.conv2d(x.bfloat16().float(), w.bfloat16().float(), b, stride=2, padding=1)).permute(0, 2, 3, 1)
    else:
        ref = F.silu(F.conv2d(x, w, b, stride=2, padding=1)).permute(0, 2, 3, 1)
    out = k_stem(x, w.reshape(32, 27).contiguous(), b, dt).float()
    torch.testing.assert_close(out, ref, **tol(dt))


@pytest.mark.parametrize("dt", [FP, BF])
@pytest.mark.parametrize("hw", [(224, 224), (64, 96)])
def test_stem_40_channels_tf_same_padding(dt, hw):
    """EfficientNet-B3 (efficientnet_pytorch) stem: 40 channels, static TF-"same" pad (0, 1) -- the SIMT kernel on the
    fp32 path, the tcgen05 implicit GEMM (N = 48) on the bf16 path."""
    from gpu_util import k_stem_ex

    g = gen(40)
    h, wd = hw
    x = torch.randn(3, 3, h, wd, device="cuda", generator=g)
    w = torch.randn(40, 3, 3, 3, device="cuda", generator=g) * 0.2
    b = torch.randn(40, device="cuda", generator=g) * 0.1
    xr, wr = (x.bfloat16().float(), w.bfloat16().float()) if dt == BF else (x, w)
    ref = F.silu(F.conv2d(F.pad(xr, (0, 1, 0, 1)), wr, b, stride=2)).permute(0, 2, 3, 1)
    out = k_stem_ex(x, w.reshape(40, 27).contiguous(), b, dt, 40, 0).float()
    torch.testing.assert_close(out, ref, **tol(dt))


# the 3x3 stride-2 depthwise layers of B3 (pad (0, 1)) + ragged / odd sizes (odd sizes take the plain kernel)
DW_TFSAME_CASES = [(144, 112), (288, 28), (40, 56), (24, 16), (8, 9), (16, 13)]


@pytest.mark.parametrize("dt", [FP, BF])
@pytest.mark.parametrize("c,h", DW_TFSAME_CASES)
def test_dwconv_tf_same_padding(c, h, dt):
    from gpu_util import k_dwconv_ex, tdtype

    g = gen(c * 7 + h)
    n = 3 if h > 60 else 5
    x = torch.randn(n, h, h, c, device="cuda", generator=g).to(tdtype(dt))
    w = torch.randn(c, 1, 3, 3, device="cuda", generator=g) * 0.3
    b = torch.randn(c, device="cuda", generator=g) * 0.1
    xp = F.pad(x.float().permute(0, 3, 1, 2), (0, 2, 0, 2))  # (0, 1) is what the layer needs at even sizes; odd sizes
    ho = (h - 1) // 2 + 1                                    # use their own last tap: crop to the kernel's output size
    ref = F.silu(F.conv2d(xp, w, b, stride=2, groups=c))[:, :, :ho, :ho]
    out, pool = k_dwconv_ex(x, w.reshape(c, 9).t().contiguous(), b, 3, 2, 0, dt)
    torch.testing.assert_close(out.float(), ref.permute(0, 2, 3, 1), **tol(dt))
    torch.testing.assert_close(pool, ref.sum((2, 3)), atol=5e-2 if dt == BF else 2e-3, rtol=2e-3 if dt == BF else 1e-4)


# (C, H, k, stride) for the 16 depthwise layers of B0 (deduplicated) + ragged extras
DW_CASES = [(32, 112, 3, 1), (96, 112, 3, 2), (144, 56, 3, 1), (144, 56, 5, 2), (240, 28, 5, 1), (240, 28, 3, 2),
            (480, 14, 3, 1), (480, 14, 5, 1), (672, 14, 5, 1), (672, 14, 5, 2), (1152, 7, 5, 1), (1152, 7, 3, 1),
            (8, 9, 3, 1), (16, 13, 5, 2), (24, 5, 3, 2),
            # EfficientNet-B3 shapes (symmetric-pad layers; the pad-(0,1) ones are in DW_TFSAME_CASES)
            (40, 112, 3, 1), (24, 112, 3, 1), (192, 56, 3, 1), (192, 56, 5, 2), (288, 28, 5, 1), (576, 14, 3, 1),
            (816, 14, 5, 1), (816, 14, 5, 2), (1392, 7, 5, 1), (1392, 7, 3, 1), (2304, 7, 3, 1)]


@pytest.mark.parametrize("dt", [FP, BF])
@pytest.mark.parametrize("c,h,k,s", DW_CASES)
def test_dwconv(c, h, k, s, dt):
    from gpu_util import k_dwconv, tdtype

    g = gen(c * 100 + h + k + s)
    n = 3 if h > 60 else 5
    x = torch.randn(n, h, h, c, device="cuda", generator=g).to(tdtype(dt))
    w = torch.randn(c, 1, k, k, device="cuda", generator=g) * 0.3
    b = torch.randn(c, device="cuda", generator=g) * 0.1
    ref = F.silu(F.conv2d(x.float().permute(0, 3, 1, 2), w, b, stride=s, padding=(k - 1) // 2, groups=c))
    out, pool = k_dwconv(x, w.reshape(c, k * k).t().contiguous(), b, k, s, dt)
    torch.testing.assert_close(out.float(), ref.permute(0, 2, 3, 1), **tol(dt))
    # SE pool sums: the reference pools the fp32 activation (timm `se` mean_hw), so compare with the
    # fp32 reference sum (fast-tanh SiLU on the bf16 path: ~1e-3 relative per element)
    torch.testing.assert_close(pool, ref.sum((2, 3)), atol=5e-2 if dt == BF else 2e-3, rtol=2e-3 if dt == BF else 1e-4)


@pytest.mark.parametrize("c,rd", [(32, 8), (96, 4), (144, 6), (240, 10), (480, 20), (672, 28), (1152, 48)])
def test_se(c, rd):
    from gpu_util import k_se

    g = gen(c)
    n, hw = 7, 196
    pool = torch.randn(n, c, device="cuda", generator=g) * hw * 0.3
    w1 = torch.randn(rd, c, device="cuda", generator=g) * 0.1
    b1 = torch.randn(rd, device="cuda", generator=g) * 0.1
    w2 = torch.randn(c, rd, device="cuda", generator=g) * 0.3
    b2 = torch.randn(c, device="cuda", generator=g) * 0.1
    ref = torch.sigmoid(F.silu((pool / hw) @ w1.t() + b1) @ w2.t() + b2)
    out = k_se(pool, 1.0 / hw, w1, b1, w2.t().contiguous(), b2)
    torch.testing.assert_close(out, ref, atol=1e-5, rtol=1e-5)


def pw_ref(x, w, bias, gate, rpi, res, act, dt):
    xf = x.float()
    if gate is not None:
        xf = xf * gate.repeat_interleave(rpi, 0)[: xf.shape[0]]
        if dt == BF:
            xf = xf.bfloat16().float()  # the tensor-core path rounds the gated operand
    y = xf @ w.float().t() + bias
    if act:
        y = F.silu(y)
    if res is not None:
        y = y + res.float()
    return y


# (M, Cin, Cout, gate?, residual?, act) -- B0 expand / project / head shapes incl. K,N tails and M tails
PW_CASES = [
    (2 * 12544, 16, 96, False, False, 1),
    (2 * 3136, 24, 144, False, False, 1),
    (3 * 784, 40, 240, False, False, 1),
    (5 * 196, 80, 480, False, False, 1),
    (5 * 196, 112, 672, False, False, 1),
    (7 * 49, 192, 1152, False, False, 1),
    (2 * 12544, 32, 16, True, False, 0),
    (2 * 3136, 96, 24, True, False, 0),
    (2 * 3136, 144, 24, True, True, 0),
    (3 * 784, 240, 40, True, True, 0),
    (5 * 196, 480, 80, True, True, 0),
    (5 * 196, 672, 112, True, True, 0),
    (7 * 49, 672, 192, True, False, 0),
    (7 * 49, 1152, 192, True, True, 0),
    (7 * 49, 1152, 320, True, False, 0),
    (7 * 49, 320, 1280, False, False, 0),
    (64, 64, 64, False, False, 0),
    (1, 16, 8, False, False, 1),
    (130, 8, 16, True, True, 1),
]


@pytest.mark.parametrize("impl,dt", [(1, FP), (1, BF), (0, BF)], ids=["simt-fp32", "simt-bf16", "tcgen05-bf16"])
@pytest.mark.parametrize("m,cin,cout,use_gate,use_res,act", PW_CASES)
def test_pwconv(m, cin, cout, use_gate, use_res, act, impl, dt):
    from gpu_util import k_pwconv, tdtype

    g = gen(m + cin * 7 + cout)
    rpi = 49 if m % 49 == 0 else (m if m < 128 else 65)
    x = torch.randn(m, cin, device="cuda", generator=g).to(tdtype(dt))
    w = (torch.randn(cout, cin, device="cuda", generator=g) / cin**0.5).to(tdtype(dt))
    bias = torch.randn(cout, device="cuda", generator=g) * 0.1
    gate = torch.rand((m + rpi - 1) // rpi, cin, device="cuda", generator=g) if use_gate else None
    res = torch.randn(m, cout, device="cuda", generator=g).to(tdtype(dt)) if use_res else None
    out = k_pwconv(x, w, bias, gate, rpi, res, act, dt, impl).float()
    ref = pw_ref(x, w, bias, gate, rpi, res, act, dt)
    torch.cuda.synchronize()
    torch.testing.assert_close(out, ref, **(dict(atol=2e-2, rtol=2e-2) if dt == BF else dict(atol=2e-4, rtol=2e-4)))


def test_tcgen05_matches_simt_bitwise_close():
    """Same bf16 inputs through both GEMMs: only accumulation order differs."""
    from gpu_util import k_pwconv

    g = gen(5)
    x = torch.randn(4 * 196, 480, device="cuda", generator=g).bfloat16()
    w = (torch.randn(80, 480, device="cuda", generator=g) / 22).bfloat16()
    bias = torch.zeros(80, device="cuda")
    a = k_pwconv(x, w, bias, impl=0).float()
    b = k_pwconv(x, w, bias, impl=1).float()
    assert (a - b).abs().max() <= 2 ** -6 * max(1.0, float(b.abs().max()))
