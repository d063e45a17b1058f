"""Parity evidence that means what it claims (VERDICT r1 "next" item 2), on a B200:

(a) argmax agreement >= 99.9 % / max |dsoftmax| <= 2e-2 on 4096 images (16 batches of 256) for B0 and B3;
(b) the bf16 stress network bound tied to the ideal-bf16 emulation (tools/bf16_sim.py) run INSIDE the test;
(c) direct tests of the exported tail kernels: softmax_pack, k_head_pool, k_cam, k_cam_upsample, head_backward;
(d) the SE-gated tcgen05 GEMM at real image sizes (several M tiles per image, cin % 16 == 8 tails);
(e) an out-of-range Grad-CAM class is rejected with B200DF_E_INVALID (pytorch-grad-cam raises IndexError, web_ui.py:278).
"""

import ctypes as C
import sys
from pathlib import Path

import pytest
import torch
from torch.nn import functional as F

from deepfakedetection_b200 import _lib, registry

pytestmark = pytest.mark.gpu
BF, FP = _lib.DT_BF16, _lib.DT_FP32
ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(autouse=True)
def _need_gpu():
    assert torch.cuda.is_available() and _lib.lib().b200df_device_count() > 0, "these tests need a B200"
    torch.cuda.set_device(0)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False


def build(oracle, name):
    model = registry.get_model_spec(name).builder(name, 2)
    model.to(torch.device("cuda")).eval()
    model.load_state_dict(oracle.state_dict(), strict=False)
    return model


def gen(seed):
    return torch.Generator(device="cuda").manual_seed(seed)


# ---- (a) -------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name,oracle_mod", [("efficientnet_b0", "oracle.effnet_b0"), ("efficientnet_b3", "oracle.effnet_b3")])
def test_bf16_argmax_and_softmax_on_4096_images(name, oracle_mod):
    import importlib

    from oracle.effnet_b0 import synth_images

    ref_model = importlib.import_module(oracle_mod).make_seeded_oracle(seed=0)   # the north_star recipe
    model = build(ref_model, name)
    worst, agree, total = 0.0, 0, 0
    for b in range(16):
        x = synth_images(256, seed=300 + b)
        with torch.no_grad():
            ref = ref_model(x)
        with torch.inference_mode():
            out = model(x.cuda()).cpu()
        worst = max(worst, (torch.softmax(out, 1) - torch.softmax(ref, 1)).abs().max().item())
        agree += (out.argmax(1) == ref.argmax(1)).sum().item()
        total += 256
    assert total == 4096
    assert worst <= 2e-2, f"{name}: max |dsoftmax| = {worst}"
    assert agree / total >= 0.999, f"{name}: argmax agreement {agree}/{total}"


# ---- (b) -------------------------------------------------------------------------------------------------------------
def test_bf16_stress_error_is_bounded_by_the_ideal_bf16_emulation(oracle_model):
    sys.path.insert(0, str(ROOT / "tools"))
    import bf16_sim

    from oracle.effnet_b0 import synth_images

    x = synth_images(64, seed=11)[:32]
    with torch.no_grad():
        ref = oracle_model(x)
        emu = bf16_sim.forward(oracle_model, x, dict(w=True, act=True, gate=True, trunk=True, A=True, tanh=True))
    model = build(oracle_model, "efficientnet_b0")
    with torch.inference_mode():
        out = model(x.cuda()).cpu()
    emu_rms = (emu - ref).pow(2).mean().sqrt().item()
    eng_rms = (out - ref).pow(2).mean().sqrt().item()
    emu_dp = (emu.softmax(1) - ref.softmax(1)).abs().max().item()
    eng_dp = (out.softmax(1) - ref.softmax(1)).abs().max().item()
    # the engine may not be worse than an ideal bf16 pipeline with the same rounding points (15 % + a small absolute slack:
    # single samples of this ill-conditioned network amplify one-ulp differences chaotically)
    assert eng_rms <= 1.15 * emu_rms + 1e-3, f"rms logit error engine {eng_rms} vs emulated {emu_rms}"
    assert eng_dp <= 1.15 * emu_dp + 5e-3, f"max |dsoftmax| engine {eng_dp} vs emulated {emu_dp}"


# ---- (c) -------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("c", [2, 5])
def test_softmax_pack_direct(c):
    from deepfakedetection_b200.evaluate import pack_records

    g = gen(c)
    logits = torch.randn(1000, c, device="cuda", generator=g) * 4
    labels = torch.randint(0, c, (1000,), device="cuda", generator=g)
    rec = pack_records(logits, labels)
    p = torch.softmax(logits, 1)
    torch.testing.assert_close(rec[:, :c], p, atol=2e-7, rtol=2e-6)
    assert torch.equal(rec[:, c].long(), torch.argmax(p, 1))
    assert torch.equal(rec[:, c + 1].long(), labels)


def _P(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("dt", [FP, BF])
def test_head_pool_and_cam_kernels_direct(dt):
    lib = _lib.lib()
    T = torch.bfloat16 if dt == BF else torch.float32
    g = gen(7)
    n, hw, c, nc = 5, 49, 1280, 2
    a = torch.randn(n, hw, c, device="cuda", generator=g).to(T)
    scale = torch.rand(c, device="cuda", generator=g) + 0.5
    shift = torch.randn(c, device="cuda", generator=g) * 0.2
    wfc = torch.randn(nc, c, device="cuda", generator=g) * 0.05
    z = a.float() * scale + shift
    pooled = torch.empty(n, c, device="cuda")
    _lib.check(lib.b200df_k_head_pool(_P(a), n, hw, c, _P(scale), _P(shift), _P(pooled), dt, _stream()))
    torch.testing.assert_close(pooled, F.silu(z).mean(1), atol=2e-6, rtol=1e-5)
    # Grad-CAM closed form: G = W[c_n] / hw * s * silu'(z); w = mean_hw G; cam = relu(sum_k w_k A_k)
    tgt = torch.tensor([0, 1, 1, 0, 1], dtype=torch.int32, device="cuda")
    sig = torch.sigmoid(z)
    dsilu = sig * (1 + z * (1 - sig))
    w = (wfc[tgt.long()][:, None, :] / hw * scale * dsilu).mean(1)           # [n, c]
    ref = (a.float() * w[:, None, :]).sum(2).clamp_min(0)                      # [n, hw]
    cam = torch.empty(n, hw, device="cuda")
    _lib.check(lib.b200df_k_cam(_P(a), n, hw, c, _P(scale), _P(shift), _P(wfc), nc, _P(tgt), _P(cam), dt, _stream()))
    torch.testing.assert_close(cam, ref, atol=1e-5 * max(1.0, float(ref.max())), rtol=1e-4)


def test_cam_upsample_direct_matches_oracle_postprocess():
    from oracle.gradcam import postprocess_cam

    g = gen(3)
    low = torch.rand(6, 7, 7, device="cuda", generator=g) * torch.tensor([1.0, 5.0, 0.1, 2.0, 1.0, 3.0], device="cuda").view(-1, 1, 1)
    low[3, 2:5] = 0
    out = torch.empty(6, 224, 224, device="cuda")
    _lib.check(_lib.lib().b200df_k_cam_upsample(_P(low), 6, 7, 224, _P(out), _stream()))
    ref = postprocess_cam(low.cpu(), (224, 224))
    assert (out.cpu() - ref).abs().max().item() <= 1e-5


def test_head_backward_matches_autograd(oracle_model):
    model = build(oracle_model, "efficientnet_b0.b200_fp32")
    with torch.inference_mode():
        model(torch.zeros(1, 3, 224, 224, device="cuda"))     # builds the engine
    g = gen(11)
    a = torch.randn(3, 1280, 7, 7, device="cuda", generator=g)
    gl = torch.randn(3, 2, device="cuda", generator=g)
    got = model._engine.head_backward(a, gl)
    bn, fc = oracle_model.bn2, oracle_model.classifier
    at = a.cpu().requires_grad_(True)
    with torch.enable_grad():
        z = F.silu(F.batch_norm(at, bn.running_mean, bn.running_var, bn.weight, bn.bias, False, 0.0, bn.eps)).mean((2, 3))
        (fc(z) * gl.cpu()).sum().backward()
    torch.testing.assert_close(got.cpu(), at.grad, atol=1e-7, rtol=1e-4)
    torch.testing.assert_close(model._engine.head(a).cpu(), fc(z).detach(), atol=1e-5, rtol=1e-5)


# ---- (d) -------------------------------------------------------------------------------------------------------------
# (images, rows_per_img, cin, cout, residual): SE-gated project GEMMs at real image sizes -- several M tiles per image,
# K > 64 with multi-tile images, and the cin % 16 == 8 tails (B3's 24 / 40 / 144-wide layers)
GATED_CASES = [(2, 12544, 32, 16, False), (2, 12544, 24, 24, True), (2, 12544, 40, 24, False), (3, 3136, 96, 24, False),
               (3, 3136, 144, 24, True), (3, 3136, 144, 32, True), (5, 784, 240, 40, True), (5, 784, 144, 40, False),
               (7, 196, 480, 80, True), (7, 196, 240, 80, False), (7, 196, 672, 112, True), (9, 49, 1152, 192, True),
               (3, 3136, 40, 24, True), (7, 196, 24, 16, False)]


@pytest.mark.parametrize("impl", [0, 1], ids=["tcgen05", "simt"])
@pytest.mark.parametrize("imgs,rpi,cin,cout,res", GATED_CASES)
def test_gated_gemm_at_real_image_sizes(imgs, rpi, cin, cout, res, impl):
    from gpu_util import k_pwconv
    from test_gpu_kernels import pw_ref

    m = imgs * rpi
    g = gen(m + cin + cout)
    x = torch.randn(m, cin, device="cuda", generator=g).bfloat16()
    w = (torch.randn(cout, cin, device="cuda", generator=g) / cin**0.5).bfloat16()
    bias = torch.randn(cout, device="cuda", generator=g) * 0.1
    gate = torch.rand(imgs, cin, device="cuda", generator=g)
    r = torch.randn(m, cout, device="cuda", generator=g).bfloat16() if res else None
    out = k_pwconv(x, w, bias, gate, rpi, r, 0, BF, impl).float()
    ref = pw_ref(x, w, bias, gate, rpi, r, 0, BF)
    assert torch.isfinite(out).all()
    torch.testing.assert_close(out, ref, atol=2e-2, rtol=2e-2)


# ---- (e) -------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["efficientnet_b0", "efficientformerv2_s1"])
def test_out_of_range_cam_class_is_rejected(name):
    import importlib

    mod = {"efficientnet_b0": "oracle.effnet_b0", "efficientformerv2_s1": "oracle.efficientformerv2"}[name]
    model = build(importlib.import_module(mod).make_seeded_oracle(seed=0), name)
    x = torch.zeros(2, 3, 224, 224, device="cuda")
    for bad in ([0, 2], [-1, 0]):
        with pytest.raises(_lib.B200DFError) as ei:
            model.forward_cam(x, torch.tensor(bad))
        assert ei.value.code == _lib.E_INVALID
    model.forward_cam(x, torch.tensor([0, 1]))   # in range: fine


# ---- metrics tail outside the binary / both-classes-present case ---------------------------------------------
@pytest.mark.parametrize("case", ["multiclass", "single_class"])
def test_device_metrics_tail_matches_reference_expressions_outside_binary(case):
    from deepfakedetection_b200 import evaluate

    g = torch.Generator().manual_seed(4)
    if case == "multiclass":
        c, n = 4, 500
        logits = torch.randn(n, c, generator=g) * 2
        targets = torch.randint(0, c, (n,), generator=g)
    else:
        c, n = 2, 64
        logits = torch.randn(n, c, generator=g)
        targets = torch.zeros(n, dtype=torch.long)
    probs = torch.softmax(logits, 1)
    rec = torch.cat([probs, probs.argmax(1, keepdim=True).float(), targets.view(-1, 1).float()], 1).cuda()
    got = evaluate.compute_metrics_device(rec, c, threshold=0.5 if c == 2 else None)
    want = evaluate.compute_metrics(probs, probs.argmax(1), targets, c, threshold=0.5 if c == 2 else None)
    assert got["confusion_matrix"] == want["confusion_matrix"]
    assert abs(got["accuracy"] - want["accuracy"]) < 1e-6
    assert ("roc_auc" in got) == ("roc_auc" in want)
    if "roc_auc" in want:
        assert abs(got["roc_auc"] - want["roc_auc"]) < 1e-9


# ---- uint8 path input validation ---------------------------------------------------------------------------------
def test_forward_u8_rejects_wrong_layouts(oracle_model):
    model = build(oracle_model, "efficientnet_b0")
    with pytest.raises(ValueError, match="uint8"):
        model.forward_u8(torch.zeros(2, 3, 224, 224, dtype=torch.uint8, device="cuda"))       # CHW instead of HWC
    with pytest.raises(ValueError, match="uint8"):
        model.forward_u8(torch.zeros(2, 160, 160, 3, dtype=torch.uint8, device="cuda"))       # wrong crop size
    with pytest.raises(ValueError, match="3 entries"):
        model.forward_u8(torch.zeros(2, 224, 224, 3, dtype=torch.uint8, device="cuda"), mean=(0.5,), std=(0.5,))
