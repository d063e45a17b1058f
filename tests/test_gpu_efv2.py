"""End-to-end parity on a B200 for EfficientFormerV2-S1 (timm layout; BASELINE.json configs[3]):
registered backend module (nn.Module -> torch.library op -> C ABI -> CUDA) vs the fp32 oracle on identical synthetic
inputs and weights.  Tolerances are BASELINE.json's: fp32 path max |dlogit| <= 1e-3; bf16 path max |dsoftmax| <= 2e-2
with >= 99.9 % argmax agreement (4096 images); Grad-CAM cosine >= 0.99 after normalisation."""

import pytest
import torch

from deepfakedetection_b200 import _lib, registry
from deepfakedetection_b200.gradcam import ClassifierOutputTarget, GradCAM

pytestmark = pytest.mark.gpu

NAME = "efficientformerv2_s1"


@pytest.fixture(autouse=True)
def _need_gpu():
    assert torch.cuda.is_available() and _lib.lib().b200df_device_count() > 0, "these tests need a B200"
    torch.cuda.set_device(0)


def build(oracle, name):
    """The reference's load_model sequence (orchestrator.py:356-375)."""
    model = registry.get_model_spec(name).builder(name, 2)
    model.to(torch.device("cuda")).eval()
    model.load_state_dict(oracle.state_dict(), strict=False)
    return model


@pytest.fixture(scope="module")
def calibrated():
    from oracle.efficientformerv2 import make_calibrated_oracle
    from oracle.effnet_b0 import synth_images

    m = make_calibrated_oracle(seed=0)
    x = synth_images(32, seed=11)
    with torch.no_grad():
        ref = m(x)
    return m, x, ref


@pytest.fixture(scope="module")
def seeded():
    from oracle.efficientformerv2 import make_seeded_oracle

    return make_seeded_oracle(seed=0)


def test_layer_by_layer_fp32(calibrated):
    """Every debug checkpoint (stem, every block, the attention internals) against the oracle's module outputs: this is
    the test that says WHERE a parity failure comes from."""
    import sys
    from pathlib import Path

    sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tools"))
    import layer_diff

    assert layer_diff.main(NAME, "fp32", 2) == 0


def test_fp32_path_logits(calibrated):
    m, x, ref = calibrated
    model = build(m, NAME + ".b200_fp32")
    with torch.inference_mode():
        out = model(x.cuda())
    assert out.shape == ref.shape and out.dtype == torch.float32 and out.is_cuda
    err = (out.cpu() - ref).abs().max().item()
    assert ref.std().item() > 0.5, "calibrated oracle should give image-dependent logits"
    assert err <= 1e-3, f"max |dlogit| = {err}"


def test_fp32_path_ragged_batches(calibrated):
    m, x, ref = calibrated
    model = build(m, NAME + ".b200_fp32")
    for n in (1, 3, 17):
        with torch.inference_mode():
            out = model(x[:n].cuda()).cpu()
        assert (out - ref[:n]).abs().max().item() <= 1e-3, n


def test_bf16_path_softmax_and_argmax_4096(seeded):
    """north_star tolerance on 16 batches of 256 (BASELINE configs' batch): max |dsoftmax| <= 2e-2, argmax agreement
    >= 99.9 % vs the fp32 forward, weights = default init + randomised BN / LayerScale / biases (SURVEY.md 8d recipe)."""
    from oracle.effnet_b0 import synth_images

    model = build(seeded, NAME)
    worst, agree, total, refs, outs = 0.0, 0, 0, [], []
    for b in range(16):
        x = synth_images(256, seed=100 + b)
        with torch.no_grad():
            ref = seeded(x)
        with torch.inference_mode():
            out = model(x.cuda()).cpu()
        worst = max(worst, (torch.softmax(out, 1) - torch.softmax(ref, 1)).abs().max().item())
        agree += (out.argmax(1) == ref.argmax(1)).sum().item()
        total += x.shape[0]
        refs.append(ref)
        outs.append(out)
    ref, out = torch.cat(refs), torch.cat(outs)
    assert worst <= 2e-2, f"max |dsoftmax| = {worst}"
    assert agree / total >= 0.999, f"argmax agreement {agree}/{total}"
    rel = ((out - ref).pow(2).mean().sqrt() / ref.std()).item()
    assert rel <= 0.05, f"rms logit error / logit std = {rel}"


def test_bf16_path_on_calibrated_weights(calibrated):
    m, x, ref = calibrated
    model = build(m, NAME)
    with torch.inference_mode():
        out = model(x.cuda()).cpu()
    rel = ((out - ref).pow(2).mean().sqrt() / ref.std()).item()
    assert rel <= 0.06, f"rms logit error / logit std = {rel}"


def test_bf16_simt_paths_agree_with_tcgen05(seeded, monkeypatch):
    from oracle.effnet_b0 import synth_images

    x = synth_images(16, seed=3).cuda()
    monkeypatch.setenv("B200DF_GEMM", "simt")
    simt = build(seeded, NAME)
    with torch.inference_mode():
        a = simt(x).cpu()
    monkeypatch.delenv("B200DF_GEMM")
    tc = build(seeded, NAME)
    with torch.inference_mode():
        b = tc(x).cpu()
    with torch.no_grad():
        ref = seeded(x.cpu())
    assert (a - b).pow(2).mean().sqrt().item() <= 0.05 * ref.std().item()


def test_u8_input_path(seeded):
    g = torch.Generator().manual_seed(9)
    u8 = torch.randint(0, 256, (8, 224, 224, 3), generator=g, dtype=torch.uint8)
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    x = (u8.float() / 255.0 - torch.tensor(mean)) / torch.tensor(std)
    x = x.permute(0, 3, 1, 2).contiguous()
    model = build(seeded, NAME + ".b200_fp32")
    with torch.no_grad():
        ref = seeded(x)
    with torch.inference_mode():
        out = model.forward_u8(u8.cuda(), mean, std).cpu()
    assert (out - ref).abs().max().item() <= 1e-3


@pytest.mark.parametrize("prec", ["", ".b200_fp32"])
def test_gradcam_fused_matches_autograd_oracle(calibrated, prec):
    """Grad-CAM on the reference's target (the LAST nn.Conv2d = stages.3.blocks.5.mlp.fc2.conv, web_ui.py:96-105)."""
    from oracle.gradcam import postprocess_cam

    m, x, ref = calibrated
    x = x[:8]
    tgt = ref[:8].argmax(1)
    target = m.cam_target()
    store = {}

    def hook(_m, _i, out):
        store["A"] = out
        out.register_hook(lambda g: store.__setitem__("G", g))

    hd = target.register_forward_hook(hook)
    with torch.enable_grad():
        logits = m(x)
        logits[torch.arange(8), tgt].sum().backward()
    hd.remove()
    w = store["G"].mean((2, 3), keepdim=True)
    cam_low = (w * store["A"].detach()).sum(1).clamp_min(0)
    cam_ref = postprocess_cam(cam_low, (224, 224))

    model = build(m, NAME + prec)
    lg, low, cam = model.forward_cam(x.cuda(), tgt)
    cos = torch.nn.functional.cosine_similarity(cam.cpu().flatten(1), cam_ref.flatten(1), dim=1).min().item()
    assert cos >= 0.99, f"Grad-CAM cosine {cos}"
    if prec:
        assert (low.cpu() - cam_low).abs().max().item() <= 1e-3 * max(1.0, cam_low.abs().max().item())


def test_gradcam_hook_protocol(calibrated):
    """pytorch-grad-cam's mechanism on the unchanged target rule: hooks on the last nn.Conv2d (web_ui.py:96-114, 275-279)."""
    import sys
    from pathlib import Path

    from oracle.gradcam import postprocess_cam

    m, x, ref = calibrated
    model = build(m, NAME + ".b200_fp32")
    last = None
    for mod in model.modules():
        if isinstance(mod, torch.nn.Conv2d):
            last = mod
    assert getattr(model, "_conv_head", None) is None and last is model.stages[3].blocks[-1].mlp.fc2.conv
    x1 = x[:2].cuda()
    cls = int(ref[0].argmax())
    with GradCAM(model=model, target_layers=[last], use_hooks=True) as cam:
        gray = cam(input_tensor=x1[:1], targets=[ClassifierOutputTarget(cls)])[0]
    lg, low, cam_fused = model.forward_cam(x1[:1], torch.tensor([cls]))
    a = torch.as_tensor(gray).flatten()
    b = cam_fused[0].cpu().flatten()
    cos = torch.nn.functional.cosine_similarity(a, b, dim=0).item()
    assert cos >= 0.999, cos


def test_bad_cam_target_is_rejected(seeded):
    model = build(seeded, NAME)
    x = torch.zeros(2, 3, 224, 224, device="cuda")
    with pytest.raises(Exception):
        model.forward_cam(x, torch.tensor([0, 2]))
