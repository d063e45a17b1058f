"""End-to-end parity on a B200 for FasterViT-0 (BASELINE.json configs[4]) and FasterViT-2 (the
reference registry's default `faster_vit` entry, model_registry.py:68-74): registered backend module vs
the fp32 oracle on identical synthetic inputs and weights.  fp32 path max |dlogit| <= 1e-3; bf16 path max |dsoftmax|
<= 2e-2 with >= 99.9 % argmax agreement."""

import pytest
import torch

from deepfakedetection_b200 import _lib, registry

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _need_gpu():
    assert torch.cuda.is_available() and _lib.lib().b200df_device_count() > 0, "these tests need a B200"
    torch.cuda.set_device(0)


def build(oracle, name):
    model = registry.get_model_spec(name).builder(name, 2)
    model.to(torch.device("cuda")).eval()
    model.load_state_dict(oracle.state_dict(), strict=False)
    return model


@pytest.fixture(scope="module", params=["faster_vit_0_224", "faster_vit_2_224"])
def calibrated(request):
    from oracle.effnet_b0 import synth_images
    from oracle.fastervit import make_calibrated_oracle

    m = make_calibrated_oracle(seed=0, variant=request.param)
    x = synth_images(16, seed=11)
    with torch.no_grad():
        ref = m(x)
    return request.param, m, x, ref


def test_layer_by_layer_fp32():
    import sys
    from pathlib import Path

    sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tools"))
    import layer_diff

    assert layer_diff.main("faster_vit_0_224", "fp32", 2) == 0


def test_fp32_path_logits(calibrated):
    name, m, x, ref = calibrated
    model = build(m, name + ".b200_fp32")
    with torch.inference_mode():
        out = model(x.cuda())
    assert out.shape == ref.shape and out.dtype == torch.float32
    assert ref.std().item() > 0.5
    err = (out.cpu() - ref).abs().max().item()
    assert err <= 1e-3, f"{name}: max |dlogit| = {err}"
    for n in (1, 5):
        with torch.inference_mode():
            o = model(x[:n].cuda()).cpu()
        assert (o - ref[:n]).abs().max().item() <= 1e-3, n


def test_bf16_path_on_calibrated_weights(calibrated):
    name, m, x, ref = calibrated
    model = build(m, name)
    with torch.inference_mode():
        out = model(x.cuda()).cpu()
    rel = ((out - ref).pow(2).mean().sqrt() / ref.std()).item()
    assert rel <= 0.06, f"{name}: rms logit error / logit std = {rel}"


def test_bf16_path_softmax_and_argmax_2048():
    """north_star tolerance on 8 batches of 256 images for FasterViT-0."""
    from oracle.effnet_b0 import synth_images
    from oracle.fastervit import make_seeded_oracle

    ref_model = make_seeded_oracle(seed=0)
    model = build(ref_model, "faster_vit_0_224")
    worst, agree, total, refs, outs = 0.0, 0, 0, [], []
    for b in range(8):
        x = synth_images(256, seed=200 + b)
        with torch.no_grad():
            ref = ref_model(x)
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


def test_bf16_simt_paths_agree_with_tcgen05(monkeypatch):
    from oracle.effnet_b0 import synth_images
    from oracle.fastervit import make_seeded_oracle

    ref_model = make_seeded_oracle(seed=0)
    x = synth_images(8, seed=3).cuda()
    monkeypatch.setenv("B200DF_GEMM", "simt")
    simt = build(ref_model, "faster_vit_0_224")
    with torch.inference_mode():
        a = simt(x).cpu()
    monkeypatch.delenv("B200DF_GEMM")
    tc = build(ref_model, "faster_vit_0_224")
    with torch.inference_mode():
        b = tc(x).cpu()
    with torch.no_grad():
        ref = ref_model(x.cpu())
    assert (a - b).pow(2).mean().sqrt().item() <= 0.05 * ref.std().item()


def test_gradcam_target_is_the_reference_rule():
    """The reference's target rule picks `levels.2.global_tokenizer.pos_embed` (web_ui.py:94-112) on the backend module
    exactly as on the fastervit package's module tree."""
    from oracle.fastervit import make_seeded_oracle

    ref_model = make_seeded_oracle(seed=0)
    model = build(ref_model, "faster_vit_0_224")
    last = None
    for name, mod in model.named_modules():
        if isinstance(mod, torch.nn.Conv2d):
            last = name
    assert last == ref_model.cam_target()[0] == "levels.2.global_tokenizer.pos_embed"
    assert model._cam_point() is model.levels[2].global_tokenizer.pos_embed


@pytest.mark.parametrize("prec", [".b200_fp32", ""])
def test_gradcam_hook_protocol_through_the_transformer_blocks(calibrated, prec):
    """pytorch-grad-cam's own mechanism (forward hook on the target layer, tensor hook on its output, loss.backward()):
    a forward with grad enabled returns logits whose autograd edge to the target activation is the engine's backward
    program.  The activation and its gradient are compared with the oracle's; the map with the fused path's."""
    from deepfakedetection_b200.gradcam import ClassifierOutputTarget, GradCAM
    from oracle.gradcam import gradcam_on_layer

    name, m, x, ref = calibrated
    n = 3
    tgt = ref[:n].argmax(1)
    tgt[0] = 1 - tgt[0]
    _, _, cam_ref, a_ref, g_ref = gradcam_on_layer(m, m.cam_target()[1], x[:n], tgt)
    model = build(m, name + prec)
    store = {}
    hd = model._cam_point().register_forward_hook(
        lambda _m, _i, out: (store.__setitem__("A", out), out.register_hook(lambda g: store.__setitem__("G", g)))[0] and None)
    with torch.enable_grad():
        logits = model(x[:n].cuda())
        logits[torch.arange(n), tgt.cuda()].sum().backward()
    hd.remove()
    a, g = store["A"].detach().cpu(), store["G"].cpu()
    assert a.shape == g.shape == a_ref.shape == (n, a_ref.shape[1], 14, 14)
    if prec:
        torch.testing.assert_close(a, a_ref, atol=2e-4 * a_ref.abs().max().item(), rtol=1e-3)
        assert (g - g_ref).abs().max().item() <= 2e-3 * g_ref.abs().max().item()
    w, w_ref = g.mean((2, 3)), g_ref.mean((2, 3))
    cosw = torch.nn.functional.cosine_similarity(w, w_ref, dim=1).min().item()
    assert cosw >= 0.99, f"channel weights cosine {cosw}"
    # the packaged front end in hook mode == the fused path
    with GradCAM(model=model, target_layers=[model._cam_point()], use_hooks=True) as cam:
        gray = cam(input_tensor=x[:n].cuda(), targets=[ClassifierOutputTarget(int(c)) for c in tgt])
    _, _, fused = model.forward_cam(x[:n].cuda(), tgt)
    keep = cam_ref.flatten(1).max(1).values > 0
    cos = torch.nn.functional.cosine_similarity(torch.as_tensor(gray).flatten(1), fused.cpu().flatten(1), dim=1)[keep].min().item()
    assert cos >= 0.999, f"hooked vs fused cosine {cos}"
    # a backward without a matching forward is refused, not answered from stale checkpoints
    with pytest.raises(Exception):
        model._engine.head_backward(torch.zeros(n + 1, a.shape[1], 14, 14, device="cuda"), torch.zeros(n + 1, 2, device="cuda"))


@pytest.mark.parametrize("prec", [".b200_fp32", ""])
def test_gradcam_matches_autograd_through_the_transformer_blocks(calibrated, prec):
    """Grad-CAM on the pos_embed conv = backward through level 2 (HAT blocks with carrier tokens),
    the downsample and level 3.  Oracle: hooks + autograd on the fp32 restatement.  fp32 engine: the 14x14 map within
    1e-3 of its maximum; both precisions: cosine of the final map >= 0.99 (the CAM tolerance of the other backbones)."""
    from oracle.gradcam import gradcam_on_layer

    name, m, x, ref = calibrated
    n = 6
    tgt = ref[:n].argmax(1)
    tgt[1] = 1 - tgt[1]                       # an explicit non-argmax class as well
    _, low_ref, cam_ref, _a, _g = gradcam_on_layer(m, m.cam_target()[1], x[:n], tgt)
    assert low_ref.shape[1:] == (14, 14)
    model = build(m, name + prec)
    lg, low, cam = model.forward_cam(x[:n].cuda(), tgt)
    assert low.shape == (n, 14, 14) and cam.shape == (n, 224, 224)
    keep = low_ref.flatten(1).max(1).values > 1e-6 * low_ref.abs().max()
    assert bool(keep.any()), "every reference map is identically zero: the case tests nothing"
    cos = torch.nn.functional.cosine_similarity(cam.cpu().flatten(1), cam_ref.flatten(1), dim=1)[keep].min().item()
    assert cos >= 0.99, f"{name}{prec}: Grad-CAM cosine {cos}"
    if prec:
        err = (low.cpu() - low_ref).abs().max().item()
        assert err <= 1e-3 * max(low_ref.abs().max().item(), 1e-12), f"{name}: low-res map error {err} vs max {low_ref.abs().max().item()}"
    # argmax targets (target_cls = NULL) give the same maps as passing the argmax explicitly
    lg2, low2, _ = model.forward_cam(x[:n].cuda(), None)
    lg3, low3, _ = model.forward_cam(x[:n].cuda(), lg2.argmax(1).int())
    torch.testing.assert_close(low2, low3, rtol=0, atol=0)
    # the forward's logits are those of the plain forward
    with torch.inference_mode():
        plain = model(x[:n].cuda())
    torch.testing.assert_close(lg2, plain, rtol=0, atol=0)
