"""CPU tests of the oracle itself: pinned against torchvision's independent EfficientNet-B0,
against the committed golden vectors, and Grad-CAM autograd route vs closed form."""

import numpy as np
import torch
import torchvision

from oracle.effnet_b0 import make_seeded_oracle, synth_images, to_torchvision_state_dict
from oracle.gradcam import _minmax, gradcam_autograd, gradcam_closed_form


def test_param_count_matches_published_b0():
    m = make_seeded_oracle(0)
    assert sum(p.numel() for p in m.parameters()) == 4_010_110  # SURVEY.md 8a (2-class head)


def test_oracle_equals_torchvision_efficientnet_b0():
    m = make_seeded_oracle(3)
    tv = torchvision.models.efficientnet_b0(weights=None, num_classes=2).eval()
    tv.load_state_dict(to_torchvision_state_dict(m.state_dict()), strict=True)
    x = synth_images(2, seed=5)
    with torch.no_grad():
        a, b = m(x), tv(x)
    assert torch.allclose(a, b, atol=1e-5, rtol=0), (a - b).abs().max()


def test_oracle_matches_golden(oracle_model, golden):
    x = synth_images(4, seed=123)
    np.testing.assert_allclose(x.double().sum((1, 2, 3)).numpy(), golden["x_checksum"], rtol=1e-9)
    sd = oracle_model.state_dict()
    w = np.array([float(sd[k].double().abs().sum()) for k in sorted(sd) if "num_batches" not in k])
    np.testing.assert_allclose(w, golden["w_checksum"], rtol=2e-4)
    with torch.no_grad():
        logits = oracle_model(x)
        a = oracle_model.forward_features_raw(x)
    np.testing.assert_allclose(logits.numpy(), golden["logits"], atol=2e-3)
    np.testing.assert_allclose(a.mean((2, 3)).numpy(), golden["a_mean"], atol=2e-3)
    _, cam_low, cam = gradcam_autograd(oracle_model, x, torch.from_numpy(golden["targets"]))
    np.testing.assert_allclose(cam_low.numpy(), golden["cam_low"], atol=1e-4)
    np.testing.assert_allclose(cam[:, ::8, ::8].numpy(), golden["cam_ds"], atol=2e-3)


def test_gradcam_autograd_equals_closed_form(oracle_model):
    x = synth_images(3, seed=9)
    t = torch.tensor([0, 1, 1])
    l1, c1, f1 = gradcam_autograd(oracle_model, x, t)
    l2, c2, f2 = gradcam_closed_form(oracle_model, x, t)
    assert torch.allclose(l1, l2, atol=1e-5)
    assert (c1 - c2).abs().max() <= 1e-5 * max(1.0, float(c1.abs().max()))
    assert (f1 - f2).abs().max() < 1e-4
    assert float(f1.min()) >= 0 and float(f1.max()) <= 1.0 + 1e-6


def test_cam_resize_matches_cv2():
    cv2 = __import__("pytest").importorskip("cv2")
    g = torch.Generator().manual_seed(0)
    low = _minmax(torch.rand(2, 7, 7, generator=g))
    ref = np.stack([cv2.resize(m.numpy(), (224, 224)) for m in low])
    ours = torch.nn.functional.interpolate(low[:, None], size=(224, 224), mode="bilinear", align_corners=False)[:, 0]
    np.testing.assert_allclose(ours.numpy(), ref, atol=1e-6)
