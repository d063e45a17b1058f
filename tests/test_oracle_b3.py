"""Pins the EfficientNet-B3 (efficientnet_pytorch layout) oracle: architecture against torchvision's independently
written efficientnet_b3, static-same padding against its formula and an explicit pad + valid conv (CPU only)."""

from pathlib import Path

import pytest
import torch
from torch.nn import functional as F

ROOT = Path(__file__).resolve().parent.parent

from oracle import effnet_b3 as o3
from oracle.effnet_b0 import synth_images


def test_compound_scaling_table():
    blocks = o3.b3_blocks()
    assert o3.STEM_CHS == 40 and o3.HEAD_CHS == 1536
    assert len(blocks) == 26
    assert sorted({b["cout"] for b in blocks}) == [24, 32, 48, 96, 136, 232, 384]
    assert [b["rd"] for b in blocks[:3]] == [10, 6, 6]
    # the only asymmetric pads are the 3x3 stride-2 layers (stem, first block of stages 2 and 4)
    asym = [i for i, b in enumerate(blocks) if b["pad"][0] != b["pad"][1]]
    assert asym == [2, 8] and all(blocks[i]["pad"] == (0, 1) and blocks[i]["k"] == 3 for i in asym)
    assert o3.STEM_PAD == (0, 1)
    assert all(b["pad"] == ((b["k"] - 1) // 2,) * 2 for i, b in enumerate(blocks) if i not in asym)


@pytest.mark.parametrize("size,k,s,expect", [(300, 3, 2, (0, 1, 150)), (150, 3, 1, (1, 1, 150)), (75, 5, 2, (2, 2, 38)),
                                             (38, 3, 2, (0, 1, 19)), (19, 5, 2, (2, 2, 10)), (10, 3, 1, (1, 1, 10))])
def test_static_same_pad_formula(size, k, s, expect):
    assert o3.static_same_pad(size, k, s) == expect


def test_static_same_conv_is_pad_then_valid_conv():
    g = torch.Generator().manual_seed(0)
    conv = o3.StaticSamePadConv2d(8, 8, 3, 2, groups=8, pad=(0, 1))
    x = torch.randn(2, 8, 12, 12, generator=g)
    want = F.conv2d(F.pad(x, (0, 1, 0, 1)), conv.weight, None, 2, 0, 1, 8)
    assert torch.equal(conv(x), want) and want.shape[-1] == 6
    # and it is NOT the symmetric conv
    assert not torch.allclose(want, F.conv2d(x, conv.weight, None, 2, 1, 1, 8))


def test_state_dict_layout():
    sd = o3.EfficientNetB3Oracle(2).state_dict()
    assert "_conv_stem.weight" in sd and "_bn0.running_var" in sd and "_fc.bias" in sd
    assert "_blocks.0._expand_conv.weight" not in sd and "_blocks.2._expand_conv.weight" in sd
    assert sd["_blocks.25._project_conv.weight"].shape == (384, 2304, 1, 1)
    assert sd["_conv_head.weight"].shape == (1536, 384, 1, 1)
    assert sd["_blocks.9._se_reduce.weight"].shape == (24, 576, 1, 1)  # int(96 * 0.25)


def test_matches_torchvision_b3_with_symmetric_padding():
    tv = pytest.importorskip("torchvision")
    model = o3.make_calibrated_oracle(seed=3, num_classes=5, calib_n=4, padding="symmetric", eps=1e-5)
    ref = tv.models.efficientnet_b3(weights=None, num_classes=5).eval()
    for m in ref.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.eps = 1e-5
    sd = o3.to_torchvision_state_dict(model.state_dict())
    sd = {k: v for k, v in sd.items()}
    missing, unexpected = ref.load_state_dict(sd, strict=False)
    assert not unexpected and all(k.endswith("num_batches_tracked") for k in missing), (missing, unexpected)
    x = synth_images(3, 11)
    with torch.no_grad():
        a, b = model(x), ref(x)
    assert torch.allclose(a, b, rtol=0, atol=2e-5), (a - b).abs().max()
    assert a.std() > 0.1


def test_b3_oracle_matches_golden():
    """Committed vectors (tests/golden/make_golden_b3.py) guard the oracle, its pads and the weight recipe."""
    import numpy as np

    from oracle.effnet_b0 import synth_images
    from oracle.effnet_b3 import make_seeded_oracle
    from oracle.gradcam import gradcam_autograd

    golden = np.load(ROOT / "tests" / "golden" / "effnet_b3_seed0.npz")
    model = make_seeded_oracle(seed=0)
    x = synth_images(4, seed=123)
    np.testing.assert_allclose(x.double().sum((1, 2, 3)).numpy(), golden["x_checksum"], rtol=1e-9)
    sd = model.state_dict()
    w = np.array([float(sd[k].double().abs().sum()) for k in sorted(sd) if "num_batches" not in k])
    np.testing.assert_allclose(w, golden["w_checksum"], rtol=2e-4)
    with torch.no_grad():
        logits = model(x)
        a = model.forward_features_raw(x)
    np.testing.assert_allclose(logits.numpy(), golden["logits"], atol=1e-4)
    np.testing.assert_allclose(a.mean((2, 3)).numpy(), golden["a_mean"], atol=1e-4)
    _, cam_low, cam = gradcam_autograd(model, x, torch.from_numpy(golden["targets"]))
    np.testing.assert_allclose(cam_low.numpy(), golden["cam_low"], atol=1e-5)
    np.testing.assert_allclose(cam[:, ::8, ::8].numpy(), golden["cam_ds"], atol=2e-3)
