"""CPU tests of the EfficientFormerV2 / FasterViT oracles (oracle/efficientformerv2.py, oracle/fastervit.py).

There is no second implementation of either network on disk, so PARITY IS UNPINNED against timm / fastervit themselves;
what can be pinned is pinned here: the parameter counts of every published variant (a wrong width, depth, expansion
ratio, head count, bias flag or attention-table size changes them), the MAC count, the state_dict layout the backend
module exposes, the Grad-CAM target the reference's rule picks, closed-form CAM vs autograd, and regression snapshots."""

from pathlib import Path

import numpy as np
import pytest
import torch
from torch import nn

from oracle import efficientformerv2 as ef
from oracle import fastervit as fv
from oracle.effnet_b0 import synth_images

GOLDEN = Path(__file__).resolve().parent / "golden"


# timm results table / the EfficientFormerV2 paper (Table 1), parameters at 1000 classes incl. the distillation head
@pytest.mark.parametrize("variant,published_m", [("S0", 3.60), ("S1", 6.19), ("S2", 12.71), ("L", 26.32)])
def test_efficientformerv2_param_counts_match_published(variant, published_m):
    m = ef.EfficientFormerV2Oracle(1000, variant)
    n = sum(p.numel() for p in m.parameters())
    assert round(n / 1e6, 2) == published_m, n


# FasterViT paper (Table 2) / NVlabs model zoo, parameters at 1000 classes
@pytest.mark.parametrize("variant,published_m", [("faster_vit_0_224", 31.4), ("faster_vit_1_224", 53.4),
                                                 ("faster_vit_2_224", 75.9), ("faster_vit_3_224", 159.5)])
def test_fastervit_param_counts_match_published(variant, published_m):
    m = fv.FasterViTOracle(1000, variant)
    n = sum(p.numel() for p in m.parameters())
    assert round(n / 1e6, 1) == published_m, n


def _macs(model, x):
    total = [0.0]

    def hook(mod, _inp, out):
        if isinstance(mod, nn.Conv2d):
            total[0] += out.numel() / out.shape[0] * mod.weight[0].numel()
        else:
            total[0] += out.numel() / x.shape[0] * mod.weight.shape[1]

    hs = [m.register_forward_hook(hook) for m in model.modules() if isinstance(m, (nn.Conv2d, nn.Linear))]
    with torch.no_grad():
        model(x)
    for h in hs:
        h.remove()
    return total[0]


def test_mac_counts_match_published():
    x = synth_images(1, seed=0)
    assert 0.60e9 <= _macs(ef.EfficientFormerV2Oracle(1000, "S1").eval(), x) <= 0.66e9    # 0.65 GMACs (paper, incl. attention)
    assert 3.2e9 <= _macs(fv.FasterViTOracle(1000, "faster_vit_0_224").eval(), x) <= 3.5e9   # 3.3 GFLOPs (paper's MAC count)


def test_reference_cam_target_rule():
    """/root/reference/web_ui.py:96-105: the last nn.Conv2d in modules() order."""
    m = ef.make_seeded_oracle(0)
    names = {id(mod): n for n, mod in m.named_modules()}
    assert names[id(m.cam_target())] == "stages.3.blocks.5.mlp.fc2.conv"
    assert fv.make_seeded_oracle(0).cam_target()[0] == "levels.2.global_tokenizer.pos_embed"


def test_efficientformerv2_split_forward_and_closed_form_cam():
    m = ef.make_calibrated_oracle(0)
    x = synth_images(3, seed=4)
    with torch.no_grad():
        full = m(x)
        a, resid = m.forward_features_raw(x)
        assert torch.allclose(m.forward_head(a, resid), full, atol=1e-5)
    tgt = full.argmax(1)
    store = {}
    hd = m.cam_target().register_forward_hook(lambda _m, _i, out: (store.__setitem__("A", out), out.register_hook(lambda g: store.__setitem__("G", g)))[0] and None)
    with torch.enable_grad():
        m(x)[torch.arange(3), tgt].sum().backward()
    hd.remove()
    w_auto = store["G"].mean((2, 3))
    w_closed = ef.cam_channel_weights(m, tgt)
    assert (store["G"] - store["G"].mean((2, 3), keepdim=True)).abs().max() <= 1e-7   # spatially constant
    assert torch.allclose(w_auto, w_closed.detach(), atol=1e-7, rtol=1e-4)


def test_window_partition_round_trip():
    x = torch.randn(2, 8, 14, 14)
    w = fv.window_partition(x, 7)
    assert w.shape == (8, 49, 8)
    assert torch.equal(fv.window_reverse(w, 7, 14, 14), x)
    ct = torch.randn(2, 16, 8)
    assert torch.equal(fv.ct_window(fv.ct_dewindow(ct, 4, 4, 2), 4, 4, 2).reshape(2, 16, 8), ct)


def test_backend_modules_expose_the_same_state_dict_layout():
    from deepfakedetection_b200 import registry

    for name, oracle in (("efficientformerv2_s1", ef.make_seeded_oracle(0)), ("faster_vit_0_224", fv.make_seeded_oracle(0)),
                         ("faster_vit_2_224", fv.make_seeded_oracle(0, variant="faster_vit_2_224"))):
        mod = registry.get_model_spec(name).builder(name, 2)
        a = {k: tuple(v.shape) for k, v in mod.state_dict().items()}
        b = {k: tuple(v.shape) for k, v in oracle.state_dict().items()}
        assert a == b, (name, set(a) ^ set(b))
        missing, unexpected = mod.load_state_dict(oracle.state_dict(), strict=False)
        assert not missing and not unexpected


def test_engine_tensor_list_is_a_subset_of_the_state_dict():
    import ctypes as C

    from deepfakedetection_b200 import _lib

    lib = _lib.lib()
    for arch, oracle in ((_lib.ARCH_EFFICIENTFORMERV2_S1, ef.make_seeded_oracle(0)), (_lib.ARCH_FASTERVIT_0, fv.make_seeded_oracle(0)),
                         (_lib.ARCH_FASTERVIT_2, fv.make_seeded_oracle(0, variant="faster_vit_2_224"))):
        h = C.c_void_p()
        _lib.check(lib.b200df_create(arch, 2, 0, C.byref(h)))
        sd = oracle.state_dict()
        name, numel = C.c_char_p(), C.c_size_t()
        seen = set()
        for i in range(lib.b200df_num_tensors(h)):
            _lib.check(lib.b200df_tensor_info(h, i, C.byref(name), C.byref(numel)))
            k = name.value.decode()
            assert k in sd and sd[k].numel() == numel.value, k
            seen.add(k)
        # everything with learnable content is consumed (what is left are constant tables / counters)
        left = [k for k in sd if k not in seen and not k.endswith(("num_batches_tracked", "relative_bias", "relative_coords_table",
                                                                   "relative_position_index"))
                and "to_global_feature.pos" not in k]
        assert not left, left[:5]
        lib.b200df_destroy(h)


@pytest.mark.parametrize("fname,maker", [("efficientformerv2_s1_seed0.npz", lambda: ef.make_calibrated_oracle(seed=0)),
                                         ("faster_vit_0_224_seed0.npz", lambda: fv.make_calibrated_oracle(seed=0))])
def test_regression_snapshots(fname, maker):
    g = np.load(GOLDEN / fname)
    m = maker()
    x = synth_images(4, seed=123)
    np.testing.assert_allclose(x.double().sum((1, 2, 3)).numpy(), g["x_checksum"], rtol=1e-9)
    sd = m.state_dict()
    w = np.array([float(sd[k].double().abs().sum()) for k in sorted(sd) if "num_batches" not in k])
    np.testing.assert_allclose(w, g["w_checksum"], rtol=5e-4)
    with torch.no_grad():
        np.testing.assert_allclose(m(x).numpy(), g["logits"], atol=5e-3)
