"""CPU checks for the efficientnet_pytorch B3 layout: the engine's tensor inventory, the
drop-in module's state_dict and the registry entry, against ``oracle/effnet_b3.py`` (no compute: no GPU here)."""

import ctypes as C

import pytest
import torch

from deepfakedetection_b200 import _lib, registry
from deepfakedetection_b200.module import B200EfficientNetB3, CamHookPoint


def _expected(arch):
    lib = _lib.lib()
    h = C.c_void_p()
    _lib.check(lib.b200df_create(arch, 2, 0, C.byref(h)))
    try:
        name, numel = C.c_char_p(), C.c_size_t()
        out = []
        for i in range(lib.b200df_num_tensors(h)):
            _lib.check(lib.b200df_tensor_info(h, i, C.byref(name), C.byref(numel)))
            out.append((name.value.decode(), int(numel.value)))
        ch, side = C.c_int(), C.c_int()
        _lib.check(lib.b200df_feature_shape(h, C.byref(ch), C.byref(side)))
        return out, (ch.value, side.value)
    finally:
        lib.b200df_destroy(h)


def test_engine_inventory_is_the_efficientnet_pytorch_state_dict():
    from oracle.effnet_b3 import EfficientNetB3Oracle

    sd = EfficientNetB3Oracle(2).state_dict()
    want = [(k, v.numel()) for k, v in sd.items() if not k.endswith("num_batches_tracked")]
    got, feat = _expected(_lib.ARCH_EFFICIENTNET_B3)
    assert got == want  # same names, sizes AND order (26 blocks, 496 tensors)
    assert feat == (1536, 7)
    assert _expected(_lib.ARCH_EFFICIENTNET_B0)[1] == (1280, 7)


def test_module_mirrors_the_layout_and_the_cam_target_rule():
    from oracle.effnet_b3 import EfficientNetB3Oracle

    ref = EfficientNetB3Oracle(2)
    m = B200EfficientNetB3(2)
    sd, msd = ref.state_dict(), m.state_dict()
    assert list(sd) == list(msd) and all(sd[k].shape == msd[k].shape for k in sd)
    m.load_state_dict(sd, strict=True)
    # web_ui.py:108-110: `_conv_head` is looked up first and must be an nn.Module
    assert isinstance(getattr(m, "_conv_head"), CamHookPoint)
    assert all(bn.eps == 1e-3 for bn in m.modules() if isinstance(bn, torch.nn.BatchNorm2d))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.zeros(1, 3, 224, 224))
    with pytest.raises(ValueError, match="efficientnet_b3 expects"):
        m(torch.zeros(1, 3, 300, 300))


def test_registry_serves_the_references_default_entry():
    spec = registry.get_model_spec("efficientnet_b3")
    # field for field the reference's exact entry (model_registry.py:50-58), builder swapped
    assert (spec.name, spec.train_module, spec.weights_key, spec.default_image_size) == (
        "efficientnet_b3", "trainers.efficientnet", "efficientnet_b3", 224)
    m = spec.builder("efficientnet_b3", 3)
    assert isinstance(m, B200EfficientNetB3) and m._fc.out_features == 3 and m.precision == "bf16"
    assert registry.get_model_spec("efficientnet_b3.b200_fp32").builder("efficientnet_b3.b200_fp32", 2).precision == "fp32"
