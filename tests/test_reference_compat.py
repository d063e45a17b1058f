"""The UNMODIFIED reference modules on top of the backend (VERDICT r1 "next" item 3).

Needs the reference checkout (`/root/reference`, or $B200DF_REFERENCE_ROOT): present in the build container, absent on
the GPU boxes -- so the CPU half (registry, load_model lifecycle, transforms, CAM target rule, error behaviour) runs here
in `-m "not gpu"`, and the GPU half (`inference.py` end to end + `web_ui.predict_and_visualize`) runs wherever both a B200
and a checkout exist (tools/run_reference_on_backend.py; its log is committed under profiles/)."""

import importlib
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
REF = Path(os.environ.get("B200DF_REFERENCE_ROOT", "/root/reference"))
needs_ref = pytest.mark.skipif(not (REF / "orchestration" / "model_registry.py").exists(), reason="no reference checkout")


@pytest.fixture(scope="module")
def reference():
    sys.path.insert(0, str(ROOT / "tools"))
    import run_reference_on_backend as tool

    tool.install_reference(REF)
    reg = importlib.import_module("orchestration.model_registry")
    orch = importlib.import_module("orchestration.orchestrator")
    return reg, orch


@needs_ref
def test_every_model_of_the_reference_config_resolves_to_a_b200_builder(reference):
    import yaml

    reg, _ = reference
    from deepfakedetection_b200.module import B200Module

    cfg = yaml.safe_load((REF / "config" / "inference.yaml").read_text())
    for name in cfg["selection"]:                       # efficientnet_b3, efficientformerv2_s1, faster_vit_2_224
        spec = reg.get_model_spec(name)
        assert spec.name == name and spec.default_image_size == 224
        assert isinstance(spec.builder(name, 2), B200Module), name
    # exact entries win over the prefix table (model_registry.py:87-96); names we do not serve still reach the prefix entry
    assert reg.get_model_spec("efficientformerv2_s1").builder.__module__ == "deepfakedetection_b200.registry"
    assert reg.get_model_spec("efficientformerv2_s2").builder is reg._build_efficientformer
    with pytest.raises(RuntimeError, match="timm.create_model"):
        reg.get_model_spec("efficientformerv2_s2").builder("efficientformerv2_s2", 2)   # the shim refuses: nothing silently falls through
    with pytest.raises(KeyError, match="Add it to model_registry.py"):
        reg.get_model_spec("resnet50")


@needs_ref
@pytest.mark.parametrize("name,oracle_mod,kw", [("efficientnet_b0", "oracle.effnet_b0", {}), ("efficientnet_b3", "oracle.effnet_b3", {}),
                                               ("efficientformerv2_s1", "oracle.efficientformerv2", {}),
                                               ("faster_vit_2_224", "oracle.fastervit", {"variant": "faster_vit_2_224"})])
def test_reference_load_model_fills_the_containers(reference, tmp_path, name, oracle_mod, kw):
    """orchestrator.load_model (:350-377) as written: builder -> .to(device) -> .eval() -> torch.load -> unwrap -> load_state_dict."""
    _, orch = reference
    oracle = importlib.import_module(oracle_mod).make_seeded_oracle(seed=0, **kw)
    path = tmp_path / f"{name}.pth"
    torch.save({"state_dict": oracle.state_dict()}, path)
    model = orch.load_model(name, 2, path, torch.device("cpu"))
    assert not model.training
    sd = model.state_dict()
    for k, v in oracle.state_dict().items():
        assert torch.equal(sd[k], v), k
    # no CPU fallback: the reference silently falls back to CPU without CUDA (orchestrator.py:463-469); the backend refuses
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        model(torch.zeros(1, 3, 224, 224))


@needs_ref
def test_reference_cam_target_rule_on_backend_modules(reference):
    """web_ui._resolve_cam_target (:96-114) needs gradio/pytorch_grad_cam shims to import; run its rule on our modules."""
    ui_src = (REF / "web_ui.py").read_text()
    ns: dict = {}
    start, end = ui_src.index("def _find_last_conv_layer"), ui_src.index("def _add_label")
    exec("from torch import nn\n" + ui_src[start:end], ns)   # the two rule functions, verbatim from the checkout
    from deepfakedetection_b200 import registry

    def name_of(model, layer):
        return {id(m): n for n, m in model.named_modules()}[id(layer)]

    m = registry.get_model_spec("efficientnet_b3").builder("efficientnet_b3", 2)
    assert name_of(m, ns["_resolve_cam_target"](m)) == "_conv_head"
    m = registry.get_model_spec("efficientnet_b0").builder("efficientnet_b0", 2)
    assert name_of(m, ns["_resolve_cam_target"](m)) == "conv_head"
    m = registry.get_model_spec("efficientformerv2_s1").builder("efficientformerv2_s1", 2)
    assert name_of(m, ns["_resolve_cam_target"](m)) == "stages.3.blocks.5.mlp.fc2.conv" and m._cam_point() is ns["_resolve_cam_target"](m)
    m = registry.get_model_spec("faster_vit_2_224").builder("faster_vit_2_224", 2)
    assert name_of(m, ns["_resolve_cam_target"](m)) == "levels.2.global_tokenizer.pos_embed"


@needs_ref
def test_reference_eval_transforms(reference):
    from PIL import Image

    _, orch = reference
    x = orch.build_eval_transforms(224)(Image.new("L", (300, 260), 128))       # defaults: ensure_rgb ... val_normalize all on
    assert x.shape == (3, 224, 224) and x.dtype == torch.float32
    y = orch.build_eval_transforms(224, toggles={"val_normalize": "off"})(Image.new("RGB", (300, 260), (255, 0, 0)))
    assert float(y.max()) == 1.0 and float(y.min()) == 0.0


@needs_ref
@pytest.mark.gpu
def test_unmodified_inference_py_and_web_ui_on_the_backend():
    proc = subprocess.run([sys.executable, str(ROOT / "tools" / "run_reference_on_backend.py"), "--reference-root", str(REF),
                           "--ui", "--images", "16"], capture_output=True, text=True, timeout=1200)
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-3000:]
