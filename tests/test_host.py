"""Host-side logic of the drop-in boundary (no GPU needed)."""

import types
from dataclasses import dataclass, replace

import numpy as np
import pytest
import torch
from torch import nn

from deepfakedetection_b200 import registry
from deepfakedetection_b200.evaluate import best_threshold, compute_metrics, shard_range, unpack_records
from deepfakedetection_b200.gradcam import ClassifierOutputTarget, show_cam_on_image
from deepfakedetection_b200.module import B200EfficientNetB0


def fake_reference_registry():
    """Same shape as /root/reference/orchestration/model_registry.py (which cannot be imported
    here: it imports timm / efficientnet_pytorch / fastervit at module top, :15-17)."""

    @dataclass(frozen=True)
    class ModelSpec:
        name: str
        train_module: str
        weights_key: str
        default_image_size: int
        builder: object

    mod = types.ModuleType("orchestration.model_registry")
    mod.ModelSpec = ModelSpec
    mod._EXACT_SPECS = {"efficientnet_b3": ModelSpec("efficientnet_b3", "trainers.efficientnet", "efficientnet_b3", 224, None)}
    mod._PREFIX_SPECS = {"efficientformer": ModelSpec("efficientformerv2_s1", "trainers.efficientformer_v2", "efficientformerv2_s1", 224, None)}

    def get_model_spec(model_name):  # lookup order as model_registry.py:87-98
        if model_name in mod._EXACT_SPECS:
            return mod._EXACT_SPECS[model_name]
        for prefix, spec in mod._PREFIX_SPECS.items():
            if model_name.startswith(prefix):
                return replace(spec, name=model_name, weights_key=model_name)
        raise KeyError(f"Unknown model '{model_name}'. Add it to model_registry.py.")

    mod.get_model_spec = get_model_spec
    return mod


def test_register_adds_exact_specs_and_keeps_existing():
    ref = fake_reference_registry()
    with pytest.raises(KeyError):
        ref.get_model_spec("efficientnet_b0")  # the reference does not know B0
    names = registry.register(ref)
    assert "efficientnet_b0" in names
    spec = ref.get_model_spec("efficientnet_b0")
    assert isinstance(spec, ref.ModelSpec) and spec.default_image_size == 224
    assert ref.get_model_spec("efficientnet_b3").name == "efficientnet_b3"
    model = spec.builder("efficientnet_b0", 2)
    assert isinstance(model, nn.Module) and model.precision == "bf16"
    assert ref.get_model_spec("efficientnet_b0.b200_fp32").builder("efficientnet_b0.b200_fp32", 3).precision == "fp32"
    registry.register(ref)  # idempotent


def test_standalone_lookup_error_matches_reference_wording():
    assert registry.get_model_spec("efficientnet_b0").weights_key == "efficientnet_b0"
    with pytest.raises(KeyError, match="Add it to model_registry.py"):
        registry.get_model_spec("resnet50")


def test_module_follows_load_model_lifecycle(tmp_path):
    """builder -> .to(device) -> .eval() -> torch.load -> unwrap -> load_state_dict(strict=False)
    (orchestrator.py:356-375), with all three checkpoint shapes."""
    from oracle.effnet_b0 import make_seeded_oracle

    src = make_seeded_oracle(1).state_dict()
    for wrap in (None, "state_dict", "model"):
        model = registry.build_efficientnet_b0("efficientnet_b0", 2)
        model.to(torch.device("cpu"))
        model.eval()
        model._dirty = False
        path = tmp_path / f"w_{wrap}.pth"
        torch.save(src if wrap is None else {wrap: src, "epoch": 3}, path)
        state = torch.load(path, map_location="cpu")
        if isinstance(state, dict) and "state_dict" in state:
            state = state["state_dict"]
        elif isinstance(state, dict) and "model" in state:
            state = state["model"]
        res = model.load_state_dict(state, strict=False)
        assert not res.missing_keys and not res.unexpected_keys
        assert model._dirty, "load_state_dict must mark the packed weights stale"
        for k, v in model.state_dict().items():
            assert torch.equal(v, src[k]), k


def test_strict_false_mismatches_are_reported_loudly():
    model = B200EfficientNetB0(2)
    sd = model.state_dict()
    sd.pop("classifier.bias")
    sd["extra.weight"] = torch.zeros(1)
    with pytest.warns(UserWarning, match="missing 1 tensors"):
        model.load_state_dict(sd, strict=False)


def test_cam_target_rule_finds_conv_head():
    """web_ui.py:96-114: `_conv_head` attr if it is an nn.Module, else the last nn.Conv2d."""
    model = B200EfficientNetB0(2)
    conv_head = getattr(model, "_conv_head", None)
    assert isinstance(conv_head, nn.Module) and conv_head is model.conv_head
    last = None
    for m in model.modules():
        if isinstance(m, nn.Conv2d):
            last = m
    assert last is model.conv_head
    assert "_conv_head.weight" not in model.state_dict()


def test_cpu_input_raises_no_fallback():
    model = B200EfficientNetB0(2).eval()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        model(torch.zeros(1, 3, 224, 224))
    with pytest.raises(ValueError, match="224"):
        model(torch.zeros(1, 3, 256, 256))
    with pytest.raises(RuntimeError, match="parameter container"):
        model.conv_stem(torch.zeros(1, 3, 8, 8))


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 256, 1001):
        for w in (1, 2, 4, 8):
            spans = [shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_metrics_match_reference_formulas():
    from sklearn.metrics import balanced_accuracy_score

    g = torch.Generator().manual_seed(0)
    probs1 = torch.rand(200, generator=g)
    targets = (torch.rand(200, generator=g) < probs1).long()
    probs = torch.stack([1 - probs1, probs1], 1)
    rec = torch.cat([probs, probs.argmax(1, keepdim=True).float(), targets[:, None].float()], 1)
    p, pr, t = unpack_records(rec)
    assert torch.equal(pr, probs.argmax(1)) and torch.equal(t, targets)
    m = compute_metrics(p, pr, t, 2, threshold=0.4)
    assert m["accuracy"] == pytest.approx(((probs1 >= 0.4).long() == targets).float().mean().item())
    assert np.array(m["confusion_matrix"]).sum() == 200 and 0.5 < m["roc_auc"] <= 1.0
    # the 501-point sweep of orchestrator.py:536-546
    scores, truth = probs1.numpy(), targets.numpy()
    best, chosen = -1.0, 0.5
    for thr in np.linspace(0.0, 1.0, 501, dtype=np.float64):
        bal = balanced_accuracy_score(truth, (scores >= thr).astype(np.int64))
        if bal > best:
            best, chosen = float(bal), float(thr)
    assert best_threshold(scores, truth) == pytest.approx(chosen)


def test_show_cam_on_image_and_target():
    img = np.full((8, 8, 3), 0.5, np.float32)
    mask = np.linspace(0, 1, 64, dtype=np.float32).reshape(8, 8)
    out = show_cam_on_image(img, mask, use_rgb=True)
    assert out.dtype == np.uint8 and out.shape == (8, 8, 3) and out.max() == 255
    assert out[0, 0, 2] > out[0, 0, 0] and out[-1, -1, 0] > out[-1, -1, 2]  # JET: blue -> red in RGB order
    with pytest.raises(ValueError):
        show_cam_on_image(img * 3, mask)
    assert float(ClassifierOutputTarget(1)(torch.tensor([[1.0, 2.0]]))) == 2.0


def _reference_threshold_loop(scores, truth):
    """orchestrator.py:536-546 verbatim (sklearn in the loop)."""
    import numpy as np
    from sklearn.metrics import balanced_accuracy_score

    thresholds = np.linspace(0.0, 1.0, 501, dtype=np.float64)
    best_bal, chosen = -1.0, 0.5
    for thr in thresholds:
        preds = (scores >= thr).astype(np.int64)
        bal = balanced_accuracy_score(truth, preds)
        if bal > best_bal:
            best_bal, chosen = float(bal), float(thr)
    return float(chosen)


def _hist_like_kernel(scores, truth, thr):
    import numpy as np

    j = (scores.astype(np.float64)[:, None] >= thr[None, :]).sum(1)   # thresholds ascending -> prefix property
    hist = np.zeros((2, thr.size + 1), dtype=np.int64)
    for c in (0, 1):
        np.add.at(hist[c], j[(truth == 1) == bool(c)], 1)
    return hist


@pytest.mark.parametrize("case", ["random", "ties", "separable", "skewed"])
def test_threshold_from_hist_matches_reference_sweep(case):
    """Host half of the on-device threshold sweep: histograms -> same threshold as 501 sklearn calls."""
    import warnings

    import numpy as np

    from deepfakedetection_b200.evaluate import best_threshold, threshold_from_hist

    rng = np.random.default_rng({"random": 1, "ties": 2, "separable": 3, "skewed": 4}[case])
    n = 700
    truth = (rng.random(n) < (0.1 if case == "skewed" else 0.5)).astype(np.int64)
    scores = rng.random(n).astype(np.float32)
    if case == "ties":
        scores = (np.round(scores * 20) / 20).astype(np.float32)     # many scores exactly on thresholds
    if case == "separable":
        scores = np.where(truth == 1, 0.7 + 0.3 * scores, 0.3 * scores).astype(np.float32)
    thr = np.linspace(0.0, 1.0, 501, dtype=np.float64)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        want = _reference_threshold_loop(scores, truth)
    assert threshold_from_hist(_hist_like_kernel(scores, truth, thr), thr) == want
    assert best_threshold(scores, truth) == want
    # one class only: the reference skips the sweep and keeps 0.5
    assert threshold_from_hist(_hist_like_kernel(scores, np.zeros(n, dtype=np.int64), thr), thr) == 0.5
