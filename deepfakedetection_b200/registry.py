"""Registration of the B200 backend in the reference's model registry.

The reference's plug-in point is ``orchestration/model_registry.py``: a frozen
``ModelSpec(name, train_module, weights_key, default_image_size, builder)`` dataclass
(/root/reference/orchestration/model_registry.py:21-29) looked up by ``get_model_spec``
(:78-98) in two module-level dicts, ``_EXACT_SPECS`` (:50-58) and ``_PREFIX_SPECS`` (:60-75);
an unknown name raises ``KeyError("... Add it to model_registry.py.")`` (:98).

``register()`` adds exact entries (exact wins over prefix, :87-88) whose ``builder`` returns the
B200 engine module, so ``load_model`` (/root/reference/orchestration/orchestrator.py:350-377),
``inference.py`` and ``web_ui.py`` run unchanged.  Names served:

* ``efficientnet_b0``            -- bf16 tcgen05 path (BASELINE.json configs 1-3)
* ``efficientnet_b0.b200_fp32``  -- strict fp32 path (|dlogit| <= 1e-3 vs the timm fp32 forward)
* ``efficientnet_b3`` / ``efficientnet_b3.b200_fp32`` -- the registry's own default entry (efficientnet_pytorch
  layout, model_registry.py:32-36, 50-58), so ``config/inference.yaml`` and the release weights run as they are
* ``efficientformerv2_s1`` (timm layout, model_registry.py:39-40) and ``faster_vit_0_224`` / ``faster_vit_2_224``
  (fastervit layout, model_registry.py:43-47): exact entries that win over the registry's prefix table (:60-75, 87-88),
  so the other two models of ``config/inference.yaml`` run on the engine too; each also as ``<name>.b200_fp32``
"""

from __future__ import annotations

import importlib
from collections.abc import Callable
from dataclasses import dataclass
from typing import Any

from torch import nn


@dataclass(frozen=True)
class ModelSpec:
    """Field-for-field mirror of the reference's ModelSpec (model_registry.py:21-29); used
    only when the reference package itself is not importable (e.g. standalone benchmarking)."""

    name: str
    train_module: str
    weights_key: str
    default_image_size: int
    builder: Callable[[str, int], nn.Module]


def _precision_for(model_name: str) -> str:
    return "fp32" if model_name.endswith(".b200_fp32") else "bf16"


def build_efficientnet_b0(model_name: str, num_classes: int) -> nn.Module:
    """``builder(model_name, num_classes)`` exactly as the registry calls it (orchestrator.py:357)."""
    from .module import B200EfficientNetB0

    return B200EfficientNetB0(num_classes=num_classes, precision=_precision_for(model_name))


def build_efficientnet_b3(model_name: str, num_classes: int) -> nn.Module:
    """Replaces ``_build_efficientnet`` (model_registry.py:32-36): efficientnet_pytorch ``efficientnet-b3`` layout."""
    from .module import B200EfficientNetB3

    return B200EfficientNetB3(num_classes=num_classes, precision=_precision_for(model_name))


def build_efficientformerv2_s1(model_name: str, num_classes: int) -> nn.Module:
    """Replaces ``_build_efficientformer`` (model_registry.py:39-40) for timm ``efficientformerv2_s1``."""
    from .module_vit import B200EfficientFormerV2S1

    return B200EfficientFormerV2S1(num_classes=num_classes, precision=_precision_for(model_name))


def build_fastervit(model_name: str, num_classes: int) -> nn.Module:
    """Replaces ``_build_fastervit`` (model_registry.py:43-47): ``faster_vit_0_224`` / ``faster_vit_2_224`` + fresh head."""
    from .module_vit import B200FasterViT

    variant = model_name[: -len(".b200_fp32")] if model_name.endswith(".b200_fp32") else model_name
    return B200FasterViT(num_classes=num_classes, precision=_precision_for(model_name), variant=variant)


def _vit_entries() -> dict[str, dict[str, Any]]:
    out: dict[str, dict[str, Any]] = {}
    # same train_module / weights_key / image size the reference's prefix table yields (model_registry.py:60-75, 90-96)
    for name, train, builder in (
        ("efficientformerv2_s1", "trainers.efficientformer_v2", build_efficientformerv2_s1),
        ("faster_vit_0_224", "trainers.fastervit", build_fastervit),
        ("faster_vit_2_224", "trainers.fastervit", build_fastervit),
    ):
        for suffix in ("", ".b200_fp32"):
            out[name + suffix] = {"train_module": train, "weights_key": name, "default_image_size": 224, "builder": builder}
    return out


SERVED: dict[str, dict[str, Any]] = {
    # the reference's own exact entry (model_registry.py:50-58), same train_module / weights_key / image size
    "efficientnet_b3": {
        "train_module": "trainers.efficientnet",
        "weights_key": "efficientnet_b3",
        "default_image_size": 224,
        "builder": build_efficientnet_b3,
    },
    "efficientnet_b3.b200_fp32": {
        "train_module": "trainers.efficientnet",
        "weights_key": "efficientnet_b3",
        "default_image_size": 224,
        "builder": build_efficientnet_b3,
    },
    "efficientnet_b0": {
        "train_module": "trainers.efficientnet",
        "weights_key": "efficientnet_b0",
        "default_image_size": 224,
        "builder": build_efficientnet_b0,
    },
    "efficientnet_b0.b200_fp32": {
        "train_module": "trainers.efficientnet",
        "weights_key": "efficientnet_b0",
        "default_image_size": 224,
        "builder": build_efficientnet_b0,
    },
}

SERVED.update(_vit_entries())

# our own table, same lookup semantics as the reference's get_model_spec
_LOCAL_SPECS: dict[str, ModelSpec] = {
    name: ModelSpec(name=name, **fields) for name, fields in SERVED.items()
}


def get_model_spec(model_name: str) -> ModelSpec:
    """Standalone lookup with the reference's error behaviour (model_registry.py:98)."""
    if model_name in _LOCAL_SPECS:
        return _LOCAL_SPECS[model_name]
    raise KeyError(f"Unknown model '{model_name}'. Add it to model_registry.py.")


def register(registry_module: Any = None) -> list[str]:
    """Insert the B200-served specs into the reference registry's ``_EXACT_SPECS``.

    ``registry_module`` defaults to ``orchestration.model_registry`` (must be importable, i.e.
    the reference checkout is on ``sys.path``).  Returns the names registered.  Idempotent.
    """
    if registry_module is None:
        registry_module = importlib.import_module("orchestration.model_registry")
    spec_cls = registry_module.ModelSpec
    table = registry_module._EXACT_SPECS
    for name, fields in SERVED.items():
        table[name] = spec_cls(name=name, **fields)
    return list(SERVED)
