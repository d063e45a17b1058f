"""Multi-GPU evaluation harness for real datasets: the reference's ``_run_inference_job``
(/root/reference/orchestration/orchestrator.py:437-658) run per rank on a shard of the ImageFolder, one process per GPU.

    torchrun --nnodes=1 --nproc-per-node N -m deepfakedetection_b200.harness --config config/inference.yaml

Same YAML, same semantics (weights path relative to CWD, ``inference.{split,batch_size,num_workers,img_size}``, the
binary validation-split threshold sweep :513-546, ``logs/metrics.jsonl`` with the same keys :610-647), with three changes
that are the point of the harness:

* **sharding** (SURVEY.md 8e mode ii): rank r evaluates the contiguous slice ``evaluate.shard_range(len(dataset), r, G)`` of
  the (unshuffled) ImageFolder order with replicated weights; the only exchange is ONE all-gather of the packed
  ``{probs, pred, label}`` records per split (NCCL on GPUs, gloo in the CPU tests), after which rank 0 sees exactly
  the tensors the single-process loop concatenates (:601-603) and computes / writes the metrics;
* **the fused input path is gated by the reference's YAML toggles** (orchestrator.py:316-347):
  ``val_to_tensor`` / ``val_normalize`` decide what the stem kernel fuses -- with ``val_to_tensor: true`` the loader
  hands over uint8 HWC crops (4x fewer host->device bytes than fp32 NCHW) and ToTensor (/255) + Normalize(mean, std)
  (``val_normalize: true``) or ToTensor alone (``false``) run inside ``forward_u8``; ``ensure_rgb`` / ``val_resize`` /
  ``val_center_crop`` keep their PIL meaning;
* the metrics tail can run on the device (``evaluate.compute_metrics_device``).

The plots (matplotlib) and the rich progress bar are the reference's UI and are not reproduced.
"""

from __future__ import annotations

import argparse
import json
import os
from datetime import datetime
from pathlib import Path
from typing import Any, Callable

import numpy as np
import torch
import torch.distributed as dist

from . import evaluate, registry

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)
_DEFAULT_TOGGLES = {"ensure_rgb": True, "val_resize": True, "val_center_crop": True, "val_to_tensor": True, "val_normalize": True}


def coerce_bool(value: Any) -> bool:
    """orchestrator.py:128-135."""
    if isinstance(value, bool):
        return value
    if isinstance(value, (int, float)):
        return bool(value)
    if isinstance(value, str):
        return value.strip().lower() in {"1", "true", "yes", "on"}
    return False


def resolve_transform_mapping(model_cfg: dict[str, Any], phase: str = "eval") -> dict[str, Any] | None:
    """orchestrator.py:162-180: `transforms.{train,eval}` mapping, a flat mapping of scalars, or `inference.transforms`."""
    tcfg = model_cfg.get("transforms")
    if isinstance(tcfg, dict):
        ph = tcfg.get(phase)
        if isinstance(ph, dict):
            return ph
        if all(isinstance(v, (bool, int, float, str)) for v in tcfg.values()):
            return tcfg
    scoped = (model_cfg.get("training" if phase == "train" else "inference") or {}).get("transforms")
    return scoped if isinstance(scoped, dict) else None


def resolve_toggles(toggles: dict[str, Any] | None) -> dict[str, bool]:
    out = dict(_DEFAULT_TOGGLES)
    for k, v in (toggles or {}).items():
        out[k] = coerce_bool(v)
    return out


def input_plan(toggles: dict[str, Any] | None) -> dict[str, Any]:
    """What the YAML toggles mean for the engine's input path: {'mean': 3 floats | None, 'std': ...} for ``forward_u8``."""
    t = resolve_toggles(toggles)
    if not t["val_to_tensor"]:
        raise ValueError("`val_to_tensor: false` hands PIL images to the DataLoader's collate in the reference as well "
                         "(orchestrator.py:339-341): no tensor pipeline exists for it")
    return {"mean": IMAGENET_MEAN if t["val_normalize"] else None, "std": IMAGENET_STD if t["val_normalize"] else None}


def build_u8_eval_transforms(image_size: int, toggles: dict[str, Any] | None = None):
    """The PIL half of ``build_eval_transforms`` (orchestrator.py:316-347), ending in a uint8 HWC tensor: ToTensor and
    Normalize are what ``forward_u8`` fuses into the stem kernel."""
    from torchvision import transforms

    t = resolve_toggles(toggles)
    ops: list[Any] = []
    if t["ensure_rgb"]:
        ops.append(transforms.Lambda(lambda im: im.convert("RGB") if im.mode != "RGB" else im))
    if t["val_resize"]:
        ops.append(transforms.Resize(image_size))
    if t["val_center_crop"]:
        ops.append(transforms.CenterCrop(image_size))
    ops.append(transforms.Lambda(lambda im: torch.from_numpy(np.asarray(im, dtype=np.uint8).copy())))
    return transforms.Compose(ops)


def _world() -> tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def evaluate_split(forward_u8: Callable[[torch.Tensor], torch.Tensor], dataset, batch_size: int, num_workers: int,
                   device: torch.device, num_classes: int) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """This rank's shard of `dataset` through the model, then the one all-gather -> (probs, preds, targets) of the WHOLE
    split in dataset order on every rank (what orchestrator.py:601-603 concatenates)."""
    from torch.utils.data import DataLoader, Subset

    rank, world = _world()
    lo, hi = evaluate.shard_range(len(dataset), rank, world)
    loader = DataLoader(Subset(dataset, range(lo, hi)), batch_size=batch_size, shuffle=False, num_workers=num_workers,
                        pin_memory=device.type == "cuda", persistent_workers=False)
    recs = []
    with torch.inference_mode():
        for images, targets in loader:
            logits = forward_u8(images.to(device, non_blocking=True))
            if logits.is_cuda:
                recs.append(evaluate.pack_records(logits, targets.to(device)))
            else:   # CPU stub models in the gloo tests
                p = torch.softmax(logits.float(), 1)
                recs.append(torch.cat([p, p.argmax(1, keepdim=True).float(), targets.view(-1, 1).float()], 1))
    local = torch.cat(recs, 0) if recs else torch.zeros((0, num_classes + 2), dtype=torch.float32, device=device)
    return evaluate.unpack_records(evaluate.gather_records(local))


def run_inference_job_sharded(config: dict[str, Any], model_name: str, device: torch.device, *, run_dir: Path | None = None,
                              model: Callable | None = None) -> dict[str, Any] | None:
    """One model of the YAML on this rank's shard; returns the metrics dict on rank 0 (None elsewhere)."""
    from torchvision import datasets

    rank, _ = _world()
    model_cfg = {"name": model_name, **config["models"][model_name]}
    data_cfg = config.get("data", {})
    infer_cfg = model_cfg.get("inference") or {}
    spec = registry.get_model_spec(model_name)
    num_classes = int(model_cfg.get("num_classes", data_cfg.get("num_classes", 2)))
    image_size = int(infer_cfg.get("img_size") or data_cfg.get("img_size", spec.default_image_size))
    batch_size = int(infer_cfg.get("batch_size", 64))
    num_workers = int(infer_cfg.get("num_workers", 4))
    split = infer_cfg.get("split") or data_cfg.get("test_split", "test")
    toggles = resolve_transform_mapping(model_cfg, "eval")
    plan = input_plan(toggles)
    tf = build_u8_eval_transforms(image_size, toggles)
    if model is None:
        net = spec.builder(model_name, num_classes)
        net.to(device).eval()
        wp = infer_cfg.get("weights")
        if wp:
            path = Path(wp).expanduser()
            path = path if path.is_absolute() else (Path.cwd() / path).resolve()
            if not path.exists():
                raise SystemExit(f"Weights not found: {path}")       # orchestrator.py:361-364
            state = torch.load(path, map_location=device)
            if isinstance(state, dict) and "state_dict" in state:
                state = state["state_dict"]
            elif isinstance(state, dict) and "model" in state:
                state = state["model"]
            net.load_state_dict(state, strict=False)

        def model(u8: torch.Tensor) -> torch.Tensor:
            return net.forward_u8(u8, plan["mean"], plan["std"])
    root = Path(data_cfg["root"]).expanduser()
    root = root if root.is_absolute() else (Path.cwd() / root).resolve()

    threshold = 0.5
    if num_classes == 2 and (root / data_cfg.get("val_split", "val")).exists():
        val = datasets.ImageFolder(root / data_cfg.get("val_split", "val"), transform=tf)
        if len(val) > 0:
            probs, _preds, truth = evaluate_split(model, val, batch_size, num_workers, device, num_classes)
            threshold = evaluate.best_threshold(probs[:, 1].numpy(), truth.numpy())
    path = root / split
    if not path.exists():
        raise SystemExit(f"Split not found: {path}")                 # orchestrator.py:549-552
    test = datasets.ImageFolder(path, transform=tf)
    if len(test.classes) != num_classes:
        raise SystemExit(f"{path}: {len(test.classes)} class folders, config says num_classes={num_classes}")
    probs, preds, truth = evaluate_split(model, test, batch_size, num_workers, device, num_classes)
    if rank != 0:
        return None
    metrics: dict[str, Any] = {"model": model_name, "split": split}
    metrics.update(evaluate.compute_metrics(probs, preds, truth, num_classes, threshold=threshold if num_classes == 2 else None))
    metrics["timestamp"] = datetime.now().isoformat()
    if run_dir is not None:
        (run_dir / "logs").mkdir(parents=True, exist_ok=True)
        with (run_dir / "logs" / "metrics.jsonl").open("a", encoding="utf-8") as fh:
            fh.write(json.dumps(metrics) + "\n")
    return metrics


def main() -> None:
    import yaml

    ap = argparse.ArgumentParser(description="batch-sharded evaluation of an orchestrator inference YAML on N B200s")
    ap.add_argument("--config", type=Path, default=Path("config/inference.yaml"))
    args = ap.parse_args()
    config = yaml.safe_load(args.config.read_text())
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("deepfakedetection_b200.harness: no CUDA device -- the b200 backend has no CPU fallback")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    stamp = datetime.now().strftime("%Y%m%d-%H%M%S")
    for name in config.get("selection") or list(config["models"]):
        out = Path((config["models"][name] or {}).get("output_dir") or f"runs/{name}") / stamp
        m = run_inference_job_sharded(config, name, device, run_dir=out)
        if m is not None:
            print(json.dumps(m))
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
