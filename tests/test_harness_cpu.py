"""Host logic of the multi-GPU dataset harness (deepfakedetection_b200/harness.py): YAML toggles -> input plan, the uint8
transform pipeline vs the reference's float pipeline, and a world_size-2 gloo run of the sharded job with a CPU stub model
(the sharding / gather / threshold / metrics code; the model call itself needs a B200)."""

import json
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from PIL import Image

from deepfakedetection_b200 import harness


def test_toggles_map_to_the_fused_input_path():
    assert harness.input_plan(None) == {"mean": harness.IMAGENET_MEAN, "std": harness.IMAGENET_STD}
    assert harness.input_plan({"val_normalize": "off"}) == {"mean": None, "std": None}
    assert harness.input_plan({"val_normalize": 1, "ensure_rgb": "no"})["mean"] == harness.IMAGENET_MEAN
    with pytest.raises(ValueError, match="val_to_tensor"):
        harness.input_plan({"val_to_tensor": False})
    cfg = {"transforms": {"eval": {"val_normalize": False}}}
    assert harness.resolve_transform_mapping(cfg) == {"val_normalize": False}
    assert harness.resolve_transform_mapping({"transforms": {"val_resize": "true"}}) == {"val_resize": "true"}
    assert harness.resolve_transform_mapping({"inference": {"transforms": {"val_center_crop": 0}}}) == {"val_center_crop": 0}
    assert harness.resolve_transform_mapping({}) is None


def test_u8_pipeline_equals_the_float_pipeline_before_totensor():
    """uint8 HWC crop -> (/255, Normalize) on the host == torchvision's ToTensor + Normalize of the same PIL crop, i.e. what
    forward_u8 fuses is exactly the tail of orchestrator.build_eval_transforms (:339-345)."""
    from torchvision import transforms

    g = np.random.default_rng(0)
    img = Image.fromarray(g.integers(0, 256, (300, 260, 3), dtype=np.uint8))
    u8 = harness.build_u8_eval_transforms(224, None)(img)
    assert u8.shape == (224, 224, 3) and u8.dtype == torch.uint8
    ref = transforms.Compose([transforms.Resize(224), transforms.CenterCrop(224), transforms.ToTensor(),
                              transforms.Normalize(mean=list(harness.IMAGENET_MEAN), std=list(harness.IMAGENET_STD))])(img)
    mine = (u8.float() / 255.0 - torch.tensor(harness.IMAGENET_MEAN)) / torch.tensor(harness.IMAGENET_STD)
    assert torch.allclose(mine.permute(2, 0, 1), ref, atol=1e-6)
    gray = harness.build_u8_eval_transforms(224, None)(Image.new("L", (250, 250), 7))
    assert gray.shape == (224, 224, 3)


def _make_dataset(root, n_per_class=7):
    g = np.random.default_rng(1)
    for split in ("val", "test"):
        for ci, cls in enumerate(("fake", "real")):
            d = root / split / cls
            d.mkdir(parents=True)
            for i in range(n_per_class + ci):   # ragged: 7 + 8 images per split
                arr = np.clip(g.normal(90 + 70 * ci, 40, (64, 64, 3)), 0, 255).astype(np.uint8)
                Image.fromarray(arr).save(d / f"{i}.png")


def _stub_model(u8: torch.Tensor) -> torch.Tensor:
    """A deterministic 'model' on the uint8 crops: brighter image -> class 1."""
    m = u8.float().mean((1, 2, 3)) / 255.0
    return torch.stack([-(m - 0.45) * 20, (m - 0.45) * 20], 1)


def _config(root):
    return {"data": {"root": str(root), "val_split": "val", "test_split": "test", "num_classes": 2, "img_size": 32},
            "models": {"efficientnet_b0": {"inference": {"batch_size": 4, "num_workers": 0, "img_size": 32}}}}


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, root, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from pathlib import Path

        m = harness.run_inference_job_sharded(_config(root), "efficientnet_b0", torch.device("cpu"), run_dir=Path(out), model=_stub_model)
        assert (m is None) == (rank != 0)
    finally:
        dist.destroy_process_group()


def test_sharded_job_world2_equals_single_process(tmp_path):
    root = tmp_path / "data"
    _make_dataset(root)
    single = harness.run_inference_job_sharded(_config(root), "efficientnet_b0", torch.device("cpu"), run_dir=tmp_path / "one",
                                               model=_stub_model)
    mp.spawn(_worker, args=(2, _free_port(), str(root), str(tmp_path / "two")), nprocs=2, join=True)
    two = json.loads((tmp_path / "two" / "logs" / "metrics.jsonl").read_text().strip())
    for k in ("model", "split", "accuracy", "roc_auc", "threshold", "confusion_matrix"):   # the reference's keys (:610-647)
        assert two[k] == single[k], k
    assert single["accuracy"] > 0.8 and np.array(single["confusion_matrix"]).sum() == 15
