"""Prepare a synthetic ImageFolder + weights + YAML for deepfakedetection_b200.harness, or compare two metrics.jsonl files.

    python tools/harness_check.py prepare DIR [models]      # data/, weights/, inference.yaml under DIR
    python tools/harness_check.py compare A/metrics.jsonl B/metrics.jsonl
"""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def prepare(work: Path, models: list[str]) -> None:
    import torch
    import yaml

    import run_reference_on_backend as tool

    work.mkdir(parents=True, exist_ok=True)
    tool.make_dataset(work / "data", 40)
    (work / "weights").mkdir(exist_ok=True)
    cfg = {"seed": 1, "device": "cuda",
           "data": {"root": str(work / "data"), "test_split": "test", "val_split": "val", "num_classes": 2, "img_size": 224},
           "models": {}, "selection": models}
    for m in models:
        torch.save(tool.oracle_for(m).state_dict(), work / "weights" / f"{m}.pth")
        cfg["models"][m] = {"output_dir": str(work / "runs" / m),
                            "transforms": {"eval": {"ensure_rgb": True, "val_resize": True, "val_center_crop": True}},
                            "inference": {"weights": str(work / "weights" / f"{m}.pth"), "split": "test", "batch_size": 16,
                                          "num_workers": 0, "img_size": 224}}
    (work / "inference.yaml").write_text(yaml.safe_dump(cfg))


def compare(a: Path, b: Path) -> int:
    la = [json.loads(x) for x in a.read_text().strip().splitlines()]
    lb = [json.loads(x) for x in b.read_text().strip().splitlines()]
    bad = 0
    for x, y in zip(la, lb):
        same = all(x[k] == y[k] for k in ("model", "split", "accuracy", "roc_auc", "threshold", "confusion_matrix"))
        print(x["model"], "1 GPU:", x["accuracy"], x["roc_auc"], x["threshold"], x["confusion_matrix"], "| N GPUs:", y["accuracy"], y["roc_auc"],
              y["threshold"], y["confusion_matrix"], "IDENTICAL" if same else "DIFFERENT")
        bad += not same
    return bad


if __name__ == "__main__":
    if sys.argv[1] == "prepare":
        prepare(Path(sys.argv[2]), sys.argv[3].split(",") if len(sys.argv) > 3 else ["efficientnet_b0", "efficientformerv2_s1"])
    else:
        sys.exit(compare(Path(sys.argv[2]), Path(sys.argv[3])))
