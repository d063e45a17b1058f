"""Generates tests/golden/efficientformerv2_s1_seed0.npz and tests/golden/faster_vit_0_224_seed0.npz from the oracles
(run once, commit the output).  REGRESSION SNAPSHOTS, not independent pins: the reference holds no golden vectors for
this path (tests/test_repo_smoke.py:10-22) and timm / fastervit are not installable here, so these files only guard the
oracles and the weight/input recipes against drift between machines and torch versions.

    python tests/golden/make_golden_vit.py
"""

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle import efficientformerv2, fastervit  # noqa: E402
from oracle.effnet_b0 import synth_images  # noqa: E402


def snapshot(model, path: Path) -> None:
    x = synth_images(4, seed=123)
    with torch.no_grad():
        logits = model(x)
    sd = model.state_dict()
    np.savez_compressed(
        path,
        logits=logits.numpy(),
        x_checksum=x.double().sum((1, 2, 3)).numpy(),
        w_checksum=np.array([float(sd[k].double().abs().sum()) for k in sorted(sd) if "num_batches" not in k]),
    )
    print(path.name, logits)


def main() -> None:
    torch.set_num_threads(4)
    g = ROOT / "tests" / "golden"
    snapshot(efficientformerv2.make_calibrated_oracle(seed=0), g / "efficientformerv2_s1_seed0.npz")
    snapshot(fastervit.make_calibrated_oracle(seed=0), g / "faster_vit_0_224_seed0.npz")


if __name__ == "__main__":
    main()
