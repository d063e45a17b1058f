"""Generates tests/golden/effnet_b0_seed0.npz from the oracle (run once, commit the output).

The reference holds no golden vectors for this path (tests/test_repo_smoke.py:10-22) and its
third-party model code is not installable here, so these pins are produced by the oracle
(`oracle/effnet_b0.py`, itself pinned against torchvision's independent EfficientNet-B0).
They guard the oracle and the weight/input recipes against drift between machines.

    python tests/golden/make_golden.py
"""

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle.effnet_b0 import make_calibrated_oracle, synth_images  # noqa: E402
from oracle.gradcam import gradcam_autograd  # noqa: E402


def main() -> None:
    torch.set_num_threads(4)
    model = make_calibrated_oracle(seed=0)
    x = synth_images(4, seed=123)
    with torch.no_grad():
        logits = model(x)
        a = model.forward_features_raw(x)
    targets = logits.argmax(1)
    _, cam_low, cam = gradcam_autograd(model, x, targets)
    sd = model.state_dict()
    np.savez_compressed(
        ROOT / "tests" / "golden" / "effnet_b0_seed0.npz",
        logits=logits.numpy(),
        targets=targets.numpy(),
        a_mean=a.mean((2, 3)).numpy(),           # [4,1280] channel means of the raw conv_head output
        a_abs_sum=a.abs().sum((1, 2, 3)).numpy(),
        cam_low=cam_low.numpy(),                 # [4,7,7]
        cam_ds=cam[:, ::8, ::8].numpy(),         # [4,28,28] subsample of the 224x224 map
        x_checksum=x.double().sum((1, 2, 3)).numpy(),
        w_checksum=np.array([float(sd[k].double().abs().sum()) for k in sorted(sd) if "num_batches" not in k]),
    )
    print("wrote golden:", logits)


if __name__ == "__main__":
    main()
