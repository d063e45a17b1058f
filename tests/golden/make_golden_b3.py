"""Generates tests/golden/effnet_b3_seed0.npz from the B3 oracle (run once, commit the output).

Same rationale as make_golden.py: the reference holds no vectors for this path and efficientnet_pytorch is not
installable here, so the pins come from `oracle/effnet_b3.py` (itself pinned against torchvision's
`efficientnet_b3`, tests/test_oracle_b3.py) on the BASELINE.json weight recipe (default init + randomised BN).
They guard the oracle, the TF-"same" pads and the recipes against drift, and the GPU engine is compared with them
in tests/test_gpu_b3.py::test_b3_golden_vectors.

    python tests/golden/make_golden_b3.py
"""

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle.effnet_b0 import synth_images  # noqa: E402
from oracle.effnet_b3 import make_seeded_oracle  # noqa: E402
from oracle.gradcam import gradcam_autograd  # noqa: E402


def main() -> None:
    torch.set_num_threads(4)
    model = make_seeded_oracle(seed=0)
    x = synth_images(4, seed=123)
    with torch.no_grad():
        logits = model(x)
        a = model.forward_features_raw(x)
    targets = logits.argmax(1)
    _, cam_low, cam = gradcam_autograd(model, x, targets)
    sd = model.state_dict()
    np.savez_compressed(
        ROOT / "tests" / "golden" / "effnet_b3_seed0.npz",
        logits=logits.numpy(),
        targets=targets.numpy(),
        a_mean=a.mean((2, 3)).numpy(),           # [4,1536] channel means of the raw _conv_head output
        cam_low=cam_low.numpy(),                 # [4,7,7]
        cam_ds=cam[:, ::8, ::8].numpy(),         # [4,28,28]
        x_checksum=x.double().sum((1, 2, 3)).numpy(),
        w_checksum=np.array([float(sd[k].double().abs().sum()) for k in sorted(sd) if "num_batches" not in k]),
    )
    print("wrote golden:", logits)


if __name__ == "__main__":
    main()
