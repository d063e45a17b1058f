"""Bisect aid: bf16 EfficientNet-B3 logits under the debugging switches (B200DF_GEMM=simt, B200DF_DW=plain)."""
import os
import subprocess
import sys

CHILD = r'''
import torch
from deepfakedetection_b200 import registry
from oracle.effnet_b0 import synth_images
from oracle.effnet_b3 import make_seeded_oracle
ref_model = make_seeded_oracle(seed=0)
x = synth_images(32, seed=11)
with torch.no_grad():
    ref = ref_model(x)
m = registry.get_model_spec("efficientnet_b3").builder("efficientnet_b3", 2)
m.to("cuda").eval(); m.load_state_dict(ref_model.state_dict(), strict=False)
for n in (8, 5, 2, 3, 4, 16, 19, 32, 8):
    with torch.inference_mode():
        out = m(x[:n].cuda()).cpu()
    bad = (~torch.isfinite(out)).any(1)
    print("n", n, "nan rows", [i for i, b in enumerate(bad.tolist()) if b], "max err", (out - ref[:n])[~bad].abs().max().item() if (~bad).any() else None)
'''
for env in ({}, {"B200DF_GEMM": "simt"}, {"B200DF_DW": "plain"}, {"B200DF_GEMM": "simt", "B200DF_DW": "plain"},
            {"B200DF_PDL": "0"}):
    e = dict(os.environ, **env)
    r = subprocess.run([sys.executable, "-c", CHILD], env=e, capture_output=True, text=True)
    print(env, "\n" + r.stdout.strip()[-3000:], r.stderr.strip()[-300:] if r.returncode else "")
