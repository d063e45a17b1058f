"""Per-launch table of one forward (dev tool, GPU): name, ms, algorithmic GB/s, TFLOP/s."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from bench import synth_state
from deepfakedetection_b200 import registry

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
prec = sys.argv[2] if len(sys.argv) > 2 else "bf16"
arch = sys.argv[3] if len(sys.argv) > 3 else "efficientnet_b0"
name = arch if prec == "bf16" else arch + ".b200_fp32"
m = registry.get_model_spec(name).builder(name, 2)
synth_state(m)
m.to("cuda").eval()
x = torch.randn(n, 3, 224, 224, device="cuda")
with torch.inference_mode():
    for _ in range(3):
        m(x)
    eng = m._engine
    eng.set_profile(True)
    reps = 5
    for _ in range(reps):
        m.forward_cam(x, None)
    torch.cuda.synchronize()
    recs = eng.profile_records()
per = len(recs) // reps
tot = 0.0
print(f"{'launch':16s} {'ms':>8s} {'GB/s':>8s} {'TF/s':>7s} {'MB':>8s}")
for i in range(per):
    nm = recs[i][0]
    ms = sum(recs[i + r * per][1] for r in range(reps)) / reps
    by, fl = recs[i][2], recs[i][3]
    tot += ms
    print(f"{nm:16s} {ms:8.4f} {by / ms / 1e6:8.1f} {fl / ms / 1e9:7.2f} {by / 1e6:8.1f}")
print(f"total {tot:.4f} ms  -> {n / tot * 1e3:.0f} img/s (sum of per-launch times)")
