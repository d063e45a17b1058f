"""GPU debug for the tensor-core depthwise kernel: one-hot taps, compare with shifted inputs."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, torch.nn.functional as F
from gpu_util import k_dwconv
from deepfakedetection_b200 import _lib
BF = _lib.DT_BF16
torch.manual_seed(0)
c, h, k = int(sys.argv[1]) if len(sys.argv) > 1 else 64, int(sys.argv[2]) if len(sys.argv) > 2 else 16, int(sys.argv[3]) if len(sys.argv) > 3 else 3
n = 1
x = torch.randn(n, h, h, c, device="cuda").bfloat16()
b = torch.zeros(c, device="cuda")
pad = (k - 1) // 2
for ky in range(k):
    for kx in range(k):
        w = torch.zeros(c, 1, k, k, device="cuda"); w[:, 0, ky, kx] = 1.0
        ref = F.silu(F.conv2d(x.float().permute(0, 3, 1, 2), w, b, padding=pad, groups=c)).permute(0, 2, 3, 1)
        out, pool = k_dwconv(x, w.reshape(c, k * k).t().contiguous(), b, k, 1, BF)
        torch.cuda.synchronize()
        err = (out.float() - ref).abs()
        # per-channel-group and per-column error
        eg = err.reshape(n, h, h, c // 16, 16).amax((0, 1, 2, 4)).tolist()
        ecol = err.amax((0, 1, 3)).tolist()
        erow = err.amax((0, 2, 3)).tolist()
        print(f"tap({ky},{kx}) max err {err.max().item():.3f} per-16ch-group {[round(v,2) for v in eg]} per-x {[round(v,1) for v in ecol[:16]]} per-y {[round(v,1) for v in erow[:16]]}")
        if err.max() > 0.05:
            # which shift does the output correspond to?
            xs = F.pad(x.float().permute(0, 3, 1, 2), (4, 4, 4, 4))
            best = None
            for dy in range(-3, 4):
                for dx in range(-3, 4):
                    cand = F.silu(xs[:, :, 4 + dy:4 + dy + h, 4 + dx:4 + dx + h]).permute(0, 2, 3, 1)
                    e = (out.float() - cand).abs().mean().item()
                    if best is None or e < best[0]:
                        best = (e, dy, dx)
            print("   best matching shift (dy,dx) =", best[1:], "mean err", round(best[0], 4), "expected", (ky - pad, kx - pad))
