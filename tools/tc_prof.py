"""Per-role wait-cycle table of every tcgen05 GEMM launch of one forward (dev tool, GPU).

Needs the instrumented library: `make prof` (libb200df_prof.so; never shipped).  For each launch prints, as a
median over CTAs: the role's lifetime in cycles and the share of it spent inside each mbarrier wait.
  producer: w0 = empty slot          mma: w0 = accumulator free, w1 = operands ready (full / scaled)
  epilogue: w0 = accumulator full, w1 = staging reusable (TMA store read), w2 = tcgen05.ld
  scaler  : w0 = operands landed, w1 = resident weights landed, w2 = scale body, w3 = proxy fence, w4 = arrive
  mma     : w2 = descriptor build + tcgen05.mma issue + commits + __syncwarp
"""
import ctypes as C
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
os.environ["B200DF_TC_PROF_PRINT"] = "1"
import numpy as np
import torch
from deepfakedetection_b200 import _lib

_lib.LIB_PATH = _lib._PKG / "libb200df_prof.so"
from bench import synth_state
from deepfakedetection_b200 import registry

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
m = registry.get_model_spec("efficientnet_b0").builder("efficientnet_b0", 2)
synth_state(m)
m.to("cuda").eval()
x = torch.randn(n, 3, 224, 224, device="cuda")
with torch.inference_mode():
    m(x)  # 32 GEMM launches -> slots 0..31
    m(x)  # slots 32..63 (warm)
torch.cuda.synchronize()
buf = np.zeros(64 * 148 * 32, dtype=np.uint64)
rc = _lib.lib().b200df_tc_prof_read(buf.ctypes.data_as(C.c_void_p))
assert rc == 0, rc
t = buf.reshape(64, 148, 4, 8).astype(np.float64)
names = ["prod", "mma", "epi", "scal"]
for slot in range(32, 64):
    row = [f"slot {slot:2d}"]
    for r in range(4):
        life = t[slot, :, r, 0]
        ok = life > 0
        if not ok.any():
            row.append(f"{names[r]}: -")
            continue
        L = np.median(life[ok])
        w = [np.median(t[slot, ok, r, 1 + i]) / L for i in range(5)]
        row.append(f"{names[r]} {L:7.0f}cy " + " ".join(f"{v:.2f}" for v in w))
    print(" | ".join(row))
