"""Every tuned / pinned depthwise plan must give bit-identical outputs (only the pooled sums may differ in the
last bits: different partial-sum grouping).  dev tool, GPU."""
import ctypes as C
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from deepfakedetection_b200 import _lib

lib = _lib.lib()
P = lambda t: C.c_void_p(t.data_ptr())
CASES = [(112, 32, 3, 1, "7,16,4,1,2"), (112, 96, 3, 2, "7,14,1,1,1"), (56, 144, 3, 1, "7,8,8,1,2"),
         (56, 144, 5, 2, "7,4,4,1,1"), (28, 240, 5, 1, "7,4,4,1,2"), (14, 672, 5, 2, "7,8,1,1,1"),
         (7, 1152, 5, 1, "7,8,1,4,2"), (7, 1152, 3, 1, "7,8,1,4,2"), (7, 1152, 5, 1, "7,8,1,2,2"),
         (14, 480, 5, 1, "7,16,1,1,2"), (14, 480, 3, 1, "7,16,2,2,2"), (28, 240, 3, 2, "7,8,1,1,1")]
N = 6
for (h, c, k, s, pin) in CASES:
    ho = (h + 2 * ((k - 1) // 2) - k) // s + 1
    torch.manual_seed(1)
    x = torch.randn(N, h, h, c, device="cuda").to(torch.bfloat16)
    w = torch.randn(k * k, c, device="cuda") * 0.2
    b = torch.randn(c, device="cuda") * 0.1
    res = []
    for plan in ("8,2,1,1,1" if ho % 8 == 0 else "7,2,1,1,1", pin):
        os.environ["B200DF_DW_PLAN"] = plan
        out = torch.zeros(N, ho, ho, c, device="cuda", dtype=torch.bfloat16)
        pool = torch.zeros(N, c, dtype=torch.int64, device="cuda")
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        _lib.check(lib.b200df_k_dwconv(P(x), N, h, h, c, k, s, P(w), P(b), P(out), P(pool), _lib.DT_BF16, st))
        torch.cuda.synchronize()
        res.append((out.clone(), pool.clone()))
    same = torch.equal(res[0][0], res[1][0])
    dp = ((res[0][1] - res[1][1]).abs().double() / 2**24).max().item()
    ref = (res[0][1].abs().double() / 2**24).max().item()
    print(f"{h}x{c} k{k} s{s} pin {pin}: outputs identical={same}  max |dpool|={dp:.3e} (max |pool| {ref:.3e})")
