"""One depthwise launch shape, a few repetitions (ncu target; dev tool, GPU).  usage: dw_one.py h c k s [N]"""
import ctypes as C
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from deepfakedetection_b200 import _lib

h, c, k, s = map(int, sys.argv[1:5])
N = int(sys.argv[5]) if len(sys.argv) > 5 else 256
lib = _lib.lib()
P = lambda t: C.c_void_p(t.data_ptr())
ho = (h + 2 * ((k - 1) // 2) - k) // s + 1
xs = [torch.randn(N, h, h, c, device="cuda").to(torch.bfloat16) for _ in range(3)]
out = torch.empty(N, ho, ho, c, device="cuda", dtype=torch.bfloat16)
w = torch.randn(k * k, c, device="cuda")
b = torch.randn(c, device="cuda")
pool = torch.zeros(N, c, dtype=torch.int64, device="cuda")
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for i in range(3):
    _lib.check(lib.b200df_k_dwconv(P(xs[i]), N, h, h, c, k, s, P(w), P(b), P(out), P(pool), _lib.DT_BF16, st))
torch.cuda.synchronize()
