"""Times every admissible depthwise tile plan (B200DF_DW_PLAN=th,otw,m,g,nc) for the 16 EfficientNet-B0 depthwise
shapes at batch N (dev tool, GPU).  Inputs rotate over > 300 MB of buffers so the 126 MB L2 cannot hold them."""
import ctypes as C
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from deepfakedetection_b200 import _lib

N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
only = sys.argv[2] if len(sys.argv) > 2 else None
# (h_in, channels, k, stride)
LAYERS = [(112, 32, 3, 1), (112, 96, 3, 2), (56, 144, 3, 1), (56, 144, 5, 2), (28, 240, 5, 1), (28, 240, 3, 2),
          (14, 480, 3, 1), (14, 480, 5, 1), (14, 672, 5, 1), (14, 672, 5, 2), (7, 1152, 5, 1), (7, 1152, 3, 1)]
if os.environ.get("DW_SWEEP_MODEL") == "b3":  # the EfficientNet-B3 (efficientnet_pytorch) depthwise shapes
    LAYERS = [(112, 40, 3, 1), (112, 24, 3, 1), (112, 144, 3, 2), (56, 192, 3, 1), (56, 192, 5, 2), (28, 288, 5, 1),
              (28, 288, 3, 2), (14, 576, 3, 1), (14, 576, 5, 1), (14, 816, 5, 1), (14, 816, 5, 2), (7, 1392, 5, 1),
              (7, 1392, 3, 1), (7, 2304, 3, 1)]
lib = _lib.lib()
P = lambda t: C.c_void_p(t.data_ptr())
dev = "cuda"


def run(xs, w, b, outs, pool, k, s, reps):
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    n, h, wd, c = xs[0].shape
    for i in range(reps):
        j = i % len(xs)
        _lib.check(lib.b200df_k_dwconv(P(xs[j]), n, h, wd, c, k, s, P(w), P(b), P(outs[j]), P(pool), _lib.DT_BF16, st))


for (h, c, k, s) in LAYERS:
    tag = f"{h}x{c}k{k}s{s}"
    if only and only != "default" and tag not in only.split(","):
        continue
    ho = (h + 2 * ((k - 1) // 2) - k) // s + 1
    nb = max(2, int(300e6 // (N * h * h * c * 2)) + 1)
    xs = [torch.randn(N, h, h, c, device=dev).to(torch.bfloat16) for _ in range(nb)]
    outs = [torch.empty(N, ho, ho, c, device=dev, dtype=torch.bfloat16) for _ in range(nb)]
    w = torch.randn(k * k, c, device=dev)
    b = torch.randn(c, device=dev)
    pool = torch.zeros(N, c, dtype=torch.int64, device=dev)
    res = []
    cands = [None]
    for nc, th in [] if only == "default" else [(a, t) for a in ((1, 2) if s == 1 else (1,)) for t in (8, 7)]:
        if nc == 2 and k == 5 and th == 8:
            continue
        strips = (ho + th - 1) // th
        for otw in range(2 * nc, 34, 2 * nc):
            tiles_x = (ho + otw - 1) // otw
            if tiles_x * otw - ho > 3:
                continue
            for m in range(1, strips + 1):
                if strips % m:
                    continue
                whole = tiles_x == 1 and m == strips
                for g in ((1, 2, 4, 8) if whole else (1,)):
                    cands.append((th, otw, m, g, nc))
    for cand in cands:
        if cand is None:
            os.environ.pop("B200DF_DW_PLAN", None)
        else:
            os.environ["B200DF_DW_PLAN"] = ",".join(map(str, cand))
        try:
            run(xs, w, b, outs, pool, k, s, 3)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 12
            e0.record()
            run(xs, w, b, outs, pool, k, s, reps)
            e1.record()
            torch.cuda.synchronize()
            res.append((e0.elapsed_time(e1) / reps, cand))
        except Exception as ex:  # noqa: BLE001
            print(tag, cand, "failed", ex)
            break
    base = res[0][0]
    res.sort(key=lambda r: r[0])
    by = N * (h * h + ho * ho) * c * 2
    print(f"{tag}: default {base:.4f} ms ({by / base / 1e6:.0f} GB/s) | best " +
          " ".join(f"{cd}:{t:.4f}" for t, cd in res[:5]), flush=True)
    del xs, outs
    torch.cuda.empty_cache()
