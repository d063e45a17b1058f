"""Is the step GPU-bound or launch-bound?  (dev tool, GPU)  Enqueues K forwards and compares the host time spent
enqueueing with the device time between two events."""
import sys
import time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from bench import synth_state
from deepfakedetection_b200 import registry

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
K = 40
m = registry.get_model_spec("efficientnet_b0").builder("efficientnet_b0", 2)
synth_state(m)
m.to("cuda").eval()
x = torch.randn(n, 3, 224, 224, device="cuda")
with torch.inference_mode():
    for _ in range(5):
        m(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(K):
        m(x)
    e1.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
print(f"n={n}: device {e0.elapsed_time(e1) / K:.4f} ms/step, host enqueue {(t1 - t0) / K * 1e3:.4f} ms/step, "
      f"wall {(t2 - t0) / K * 1e3:.4f} ms/step")
if len(sys.argv) > 2 and sys.argv[2] == "prof":
    import cProfile
    import pstats
    pr = cProfile.Profile()
    with torch.inference_mode():
        pr.enable()
        for _ in range(K):
            m(x)
        pr.disable()
        torch.cuda.synchronize()
    pstats.Stats(pr).sort_stats("tottime").print_stats(12)
