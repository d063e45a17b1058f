"""One forward step bracketed by cudaProfilerStart/Stop (run under `ncu --profile-from-start off`)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from bench import synth_state
from deepfakedetection_b200 import registry

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
cam = len(sys.argv) > 2 and sys.argv[2] == "cam"
arch = sys.argv[3] if len(sys.argv) > 3 else "efficientnet_b0"
m = registry.get_model_spec(arch).builder(arch, 2)
synth_state(m)
m.to("cuda").eval()
x = torch.randn(n, 3, 224, 224, device="cuda")
with torch.inference_mode():
    for _ in range(3):
        m(x)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    out = m.forward_cam(x, None) if cam else m(x)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
print("ok", float(out[0].float().sum()) if cam else float(out.float().sum()))
