"""GPU debug: compare bf16 engine variants (tcgen05 vs SIMT GEMM, TMA vs plain depthwise) per sample."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from deepfakedetection_b200 import registry
from oracle.effnet_b0 import synth_images, make_calibrated_oracle


def build(sd, name, **env):
    for k, v in env.items():
        os.environ[k] = v
    m = registry.get_model_spec(name).builder(name, 2)
    m.to("cuda").eval()
    m.load_state_dict(sd, strict=False)
    with torch.inference_mode():
        m(torch.zeros(1, 3, 224, 224, device="cuda"))  # force pack while env is set
    for k in env:
        del os.environ[k]
    return m


om = make_calibrated_oracle(seed=0)
x = synth_images(64, seed=11)
with torch.no_grad():
    ref = om(x)
sd = om.state_dict()
variants = {
    "tc+tma": build(sd, "efficientnet_b0"),
    "simt+tma": build(sd, "efficientnet_b0", B200DF_GEMM="simt"),
    "tc+plain": build(sd, "efficientnet_b0", B200DF_DW="plain"),
    "simt+plain": build(sd, "efficientnet_b0", B200DF_GEMM="simt", B200DF_DW="plain"),
}
torch.set_printoptions(linewidth=200, precision=3)
for n in (16, 64):
    print("== n", n)
    outs = {}
    for nm, m in variants.items():
        with torch.inference_mode():
            a = m(x[:n].cuda()).cpu()
            b = m(x[:n].cuda()).cpu()
        outs[nm] = a
        print(nm, "rerun-diff", (a - b).abs().max().item(), "vs ref per-sample max:", (a - ref[:n]).abs().max(1).values[:16])
    for nm in variants:
        print(nm, "vs tc+tma", (outs[nm] - outs["tc+tma"]).abs().max().item())
