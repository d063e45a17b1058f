"""Layer-by-layer diff of a layer-program engine (EfficientFormerV2 / FasterViT) against its oracle, on a B200.

For every debug checkpoint the engine exposes (`b200df_debug_mark`, include/b200df.h) the forward is re-run up to that
op (`b200df_debug_stop_after`), the named activation buffer is read back and compared with the output of the oracle
module of the same name (forward hooks).  Prints the first checkpoint whose error exceeds the threshold -- the tool used
to localise a wrong kernel / wrong fold when a whole-net parity test fails.

    python tools/layer_diff.py efficientformerv2_s1 [fp32|bf16] [n]
"""

from __future__ import annotations

import ctypes as C
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from deepfakedetection_b200 import _lib, registry  # noqa: E402


def oracle_for(name: str):
    from oracle import efficientformerv2, fastervit

    if name.startswith("efficientformer"):
        return efficientformerv2.make_calibrated_oracle(0)
    return fastervit.make_calibrated_oracle(0, variant=name)


def reference_tensors(oracle, x: torch.Tensor) -> dict[str, torch.Tensor]:
    """name -> reference activation as [n, rows, c] (token-major / NHWC)."""
    store: dict[str, torch.Tensor] = {}

    def nhwc(t):
        return t.permute(0, 2, 3, 1).reshape(t.shape[0], -1, t.shape[1])

    hooks = []
    for name, mod in oracle.named_modules():
        if not name:
            continue

        def hook(m, inp, out, name=name):
            if isinstance(out, tuple):  # HAT block: (window tokens [B*nw, 49, C], carrier tokens [B, 16, C] or None)
                xw, ct = out
                store[name] = xw.detach()
                if ct is not None:
                    store[name + "#ct"] = ct.detach()
            elif out.ndim == 4:
                store[name] = nhwc(out.detach())
            elif out.ndim == 3:
                store[name] = out.detach()
        hooks.append(mod.register_forward_hook(hook))
        if type(mod).__name__ == "Block" and getattr(mod, "token_mixer", None) is not None:
            def pre(m, inp, name=name):
                xin = inp[0]
                store[name + ".+attn"] = nhwc((xin + m.ls1(m.token_mixer(xin))).detach())
            hooks.append(mod.register_forward_pre_hook(pre))
    with torch.no_grad():
        oracle(x)
    for h in hooks:
        h.remove()
    return store


def main(name: str = "efficientformerv2_s1", precision: str = "fp32", n: int = 2, thresh: float | None = None) -> int:
    from oracle.effnet_b0 import synth_images

    import os

    os.environ["B200DF_GRAPH_MAX_N"] = "0"   # every forward must really run (and stop where asked), not replay a graph
    lib = _lib.lib()
    oracle = oracle_for(name)
    x = synth_images(n, seed=5)
    ref = reference_tensors(oracle, x)
    reg_name = name + (".b200_fp32" if precision == "fp32" else "")
    model = registry.get_model_spec(reg_name).builder(reg_name, 2)
    model.to("cuda").eval()
    model.load_state_dict(oracle.state_dict(), strict=False)
    xg = x.cuda()
    with torch.inference_mode():
        full = model(xg).cpu()
    eng = model._engine
    h = eng._h
    ws_base, _ = eng._workspace(n)
    thresh = thresh if thresh is not None else (2e-3 if precision == "fp32" else 0.08)
    dt = torch.float32 if precision == "fp32" else torch.bfloat16
    es = 4 if precision == "fp32" else 2
    nm = lib.b200df_debug_num_marks(h)
    bad = 0
    for i in range(nm):
        nmp, op, dims, off = C.c_char_p(), C.c_int(), (C.c_int * 4)(), C.c_size_t()
        _lib.check(lib.b200df_debug_mark(h, i, n, C.byref(nmp), C.byref(op), C.byref(dims), C.byref(off)))
        mname = nmp.value.decode()
        seq_rows, row_off, rows, c = dims[0], dims[1], dims[2], dims[3]
        key = mname
        if key not in ref:
            print(f"[{i:3d}] op {op.value:3d} {mname}: no reference tensor (skipped)")
            continue
        r = ref[key]
        _lib.check(lib.b200df_debug_stop_after(h, op.value))
        with torch.inference_mode():
            model(xg)
        torch.cuda.synchronize()
        nseq = r.numel() // (rows * c)
        ws_buf = next(v[0] for k, v in eng._ws.items() if k[0] == n and v[1] == ws_base)   # the uint8 tensor behind `ws_base`
        o = ws_base - ws_buf.data_ptr() + off.value
        t = ws_buf[o:o + nseq * seq_rows * c * es].view(dt)
        got = t.view(nseq, seq_rows, c)[:, row_off:row_off + rows].float().cpu()
        r = r.reshape(nseq, rows, c)
        err = (got - r).abs().max().item()
        scale = r.abs().max().item()
        flag = "" if err <= thresh * max(scale, 1.0) else "   <-- MISMATCH"
        bad += bool(flag)
        print(f"[{i:3d}] op {op.value:3d} {mname:55s} max|d| {err:.3e}  (ref max {scale:.3e}){flag}")
    _lib.check(lib.b200df_debug_stop_after(h, -1))
    with torch.no_grad():
        want = oracle(x)
    print("logits engine:", full.flatten().tolist())
    print("logits oracle:", want.flatten().tolist())
    print(f"{bad} checkpoint(s) out of tolerance")
    return bad


if __name__ == "__main__":
    a = sys.argv[1:]
    sys.exit(1 if main(a[0] if a else "efficientformerv2_s1", a[1] if len(a) > 1 else "fp32", int(a[2]) if len(a) > 2 else 2) else 0)
