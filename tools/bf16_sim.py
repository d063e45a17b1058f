"""CPU emulation of the bf16 engine's rounding points (dev tool; uses the oracle => not product code).
Answers: which rounding points dominate |dsoftmax| vs the fp32 oracle?"""
import sys, itertools
import torch
from torch.nn import functional as F
sys.path.insert(0, "/root/repo")
from oracle.effnet_b0 import make_calibrated_oracle, synth_images

def r(x, on=True):
    return x.bfloat16().float() if on else x

def fold(conv_w, bn):
    s = bn.weight / torch.sqrt(bn.running_var + bn.eps)
    return conv_w * s.view(-1, 1, 1, 1), bn.bias - bn.running_mean * s

def tanh_approx(x):
    # model MUFU.TANH: ~2^-11 relative error; emulate by rounding tanh to 11 bits of mantissa
    t = torch.tanh(x)
    return t  # placeholder (exact)

def silu_fast(x, approx):
    if not approx:
        return F.silu(x)
    h = 0.5 * x
    t = torch.tanh(h)
    # quantise tanh to ~2^-11 relative
    m, e = torch.frexp(t)
    t = torch.ldexp(torch.round(m * 2048) / 2048, e)
    return h * t + h

def forward(model, x, cfg):
    rw, ract, rgate, rtrunk, rA, approx = cfg["w"], cfg["act"], cfg["gate"], cfg["trunk"], cfg["A"], cfg["tanh"]
    w, b = fold(model.conv_stem.weight, model.bn1)
    y = silu_fast(F.conv2d(x, w, b, 2, 1), approx)   # stem weights fp32
    y = r(y, rtrunk)
    for stage in model.blocks:
        for blk in stage:
            inp = y
            if hasattr(blk, "conv_pwl"):
                w, b = fold(blk.conv_pw.weight, blk.bn1)
                y = r(silu_fast(F.conv2d(y, r(w, rw), b), approx), ract)
                w, b = fold(blk.conv_dw.weight, blk.bn2)
                pw, pbn = blk.conv_pwl, blk.bn3
            else:
                w, b = fold(blk.conv_dw.weight, blk.bn1)
                pw, pbn = blk.conv_pw, blk.bn2
            d = silu_fast(F.conv2d(y, w, b, blk.conv_dw.stride, blk.conv_dw.padding, groups=w.shape[0]), approx)
            pool = d.mean((2, 3), keepdim=True)       # pooled from fp32 pre-rounding values
            d = r(d, ract)
            g = torch.sigmoid(blk.se.conv_expand(F.silu(blk.se.conv_reduce(pool))))
            d = r(d * g, rgate)
            w, b = fold(pw.weight, pbn)
            y = F.conv2d(d, r(w, rw), b)
            if blk.has_skip:
                y = y + inp
            y = r(y, rtrunk)
    A = r(F.conv2d(y, r(model.conv_head.weight, rw)), rA)
    return model.forward_head(A)

if __name__ == "__main__":
    torch.set_num_threads(8)
    model = make_calibrated_oracle(seed=0)
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    x = synth_images(64, seed=11)[:n]
    with torch.no_grad():
        ref = model(x)
        base = dict(w=True, act=True, gate=True, trunk=True, A=True, tanh=True)
        variants = {"all": base, "none": {k: False for k in base}}
        for k in base:
            v = dict(base); v[k] = False
            variants["no_" + k] = v
        for k in base:
            v = {kk: False for kk in base}; v[k] = True
            variants["only_" + k] = v
        for name, cfg in variants.items():
            out = forward(model, x, cfg)
            dl = (out - ref).abs().max().item()
            dp = (out.softmax(1) - ref.softmax(1)).abs().max().item()
            print(f"{name:12s} max|dlogit| {dl:.4f}  max|dsoftmax| {dp:.4f}")
