"""Thin torch-tensor wrappers over the per-kernel C-ABI entry points (GPU tests only)."""

import ctypes as C

import torch

from deepfakedetection_b200 import _lib


def P(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def tdtype(dt):
    return torch.bfloat16 if dt == _lib.DT_BF16 else torch.float32


def k_stem(x, w, b, dt):
    n, _, h, wd = x.shape
    out = torch.empty((n, h // 2, wd // 2, 32), dtype=tdtype(dt), device=x.device)
    _lib.check(_lib.lib().b200df_k_stem(P(x), n, h, wd, P(w), P(b), P(out), dt, stream()))
    return out


def k_stem_ex(x, w, b, dt, oc, pad):
    n, _, h, wd = x.shape
    out = torch.empty((n, h // 2, wd // 2, oc), dtype=tdtype(dt), device=x.device)
    _lib.check(_lib.lib().b200df_k_stem_ex(P(x), n, h, wd, P(w), P(b), P(out), dt, oc, pad, stream()))
    return out


def k_dwconv_ex(x, w_tap_major, bias, k, stride, pad_lo, dt):
    n, h, wd, c = x.shape
    ho, wo = (h - 1) // stride + 1, (wd - 1) // stride + 1
    out = torch.empty((n, ho, wo, c), dtype=tdtype(dt), device=x.device)
    pl = torch.zeros((n, c), dtype=torch.int64, device=x.device)
    _lib.check(_lib.lib().b200df_k_dwconv_ex(P(x), n, h, wd, c, k, stride, pad_lo, P(w_tap_major), P(bias), P(out),
                                             P(pl), dt, stream()))
    return out, (pl.double() / 2.0**24).float()


def k_dwconv(x, w_tap_major, bias, k, stride, dt, pool=True):
    n, h, wd, c = x.shape
    pad = (k - 1) // 2
    ho, wo = (h + 2 * pad - k) // stride + 1, (wd + 2 * pad - k) // stride + 1
    out = torch.empty((n, ho, wo, c), dtype=tdtype(dt), device=x.device)
    pl = torch.zeros((n, c), dtype=torch.int64, device=x.device) if pool else None  # Q24 fixed point
    _lib.check(_lib.lib().b200df_k_dwconv(P(x), n, h, wd, c, k, stride, P(w_tap_major), P(bias), P(out), P(pl), dt, stream()))
    return out, (pl.double() / 2.0**24).float() if pool else None


def k_se(pool, inv_hw, w1, b1, w2t, b2):
    n, c = pool.shape
    gate = torch.empty_like(pool)
    q = (pool.double() * 2.0**24).round().to(torch.int64)
    _lib.check(_lib.lib().b200df_k_se(P(q), n, c, w1.shape[0], C.c_float(inv_hw), P(w1), P(b1), P(w2t), P(b2), P(gate), stream()))
    return gate


def k_pwconv(x, w, bias, gate=None, rows_per_img=1, residual=None, act=0, dt=_lib.DT_BF16, impl=0):
    m, cin = x.shape
    cout = w.shape[0]
    out = torch.empty((m, cout), dtype=tdtype(dt), device=x.device)
    _lib.check(_lib.lib().b200df_k_pwconv(P(x), m, cin, cout, P(w), P(bias), P(gate), rows_per_img, P(residual), act, P(out), dt, impl, stream()))
    return out


# ---- kernels of the transformer-style backbones ----------------------------------------------------------------------
def k_stem_act(x, w, b, dt, oc, pad, act):
    n, _, h, wd = x.shape
    out = torch.empty((n, h // 2, wd // 2, oc), dtype=tdtype(dt), device=x.device)
    _lib.check(_lib.lib().b200df_k_stem_act(P(x), n, h, wd, P(w), P(b), P(out), dt, oc, pad, act, stream()))
    return out


def k_dwconv_act(x, w_tap_major, bias, k, stride, act, dt):
    n, h, wd, c = x.shape
    ho, wo = (h - 1) // stride + 1, (wd - 1) // stride + 1
    out = torch.empty((n, ho, wo, c), dtype=tdtype(dt), device=x.device)
    _lib.check(_lib.lib().b200df_k_dwconv_act(P(x), n, h, wd, c, k, stride, P(w_tap_major), P(bias), P(out), act, dt, stream()))
    return out


def k_conv3x3(x, w_k, bias, stride, act, dt, impl, residual=None):
    """x NHWC; w_k [cout, 9*cin] with k = (ky*3+kx)*cin + c."""
    n, h, wd, cin = x.shape
    cout = w_k.shape[0]
    ho, wo = (h - 1) // stride + 1, (wd - 1) // stride + 1
    out = torch.empty((n, ho, wo, cout), dtype=tdtype(dt), device=x.device)
    _lib.check(_lib.lib().b200df_k_conv3x3(P(x), n, h, wd, cin, cout, stride, P(w_k), P(bias), P(residual), act, P(out), dt,
                                          impl, stream()))
    return out


def k_attention(q, k, v, nseq, nq, nk, dk, dv, heads, scale, bias=None, th=None, vlocal=None, act=0, dt=_lib.DT_FP32,
                q_off=0, k_off=0, v_off=0):
    """q/k/v: 2-D [tokens, pitch] tensors (possibly the same fused buffer, columns selected by the offsets)."""
    out = torch.empty((nseq * nq, heads * dv), dtype=tdtype(dt), device=q.device)
    t = th or (None, None, None, None)
    _lib.check(_lib.lib().b200df_k_attention(P(q), P(k), P(v), q.shape[1], k.shape[1], v.shape[1], q_off, k_off, v_off, nseq,
                                            nq, nk, dk, dv, heads, C.c_float(scale), P(bias), P(t[0]), P(t[1]), P(t[2]),
                                            P(t[3]), P(vlocal), act, P(out), dt, stream()))
    return out


def k_layernorm(x, gamma, beta, eps, dt):
    y = torch.empty_like(x)
    _lib.check(_lib.lib().b200df_k_layernorm(P(x), P(y), P(gamma), P(beta), x.shape[0], x.shape[1], C.c_float(eps), dt, stream()))
    return y


def k_upsample2x_act(x, act, dt):
    n, side, _, c = x.shape
    y = torch.empty((n, 2 * side, 2 * side, c), dtype=tdtype(dt), device=x.device)
    _lib.check(_lib.lib().b200df_k_upsample2x_act(P(x), P(y), n, side, c, act, dt, stream()))
    return y
