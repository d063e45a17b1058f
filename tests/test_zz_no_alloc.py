"""`b200df_head_static`: the allocation-free variant of the split head (include/b200df.h), called through the C-ABI and
checked against the existing entry point `b200df_head` on two calls in a row.  Its pooled scratch is a file-scope
`static __device__` array whose address the library resolves with cudaGetSymbolAddress; a wrong address would fault or
give garbage here.  The file is named so that pytest collects it LAST."""

import ctypes as C

import pytest
import torch

from deepfakedetection_b200 import _lib, registry

pytestmark = pytest.mark.gpu
MAX_FLOATS = 256 * 1536   # B200DF_HEAD_STATIC_MAX_FLOATS


@pytest.fixture(autouse=True)
def _need_gpu():
    assert torch.cuda.is_available() and _lib.lib().b200df_device_count() > 0, "these tests need a B200"
    torch.cuda.set_device(0)


def _P(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


@pytest.fixture(scope="module")
def engine(oracle_model):
    name = "efficientnet_b0.b200_fp32"
    model = registry.get_model_spec(name).builder(name, 2)
    model.to(torch.device("cuda", 0)).eval()
    model.load_state_dict(oracle_model.state_dict(), strict=False)
    with torch.inference_mode():
        model(torch.zeros(1, 3, 224, 224, device="cuda"))     # builds the engine
    return model._engine


def _head(fn, eng, a):
    logits = torch.full((a.shape[0], 2), float("nan"), dtype=torch.float32, device=a.device)
    _lib.check(fn(eng._h, _P(a), a.shape[0], _P(logits), _stream()))
    torch.cuda.synchronize()
    return logits


def test_head_static_matches_head_on_two_calls_in_a_row(engine):
    lib = _lib.lib()
    g = torch.Generator(device="cuda").manual_seed(5)
    # second call: other data and a SMALLER batch, so the array still holds rows of the first call behind it
    batches = [torch.randn(64, 1280, 7, 7, device="cuda", generator=g), 3.0 * torch.randn(5, 1280, 7, 7, device="cuda", generator=g)]
    for i, a in enumerate(batches):
        want = _head(lib.b200df_head, engine, a)
        free_before = torch.cuda.mem_get_info()[0]
        got = _head(lib.b200df_head_static, engine, a)
        free_after = torch.cuda.mem_get_info()[0]
        assert torch.isfinite(got).all(), f"call {i}: non-finite logits"
        # same two kernels on the same inputs; the tolerance is the split-head test's (test_gpu_parity_hardening.py)
        torch.testing.assert_close(got, want, atol=1e-5, rtol=1e-5)
        if i > 0:   # (the first call may page the module in: CUDA loads it lazily)
            assert free_after >= free_before, "b200df_head_static took device memory"


def test_head_static_rejects_sizes_beyond_the_static_array(engine):
    lib = _lib.lib()
    n = MAX_FLOATS // 1280 + 1
    a = torch.zeros(8, device="cuda")          # never read: the size check comes first
    out = torch.zeros(2 * n, device="cuda")
    rc = lib.b200df_head_static(engine._h, _P(a), n, _P(out), _stream())
    assert rc == _lib.E_INVALID
    assert b"static scratch" in lib.b200df_last_error()
    # the largest batch that fits still works
    n_ok = MAX_FLOATS // 1280
    a = torch.randn(n_ok, 1280, 7, 7, device="cuda")
    torch.testing.assert_close(_head(lib.b200df_head_static, engine, a), _head(lib.b200df_head, engine, a), atol=1e-5, rtol=1e-5)
