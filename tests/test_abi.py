"""The C-ABI shared library loads without a GPU and exports exactly what include/b200df.h declares."""

import ctypes as C
import re
import subprocess
from pathlib import Path

import pytest

from deepfakedetection_b200 import _lib

ROOT = Path(__file__).resolve().parent.parent


def header_symbols() -> list[str]:
    text = (ROOT / "include" / "b200df.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200df_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _lib.lib()
    syms = header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/b200df.h but not exported"
    assert sorted(_lib.SIGNATURES) == syms, "ctypes table and header disagree"
    out = subprocess.run(["nm", "-D", "--defined-only", str(_lib.LIB_PATH)], capture_output=True, text=True).stdout
    exported = sorted(set(re.findall(r" T (b200df_[a-z0-9_]+)", out)))
    assert exported == syms, "library exports symbols the header does not declare (or vice versa)"


def test_library_is_self_contained():
    # statically linked cudart, no libcuda/libtorch dependency: loads on a box without a driver
    out = subprocess.run(["ldd", str(_lib.LIB_PATH)], capture_output=True, text=True).stdout
    assert "libtorch" not in out and "libcuda.so" not in out and "libcudart" not in out


def test_describe_matches_oracle_state_dict():
    from oracle.effnet_b0 import make_seeded_oracle

    lib = _lib.lib()
    h = C.c_void_p()
    _lib.check(lib.b200df_create(_lib.ARCH_EFFICIENTNET_B0, 2, 0, C.byref(h)))
    try:
        sd = {k: v for k, v in make_seeded_oracle(0).state_dict().items() if "num_batches_tracked" not in k}
        name, numel = C.c_char_p(), C.c_size_t()
        seen = {}
        for i in range(lib.b200df_num_tensors(h)):
            _lib.check(lib.b200df_tensor_info(h, i, C.byref(name), C.byref(numel)))
            seen[name.value.decode()] = numel.value
        assert set(seen) == set(sd)
        for k, v in sd.items():
            assert seen[k] == v.numel(), k
    finally:
        lib.b200df_destroy(h)


def test_error_paths_without_compute():
    lib = _lib.lib()
    h = C.c_void_p()
    assert lib.b200df_create(99, 2, 0, C.byref(h)) == _lib.E_INVALID
    assert b"unknown arch" in lib.b200df_last_error()
    _lib.check(lib.b200df_create(0, 2, 0, C.byref(h)))
    try:
        buf = (C.c_float * 4)()
        assert lib.b200df_set_tensor(h, b"nope.weight", buf, 4, 0) == _lib.E_INVALID
        assert lib.b200df_set_tensor(h, b"bn1.weight", buf, 4, 0) == _lib.E_INVALID  # wrong numel
        assert lib.b200df_finalize(h, 0) == _lib.E_STATE  # tensors missing
        need = C.c_size_t()
        assert lib.b200df_workspace_bytes(h, 8, 0, C.byref(need)) == _lib.E_STATE
        assert lib.b200df_forward(h, None, 1, None, None, 0, None) == _lib.E_STATE
    finally:
        lib.b200df_destroy(h)


def test_no_device_means_error_not_fallback():
    lib = _lib.lib()
    if lib.b200df_device_count() > 0:
        pytest.skip("a B200 is present")
    from oracle.effnet_b0 import make_seeded_oracle

    h = C.c_void_p()
    _lib.check(lib.b200df_create(0, 2, 0, C.byref(h)))
    try:
        for k, v in make_seeded_oracle(0).state_dict().items():
            if "num_batches_tracked" in k:
                continue
            t = v.float().contiguous()
            _lib.check(lib.b200df_set_tensor(h, k.encode(), C.c_void_p(t.data_ptr()), t.numel(), 0))
        assert lib.b200df_finalize(h, _lib.PREC_BF16) == _lib.E_NO_DEVICE
        assert b"no CPU fallback" in lib.b200df_last_error()
    finally:
        lib.b200df_destroy(h)
