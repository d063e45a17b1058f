"""ctypes binding of ``libb200df.so`` (the C ABI declared in ``include/b200df.h``).

There is deliberately no fallback: if the shared library has not been built
(``python __graft_entry__.py build`` / ``make``) importing the compute path
raises, and every compute call raises if no sm_100 device is present.
"""

from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

_PKG = Path(__file__).resolve().parent
REPO_ROOT = _PKG.parent
# B200DF_LIB: another build of the SAME library (A/B timing runs, the `make prof` build); never a fallback
LIB_PATH = Path(os.environ["B200DF_LIB"]).resolve() if os.environ.get("B200DF_LIB") else _PKG / "libb200df.so"

OK = 0
E_INVALID, E_STATE, E_CUDA, E_NO_DEVICE, E_WORKSPACE = -1, -2, -3, -4, -5
ARCH_EFFICIENTNET_B0 = 0
ARCH_EFFICIENTNET_B3 = 1
ARCH_EFFICIENTFORMERV2_S1 = 2
ARCH_FASTERVIT_0 = 3
ARCH_FASTERVIT_2 = 4
PREC_BF16, PREC_FP32 = 0, 1
DT_BF16, DT_FP32 = 0, 1

# name -> (restype, argtypes); must list every symbol include/b200df.h declares
# (tests/test_abi.py parses the header and checks this table and the .so against it).
_vp, _i, _ll, _sz, _f = C.c_void_p, C.c_int, C.c_longlong, C.c_size_t, C.c_float
SIGNATURES: dict[str, tuple] = {
    "b200df_last_error": (C.c_char_p, []),
    "b200df_abi_version": (_i, []),
    "b200df_device_count": (_i, []),
    "b200df_create": (_i, [_i, _i, _i, C.POINTER(_vp)]),
    "b200df_destroy": (_i, [_vp]),
    "b200df_set_tensor": (_i, [_vp, C.c_char_p, _vp, _sz, _i]),
    "b200df_num_tensors": (_i, [_vp]),
    "b200df_tensor_info": (_i, [_vp, _i, C.POINTER(C.c_char_p), C.POINTER(_sz)]),
    "b200df_feature_shape": (_i, [_vp, C.POINTER(_i), C.POINTER(_i)]),
    "b200df_finalize": (_i, [_vp, _i]),
    "b200df_workspace_bytes": (_i, [_vp, _i, _i, C.POINTER(_sz)]),
    "b200df_forward": (_i, [_vp, _vp, _i, _vp, _vp, _sz, _vp]),
    "b200df_forward_u8": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "b200df_forward_cam": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "b200df_trunk": (_i, [_vp, _vp, _i, _vp, _vp, _sz, _vp]),
    "b200df_head": (_i, [_vp, _vp, _i, _vp, _vp]),
    "b200df_head_backward": (_i, [_vp, _vp, _vp, _i, _vp, _vp]),
    "b200df_trunk_ex": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _sz, _vp]),
    "b200df_head_ex": (_i, [_vp, _vp, _vp, _i, _vp, _vp]),
    "b200df_head_static": (_i, [_vp, _vp, _i, _vp, _vp]),
    "b200df_softmax_pack": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "b200df_cam_overlay": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp, C.c_float, _vp, _vp]),
    "b200df_threshold_hist": (_i, [_vp, _vp, _i, _vp, _i, _vp, _vp]),
    "b200df_confusion_matrix": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "b200df_auc_counts": (_i, [_vp, _i, _vp, _i, _vp, _vp]),
    "b200df_debug_num_marks": (_i, [_vp]),
    "b200df_debug_mark": (_i, [_vp, _i, _i, C.POINTER(C.c_char_p), C.POINTER(_i), C.POINTER(_i * 4), C.POINTER(_sz)]),
    "b200df_debug_stop_after": (_i, [_vp, _i]),
    "b200df_launch_count": (_ll, []),
    "b200df_reset_launch_count": (None, []),
    "b200df_set_profile": (_i, [_vp, _i]),
    "b200df_profile_count": (_i, [_vp]),
    "b200df_profile_get": (_i, [_vp, _i, C.POINTER(C.c_char_p), C.POINTER(_f), C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "b200df_profile_reset": (_i, [_vp]),
    "b200df_k_stem": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _i, _vp]),
    "b200df_k_stem_ex": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "b200df_k_dwconv_ex": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _vp]),
    "b200df_k_dwconv": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _vp]),
    "b200df_k_se": (_i, [_vp, _i, _i, _i, _f, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b200df_k_pwconv": (_i, [_vp, _ll, _i, _i, _vp, _vp, _vp, _i, _vp, _i, _vp, _i, _i, _vp]),
    "b200df_k_head_pool": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _i, _vp]),
    "b200df_k_cam": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp]),
    "b200df_k_cam_upsample": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "b200df_k_stem_act": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "b200df_k_dwconv_act": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _i, _i, _vp]),
    "b200df_k_conv3x3": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _i, _vp, _i, _i, _vp]),
    "b200df_k_attention": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _f, _vp, _vp, _vp, _vp, _vp,
                                _vp, _i, _vp, _i, _vp]),
    "b200df_k_layernorm": (_i, [_vp, _vp, _vp, _vp, _ll, _i, _f, _i, _vp]),
    "b200df_k_attention_bwd": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _f, _vp, _vp, _vp, _i, _vp]),
    "b200df_k_layernorm_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _ll, _i, _f, _i, _vp]),
    "b200df_k_gelu_bwd": (_i, [_vp, _vp, _ll, _i, _vp]),
    "b200df_k_conv3x3_s2_bwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "b200df_k_ct_pool_bwd": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "b200df_k_upsample2x_act": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
}

_lib = None


class B200DFError(RuntimeError):
    """A libb200df call failed (carries the C-side status code)."""

    def __init__(self, code: int, message: str) -> None:
        super().__init__(f"libb200df error {code}: {message}")
        self.code = code


def build(verbose: bool = False) -> Path:
    """Compile libb200df.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", str(REPO_ROOT), "-j", str(min(8, os.cpu_count() or 1))]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"building libb200df.so failed:\n{proc.stdout[-4000:]}\n{proc.stderr[-4000:]}")
    if verbose:
        print(proc.stdout[-2000:])
    return LIB_PATH


def lib() -> C.CDLL:
    """Load the library once and install the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python __graft_entry__.py build` (or `make`). "
            "deepfakedetection_b200 has no CPU or PyTorch fallback for its compute path."
        )
    handle = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(handle, name)
        fn.restype = res
        fn.argtypes = args
    if handle.b200df_abi_version() != 1:
        raise ImportError("libb200df.so ABI version mismatch; rebuild it")
    _lib = handle
    return handle


def check(code: int) -> None:
    if code != OK:
        msg = lib().b200df_last_error()
        raise B200DFError(code, msg.decode("utf-8", "replace") if msg else "")


def launch_count() -> int:
    return int(lib().b200df_launch_count())


def reset_launch_count() -> None:
    lib().b200df_reset_launch_count()
