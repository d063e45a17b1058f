"""Python handle around one ``b200df_engine`` (C ABI in ``include/b200df.h``).

PyTorch is used here only for device memory, streams and dtype bookkeeping; all
arithmetic happens inside ``libb200df.so``.
"""

from __future__ import annotations

import collections
import ctypes as C
import os
import threading

import torch

from . import _lib

_ALIGN = 1024


def _ptr(t: torch.Tensor | None) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def _stream_ptr(device: torch.device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda_f32(t: torch.Tensor, what: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(
            f"{what} is on '{t.device}': the b200 backend has no CPU fallback -- put the model and "
            "its inputs on a CUDA (sm_100) device (set `device: cuda` in the YAML)."
        )
    if t.dtype != torch.float32:
        raise RuntimeError(f"{what} must be float32, got {t.dtype}")


class Engine:
    """One model instance on one GPU: weights (folded/packed) + a workspace per batch size."""

    def __init__(self, arch: int, num_classes: int, device: torch.device) -> None:
        self._lib = _lib.lib()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError(f"b200 backend needs a CUDA device, got '{device}' (no CPU fallback)")
        index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", index)
        self.num_classes = num_classes
        self._h = C.c_void_p()
        _lib.check(self._lib.b200df_create(arch, num_classes, index, C.byref(self._h)))
        ch, side = C.c_int(), C.c_int()
        _lib.check(self._lib.b200df_feature_shape(self._h, C.byref(ch), C.byref(side)))
        self.feature_channels, self.feature_side = int(ch.value), int(side.value)  # Grad-CAM activation A
        self.precision: int | None = None
        # workspaces keyed by (batch size, stream): two forwards of the same n on different streams never share scratch
        self._ws: "collections.OrderedDict[tuple[int, int], tuple[torch.Tensor, int, int, int]]" = collections.OrderedDict()
        self._lock = threading.Lock()
        # CUDA-graph replay for small batches (launch-bound regime: the UI's n = 1, strong scaling's 32 images per GPU):
        # one graph per batch size, captured on first use; B200DF_GRAPH_MAX_N=0 disables it
        self.graph_max_n = int(os.environ.get("B200DF_GRAPH_MAX_N", "64"))
        self._graphs: dict[int, tuple[torch.cuda.CUDAGraph, torch.Tensor, torch.Tensor]] = {}
        self._profiling = False

    # -- description ---------------------------------------------------------------------------------
    def expected_tensors(self) -> list[tuple[str, int]]:
        out = []
        name = C.c_char_p()
        numel = C.c_size_t()
        for i in range(self._lib.b200df_num_tensors(self._h)):
            _lib.check(self._lib.b200df_tensor_info(self._h, i, C.byref(name), C.byref(numel)))
            out.append((name.value.decode(), int(numel.value)))
        return out

    # -- weights -------------------------------------------------------------------------------------
    def load_state(self, state: dict[str, torch.Tensor], precision: int) -> None:
        """Push every expected tensor (timm names) and fold/pack for ``precision``."""
        with self._lock:
            for name, numel in self.expected_tensors():
                if name not in state:
                    raise KeyError(f"state is missing '{name}'")
                t = state[name].detach().to(device="cpu", dtype=torch.float32).contiguous()
                if t.numel() != numel:
                    raise ValueError(f"'{name}': {tuple(t.shape)} has {t.numel()} elements, engine expects {numel}")
                _lib.check(self._lib.b200df_set_tensor(self._h, name.encode(), _ptr(t), numel, 0))
            with torch.cuda.device(self.device):
                _lib.check(self._lib.b200df_finalize(self._h, precision))
            self.precision = precision
            self._drop_workspaces()

    # -- workspace -----------------------------------------------------------------------------------
    def _drop_workspaces(self) -> None:
        cur = torch.cuda.current_stream(self.device)
        for buf, _base, _size, _mode in self._ws.values():
            buf.record_stream(cur)   # kernels launched on another stream may still be using it
        self._ws.clear()
        self._graphs.clear()

    def _workspace(self, n: int, stream_id: int | None = None, mode: int = 0) -> tuple[int, int]:
        """mode 0: forward; mode 1: forward + Grad-CAM (a superset: FasterViT's CAM adds checkpoint / gradient buffers)."""
        if stream_id is None:
            stream_id = torch.cuda.current_stream(self.device).cuda_stream
        key = (n, stream_id)
        hit = self._ws.get(key)
        if hit is not None and hit[3] < mode:      # a forward-only workspace cannot serve a CAM call: replace it
            hit[0].record_stream(torch.cuda.current_stream(self.device))
            del self._ws[key]
            hit = None
        if hit is None:
            need = C.c_size_t()
            _lib.check(self._lib.b200df_workspace_bytes(self._h, n, mode, C.byref(need)))
            buf = torch.empty(int(need.value) + _ALIGN, dtype=torch.uint8, device=self.device)
            base = (buf.data_ptr() + _ALIGN - 1) // _ALIGN * _ALIGN
            while len(self._ws) >= 6:  # a few (batch size, stream) pairs: full batch, ragged tail, n = 1 from the UI ...
                _k, old = self._ws.popitem(last=False)   # least recently used
                old[0].record_stream(torch.cuda.current_stream(self.device))
            hit = (buf, base, int(need.value), mode)
            self._ws[key] = hit
        else:
            self._ws.move_to_end(key)
        return hit[1], hit[2]

    # -- hot path ------------------------------------------------------------------------------------
    def forward(self, images: torch.Tensor) -> torch.Tensor:
        require_cuda_f32(images, "images")
        n = images.shape[0]
        images = images.contiguous()
        logits = torch.empty((n, self.num_classes), dtype=torch.float32, device=images.device)
        if n == 0:
            return logits
        with self._lock, torch.cuda.device(images.device):
            if 0 < n <= self.graph_max_n and not self._profiling and not torch.cuda.is_current_stream_capturing():
                if self._forward_graphed(images, logits):
                    return logits
            base, size = self._workspace(n)
            _lib.check(self._lib.b200df_forward(self._h, _ptr(images), n, _ptr(logits), C.c_void_p(base), size,
                                               _stream_ptr(images.device)))
        return logits

    def _forward_graphed(self, images: torch.Tensor, logits: torch.Tensor) -> bool:
        """Replay (capturing on first use) the CUDA graph of one n-image forward: ~70-300 launches become one.
        Returns False (and switches graphs off) if capture is not possible on this driver."""
        n = images.shape[0]
        hit = self._graphs.get(n)
        if hit is None:
            try:
                cur = torch.cuda.current_stream(self.device)
                side = torch.cuda.Stream(self.device)
                with torch.inference_mode(False):   # the static buffers outlive the caller's inference_mode block
                    x = torch.empty(images.shape, dtype=images.dtype, device=images.device)
                    y = torch.empty(logits.shape, dtype=logits.dtype, device=logits.device)
                base, size = self._workspace(n, stream_id=-n)   # scratch owned by this graph
                side.wait_stream(cur)
                with torch.cuda.stream(side):
                    x.copy_(images)
                    # un-captured warm-up: encodes the tensor maps / tile plans of this (n, workspace) on the host
                    _lib.check(self._lib.b200df_forward(self._h, _ptr(x), n, _ptr(y), C.c_void_p(base), size,
                                                       _stream_ptr(self.device)))
                side.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=side):
                    _lib.check(self._lib.b200df_forward(self._h, _ptr(x), n, _ptr(y), C.c_void_p(base), size,
                                                       _stream_ptr(self.device)))
                cur.wait_stream(side)
                hit = (g, x, y)
                self._graphs[n] = hit
            except Exception as ex:  # noqa: BLE001 - capture is an optimisation, never a requirement
                import warnings

                warnings.warn(f"b200 backend: CUDA-graph capture failed ({ex}); small batches take the direct path", stacklevel=2)
                self.graph_max_n = 0
                return False
        g, x, y = hit
        x.copy_(images, non_blocking=True)
        g.replay()
        logits.copy_(y, non_blocking=True)
        return True

    def forward_u8(self, images_hwc: torch.Tensor, mean, std) -> torch.Tensor:
        if not images_hwc.is_cuda or images_hwc.dtype != torch.uint8:
            raise RuntimeError("forward_u8 expects a CUDA uint8 [n,224,224,3] tensor")
        n = images_hwc.shape[0]
        images_hwc = images_hwc.contiguous()
        logits = torch.empty((n, self.num_classes), dtype=torch.float32, device=images_hwc.device)
        if n == 0:
            return logits
        m = (C.c_float * 3)(*mean) if mean is not None else None
        s = (C.c_float * 3)(*std) if std is not None else None
        with self._lock, torch.cuda.device(images_hwc.device):
            base, size = self._workspace(n)
            _lib.check(self._lib.b200df_forward_u8(self._h, _ptr(images_hwc), n, m, s, _ptr(logits),
                                                  C.c_void_p(base), size, _stream_ptr(images_hwc.device)))
        return logits

    def forward_cam(self, images: torch.Tensor, targets: torch.Tensor | None, want_map: bool = True):
        """-> (logits [n,C], cam_lowres [n,7,7], cam [n,224,224] or None)."""
        require_cuda_f32(images, "images")
        n = images.shape[0]
        images = images.contiguous()
        dev = images.device
        logits = torch.empty((n, self.num_classes), dtype=torch.float32, device=dev)
        low = torch.empty((n, self.feature_side, self.feature_side), dtype=torch.float32, device=dev)
        cam = torch.empty((n, images.shape[-2], images.shape[-1]), dtype=torch.float32, device=dev) if want_map else None
        if n == 0:
            return logits, low, cam
        tgt = None
        if targets is not None:
            tgt = targets.to(device=dev, dtype=torch.int32).contiguous()
        with self._lock, torch.cuda.device(dev):
            base, size = self._workspace(n, mode=1)
            _lib.check(self._lib.b200df_forward_cam(self._h, _ptr(images), n, _ptr(tgt), _ptr(logits), _ptr(low),
                                                   _ptr(cam), C.c_void_p(base), size, _stream_ptr(dev)))
        return logits, low, cam

    def trunk(self, images: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
        """-> (A [n,C,s,s] fp32 NCHW: the raw output of the reference's Grad-CAM target conv, aux [n,num_classes])."""
        require_cuda_f32(images, "images")
        n = images.shape[0]
        images = images.contiguous()
        a = torch.empty((n, self.feature_channels, self.feature_side, self.feature_side), dtype=torch.float32,
                        device=images.device)
        aux = torch.zeros((n, self.num_classes), dtype=torch.float32, device=images.device)
        if n == 0:
            return a, aux
        with self._lock, torch.cuda.device(images.device):
            # mode 1: on FasterViT the trunk leaves the activation checkpoints of the Grad-CAM backward program in this
            # workspace; head_backward runs on them (the workspace stays cached per (n, stream) until evicted)
            base, size = self._workspace(n, mode=1)
            _lib.check(self._lib.b200df_trunk_ex(self._h, _ptr(images), n, _ptr(a), _ptr(aux), C.c_void_p(base), size,
                                                _stream_ptr(images.device)))
        return a, aux

    def head(self, a: torch.Tensor, aux: torch.Tensor | None = None) -> torch.Tensor:
        require_cuda_f32(a, "activation")
        a = a.contiguous()
        n = a.shape[0]
        logits = torch.empty((n, self.num_classes), dtype=torch.float32, device=a.device)
        if aux is None:
            aux = torch.zeros((n, self.num_classes), dtype=torch.float32, device=a.device)
        aux = aux.to(dtype=torch.float32).contiguous()
        with self._lock, torch.cuda.device(a.device):
            _lib.check(self._lib.b200df_head_ex(self._h, _ptr(a), _ptr(aux), n, _ptr(logits), _stream_ptr(a.device)))
        return logits

    def head_backward(self, a: torch.Tensor, grad_logits: torch.Tensor) -> torch.Tensor:
        require_cuda_f32(a, "activation")
        a = a.contiguous()
        g = grad_logits.to(dtype=torch.float32).contiguous()
        out = torch.empty_like(a)
        with self._lock, torch.cuda.device(a.device):
            _lib.check(self._lib.b200df_head_backward(self._h, _ptr(a), _ptr(g), a.shape[0], _ptr(out),
                                                     _stream_ptr(a.device)))
        return out

    # -- per-launch profiling (bench.py) -----------------------------------------------------------------
    def set_profile(self, enable: bool) -> None:
        self._profiling = bool(enable)
        _lib.check(self._lib.b200df_set_profile(self._h, 1 if enable else 0))

    def profile_records(self) -> list[tuple[str, float, float, float]]:
        """[(name, ms, algorithmic bytes, flops)] for every launch since the last reset (synchronises)."""
        out = []
        name, ms, by, fl = C.c_char_p(), C.c_float(), C.c_double(), C.c_double()
        for i in range(self._lib.b200df_profile_count(self._h)):
            _lib.check(self._lib.b200df_profile_get(self._h, i, C.byref(name), C.byref(ms), C.byref(by), C.byref(fl)))
            out.append((name.value.decode(), float(ms.value), float(by.value), float(fl.value)))
        _lib.check(self._lib.b200df_profile_reset(self._h))
        return out

    def close(self) -> None:
        if self._h:
            self._lib.b200df_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self) -> None:  # noqa: D105
        try:
            self.close()
        except Exception:  # pragma: no cover - interpreter teardown
            pass
