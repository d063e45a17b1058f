"""``torch.library`` custom ops over the C ABI (``b200df::forward``, ``::trunk``, ``::head``,
``::head_backward``, ``::forward_cam``).

The ops take an integer engine id (see :func:`register_engine`) so they stay pure tensor
functions from PyTorch's point of view.  They are registered for CUDA only: there is no CPU
kernel and no fallback.  ``head`` carries an autograd formula (the closed-form Grad-CAM
gradient, ``b200df_head_backward``) so pytorch-grad-cam's hook protocol
(/root/reference/web_ui.py:275-279) works on the split trunk/head path.
"""

from __future__ import annotations

import itertools
import weakref

import torch

from .engine import Engine

_ENGINES: "weakref.WeakValueDictionary[int, Engine]" = weakref.WeakValueDictionary()
_ids = itertools.count(1)


def register_engine(engine: Engine) -> int:
    eid = next(_ids)
    _ENGINES[eid] = engine
    return eid


def _engine(eid: int) -> Engine:
    try:
        return _ENGINES[eid]
    except KeyError:
        raise RuntimeError(f"b200df engine {eid} no longer exists") from None


@torch.library.custom_op("b200df::forward", mutates_args=(), device_types="cuda")
def forward(images: torch.Tensor, engine: int) -> torch.Tensor:
    return _engine(engine).forward(images)


@forward.register_fake
def _(images, engine):
    return images.new_empty((images.shape[0], _engine(engine).num_classes), dtype=torch.float32)


@torch.library.custom_op("b200df::trunk", mutates_args=(), device_types="cuda")
def trunk(images: torch.Tensor, engine: int) -> tuple[torch.Tensor, torch.Tensor]:
    a, aux = _engine(engine).trunk(images)
    return a, aux


@trunk.register_fake
def _(images, engine):
    e = _engine(engine)
    return (images.new_empty((images.shape[0], e.feature_channels, e.feature_side, e.feature_side), dtype=torch.float32),
            images.new_empty((images.shape[0], e.num_classes), dtype=torch.float32))


@torch.library.custom_op("b200df::head", mutates_args=(), device_types="cuda")
def head(a: torch.Tensor, aux: torch.Tensor, engine: int) -> torch.Tensor:
    return _engine(engine).head(a, aux)


@head.register_fake
def _(a, aux, engine):
    return a.new_empty((a.shape[0], _engine(engine).num_classes), dtype=torch.float32)


@torch.library.custom_op("b200df::head_backward", mutates_args=(), device_types="cuda")
def head_backward(a: torch.Tensor, grad_logits: torch.Tensor, engine: int) -> torch.Tensor:
    return _engine(engine).head_backward(a, grad_logits)


@head_backward.register_fake
def _(a, grad_logits, engine):
    return torch.empty_like(a)


def _head_setup(ctx, inputs, output):
    a, _aux, engine = inputs
    ctx.save_for_backward(a)
    ctx.engine = engine


def _head_bwd(ctx, grad_logits):
    (a,) = ctx.saved_tensors
    return head_backward(a, grad_logits.contiguous(), ctx.engine), None, None


head.register_autograd(_head_bwd, setup_context=_head_setup)


@torch.library.custom_op("b200df::forward_cam", mutates_args=(), device_types="cuda")
def forward_cam(images: torch.Tensor, targets: torch.Tensor | None, engine: int) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    logits, low, cam = _engine(engine).forward_cam(images, targets, want_map=True)
    return logits, low, cam


@forward_cam.register_fake
def _(images, targets, engine):
    n = images.shape[0]
    return (
        images.new_empty((n, _engine(engine).num_classes), dtype=torch.float32),
        images.new_empty((n, _engine(engine).feature_side, _engine(engine).feature_side), dtype=torch.float32),
        images.new_empty((n, images.shape[-2], images.shape[-1]), dtype=torch.float32),
    )
