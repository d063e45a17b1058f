"""Batch-sharded multi-GPU evaluation (one process per GPU, torchrun-style).

The reference evaluates on a single device (/root/reference/orchestration/orchestrator.py:583-599)
and then computes accuracy / ROC-AUC / confusion matrix on the concatenated per-batch results
(orchestrator.py:601-647).  Images are independent in eval mode, so the path shards perfectly:
rank r of G evaluates a contiguous slice of the dataset with replicated weights and there is no
data-path collective.  The only exchange is at the end: every rank contributes its packed
``{probs[C], pred, label}`` records (``b200df_softmax_pack``) to ONE all-gather (NCCL over
NVLink on GPUs, gloo in the CPU tests) on padded equal-length buffers, then trims -- so the
metrics block sees exactly the tensors the single-process loop would have produced.
"""

from __future__ import annotations

import ctypes as C
from typing import Any

import numpy as np
import torch
import torch.distributed as dist


def shard_range(n_total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced slice [lo, hi) of ``n_total`` samples for ``rank`` (sizes differ by <= 1)."""
    base, rem = divmod(n_total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def pack_records(logits: torch.Tensor, labels: torch.Tensor | None) -> torch.Tensor:
    """softmax + argmax + label -> float32 [n, C+2] records, on the device, via the C ABI."""
    from . import _lib
    from .engine import _ptr, _stream_ptr

    n, c = logits.shape
    out = torch.empty((n, c + 2), dtype=torch.float32, device=logits.device)
    if n == 0:
        return out
    lab = labels.to(device=logits.device, dtype=torch.int64).contiguous() if labels is not None else None
    with torch.cuda.device(logits.device):
        _lib.check(_lib.lib().b200df_softmax_pack(_ptr(logits.contiguous()), _ptr(lab), n, c, _ptr(out),
                                                  _stream_ptr(logits.device)))
    return out


def gather_records(local: torch.Tensor, group=None) -> torch.Tensor:
    """All-gather variable-length [n_r, W] record blocks into rank order [sum n_r, W] (one collective
    for the sizes, one for the payload).  Works on CUDA/NCCL and CPU/gloo."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    width = local.shape[1]
    sizes = torch.zeros(world, dtype=torch.int64, device=local.device)
    mine = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
    dist.all_gather_into_tensor(sizes, mine, group=group)
    sizes_l = [int(s) for s in sizes.cpu()]
    cap = max(max(sizes_l), 1)
    padded = torch.zeros((cap, width), dtype=local.dtype, device=local.device)
    padded[: local.shape[0]] = local
    gathered = torch.empty((world * cap, width), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(gathered, padded, group=group)
    parts = [gathered[r * cap : r * cap + sizes_l[r]] for r in range(world)]
    return torch.cat(parts, 0)


def unpack_records(records: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """records [n, C+2] -> (probs float32 [n,C], preds int64 [n], targets int64 [n]) on the CPU,
    the three tensors orchestrator.py:601-603 concatenates."""
    r = records.detach().cpu()
    c = r.shape[1] - 2
    return r[:, :c].contiguous(), r[:, c].round().long(), r[:, c + 1].round().long()


def compute_metrics(probs: torch.Tensor, preds: torch.Tensor, targets: torch.Tensor, num_classes: int,
                    threshold: float | None = None) -> dict[str, Any]:
    """accuracy / roc_auc / confusion matrix exactly as orchestrator.py:605-635."""
    from sklearn.metrics import confusion_matrix, roc_auc_score

    if num_classes == 2 and threshold is not None:
        preds = (probs[:, 1] >= threshold).long()
    out: dict[str, Any] = {"accuracy": (preds == targets).float().mean().item() if len(targets) else float("nan")}
    if torch.unique(targets).numel() > 1:
        try:
            if num_classes == 2:
                out["roc_auc"] = float(roc_auc_score(targets.numpy(), probs[:, 1].numpy()))
            else:
                out["roc_auc"] = float(roc_auc_score(targets.numpy(), probs.numpy(), multi_class="ovr"))
        except ValueError:
            pass
    if num_classes == 2 and threshold is not None:
        out["threshold"] = float(threshold)
    out["confusion_matrix"] = confusion_matrix(targets.numpy(), preds.numpy()).tolist()
    return out


def best_threshold(scores: np.ndarray, truth: np.ndarray) -> float:
    """The reference's 501-point balanced-accuracy sweep (orchestrator.py:536-546), vectorised."""
    if scores.size == 0 or np.unique(truth).size < 2:
        return 0.5
    thr = np.linspace(0.0, 1.0, 501, dtype=np.float64)
    pred = scores[None, :] >= thr[:, None]
    pos, neg = truth == 1, truth == 0
    tpr = (pred & pos[None]).sum(1) / max(pos.sum(), 1)
    tnr = (~pred & neg[None]).sum(1) / max(neg.sum(), 1)
    bal = 0.5 * (tpr + tnr)
    return float(thr[int(np.argmax(bal))])  # first maximum, like the reference's strict `>`


@torch.inference_mode()
def evaluate_sharded(model, images: torch.Tensor, labels: torch.Tensor, batch_size: int = 256, group=None,
                     threshold: float | None = None) -> dict[str, Any]:
    """Evaluate ``images`` [n,3,224,224] / ``labels`` [n] (host or device tensors holding the FULL
    dataset on every rank) with the dataset sharded over the process group; returns the metrics
    (identical on every rank) plus the gathered tensors."""
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    lo, hi = shard_range(images.shape[0], rank, world)
    dev = next(model.parameters()).device
    recs = []
    for s in range(lo, hi, batch_size):
        e = min(s + batch_size, hi)
        x = images[s:e].to(dev, non_blocking=True)
        logits = model(x)
        recs.append(pack_records(logits, labels[s:e]))
    local = torch.cat(recs, 0) if recs else torch.empty((0, model.num_classes + 2), dtype=torch.float32, device=dev)
    probs, preds, targets = unpack_records(gather_records(local, group))
    out = compute_metrics(probs, preds, targets, model.num_classes, threshold)
    out.update(probs=probs, preds=preds, targets=targets)
    return out
