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


class HostBatchFeeder:
    """`images.to(device, non_blocking=True)` (/root/reference/orchestration/orchestrator.py:588) for pinned host
    batches, split over ``parts`` copy streams.  One cudaMemcpyAsync keeps a single copy engine busy (47 GB/s
    measured on the B200 boxes for a 154 MB fp32 batch); two halves on two streams reach 54 GB/s, which is
    what bounds the fp32-input end-to-end rate (tools/h2d_bench.py).  Double-buffered: ``feed(i, host)`` starts
    the copy of batch ``i`` into buffer ``i & 1`` and makes the current (compute) stream wait for it;
    ``release(i)`` marks the buffer free once the consumer work has been enqueued."""

    def __init__(self, example: torch.Tensor, device, parts: int = 2) -> None:
        self.device = torch.device(device)
        self.parts = max(1, min(parts, example.shape[0]))
        self.bufs = [torch.empty(example.shape, dtype=example.dtype, device=self.device) for _ in range(2)]
        self.streams = [torch.cuda.Stream(self.device) for _ in range(self.parts)]
        self.copied = [[torch.cuda.Event() for _ in range(self.parts)] for _ in range(2)]
        self.consumed = [torch.cuda.Event() for _ in range(2)]
        self.used = [False, False]

    def wait_event_on_copy_streams(self, ev) -> None:
        for s in self.streams:
            s.wait_event(ev)

    def feed(self, i: int, host: torch.Tensor) -> torch.Tensor:
        j = i & 1
        buf = self.bufs[j]
        comp = torch.cuda.current_stream(self.device)
        for s, ev, h, d in zip(self.streams, self.copied[j], host.chunk(self.parts), buf.chunk(self.parts)):
            with torch.cuda.stream(s):
                if self.used[j]:
                    s.wait_event(self.consumed[j])
                d.copy_(h, non_blocking=True)
                ev.record(s)
        for ev in self.copied[j]:
            comp.wait_event(ev)
        return buf

    def release(self, i: int) -> None:
        j = i & 1
        self.consumed[j].record(torch.cuda.current_stream(self.device))
        self.used[j] = True


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


def gather_records(local: torch.Tensor, group=None, equal_shards: bool = False) -> torch.Tensor:
    """All-gather variable-length [n_r, W] record blocks into rank order [sum n_r, W] (one collective
    for the sizes, one for the payload).  Works on CUDA/NCCL and CPU/gloo.  ``equal_shards=True`` (every rank holds
    the same number of records, e.g. fixed per-GPU batches) skips the size exchange and its host sync: ONE collective."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    width = local.shape[1]
    if equal_shards:
        gathered = torch.empty((world * local.shape[0], width), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(gathered, local.contiguous(), group=group)
        return gathered
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


# ---- on-device metrics tail ---------------------------------------------------------
def threshold_from_hist(hist: np.ndarray, thresholds: np.ndarray) -> float:
    """Histograms of `b200df_threshold_hist` -> the reference's best threshold (orchestrator.py:536-546).

    hist[c, j] = number of class-c samples whose score is >= exactly j thresholds.  For threshold k the
    predicted positives are the samples with j > k, so TP/FP are suffix sums; balanced accuracy is
    sklearn's mean of per-class recalls in float64; the first strict maximum wins, like the reference loop.
    Pure host arithmetic on 2 x 502 integers (testable without a GPU)."""
    hist = np.asarray(hist, dtype=np.int64).reshape(2, -1)
    nthr = hist.shape[1] - 1
    n_neg, n_pos = int(hist[0].sum()), int(hist[1].sum())
    if n_neg == 0 or n_pos == 0:
        return 0.5  # the reference skips the sweep when only one class is present
    suffix = np.cumsum(hist[:, ::-1], axis=1)[:, ::-1]      # suffix[c, j] = sum_{i >= j} hist[c, i]
    best_bal, chosen = -1.0, 0.5
    for k in range(nthr):
        tp = int(suffix[1, k + 1])
        fp = int(suffix[0, k + 1])
        recalls = np.array([(n_neg - fp) / n_neg, tp / n_pos], dtype=np.float64)
        bal = float(np.mean(recalls))
        if bal > best_bal:
            best_bal, chosen = bal, float(thresholds[k])
    return chosen


def best_threshold_device(scores: torch.Tensor, labels: torch.Tensor) -> float:
    """The 501-point sweep with the O(N) part on the GPU: scores float32 [n] (probs[:, 1]), labels int64 [n]."""
    from . import _lib
    from .engine import _ptr, _stream_ptr

    thr = np.linspace(0.0, 1.0, 501, dtype=np.float64)
    n = scores.shape[0]
    if n == 0:
        return 0.5
    dev = scores.device
    scores = scores.contiguous().float()
    labels = labels.to(device=dev, dtype=torch.int64).contiguous()
    thr_d = torch.from_numpy(thr).to(dev)
    hist = torch.zeros(2 * (thr.size + 1), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().b200df_threshold_hist(_ptr(scores), _ptr(labels), n, _ptr(thr_d), thr.size, _ptr(hist),
                                                    _stream_ptr(dev)))
    return threshold_from_hist(hist.cpu().numpy(), thr)


def confusion_matrix_device(targets: torch.Tensor, preds: torch.Tensor, num_classes: int) -> np.ndarray:
    """counts[target, pred] (sklearn.metrics.confusion_matrix with labels 0..C-1) on the GPU."""
    from . import _lib
    from .engine import _ptr, _stream_ptr

    dev = targets.device
    cm = torch.zeros((num_classes, num_classes), dtype=torch.int64, device=dev)
    n = targets.shape[0]
    if n:
        t = targets.to(torch.int64).contiguous()
        p = preds.to(device=dev, dtype=torch.int64).contiguous()
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().b200df_confusion_matrix(_ptr(t), _ptr(p), n, num_classes, _ptr(cm), _stream_ptr(dev)))
    return cm.cpu().numpy()


def roc_auc_device(targets: torch.Tensor, scores: torch.Tensor) -> float:
    """Binary ROC-AUC = P(score_pos > score_neg) + P(equal)/2 (what roc_auc_score's trapezoid integrates to).
    torch.sort orders the negatives (plumbing); the pair counting is `b200df_auc_counts`."""
    from . import _lib
    from .engine import _ptr, _stream_ptr

    dev = scores.device
    scores = scores.contiguous().float()
    pos = scores[targets == 1].contiguous()
    neg = torch.sort(scores[targets != 1]).values.contiguous()
    if pos.numel() == 0 or neg.numel() == 0:
        raise ValueError("Only one class present in y_true. ROC AUC score is not defined in that case.")
    out = torch.zeros(2, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().b200df_auc_counts(_ptr(pos), pos.numel(), _ptr(neg), neg.numel(), _ptr(out), _stream_ptr(dev)))
    less, equal = (int(v) for v in out.cpu())
    return (less + 0.5 * equal) / (pos.numel() * neg.numel())


def compute_metrics_device(records: torch.Tensor, num_classes: int, threshold: float | None = None) -> dict[str, Any]:
    """`compute_metrics` (orchestrator.py:605-635) on the gathered device records [n, C+2] without the O(N)
    host work: only the confusion matrix / AUC counts cross PCIe.

    Device scope = the reference's deployed case: binary heads with both classes present.  Everything else -- more than
    two classes (sklearn's one-vs-rest AUC, :624-626) or data in which a class never occurs (sklearn then returns a
    confusion matrix over the OBSERVED labels only) -- is handed to `compute_metrics` on the host, so the numbers are the
    reference's in every case."""
    c = records.shape[1] - 2
    probs = records[:, :c]
    preds = records[:, c].round().long()
    targets = records[:, c + 1].round().long()
    seen = torch.unique(torch.cat([targets, preds if not (num_classes == 2 and threshold is not None) else
                                   (probs[:, 1] >= threshold).long()]))
    if records.shape[0] == 0 or num_classes != 2 or seen.numel() < num_classes or int(targets.min()) < 0 or int(targets.max()) >= num_classes:
        p, q, t = unpack_records(records)
        return compute_metrics(p, q, t, num_classes, threshold)
    if num_classes == 2 and threshold is not None:
        preds = (probs[:, 1] >= threshold).long()   # the reference's own expression (orchestrator.py:609)
    cm = confusion_matrix_device(targets, preds, num_classes)
    n = int(cm.sum())
    out: dict[str, Any] = {"accuracy": float(np.float32(np.trace(cm)) / np.float32(n)) if n else float("nan")}
    if (cm.sum(1) > 0).sum() > 1 and num_classes == 2:
        out["roc_auc"] = roc_auc_device(targets, probs[:, 1])
    if num_classes == 2 and threshold is not None:
        out["threshold"] = float(threshold)
    out["confusion_matrix"] = cm.tolist()
    return out


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
