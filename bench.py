#!/usr/bin/env python
"""bench.py -- EfficientNet-B0 224x224 eval images/s on N B200s (BASELINE.json metric, configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch 256] [--workload eval|cam] [--model NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the reference's CPU path (oracle port) on the host cores

The default line is BASELINE.json's headline (configs[1]); it also carries, as EXTRA keys measured in the same run:
`cam` (configs[2]: Grad-CAM batch 64, maps/s), `models` (configs[3], [4]: EfficientFormerV2-S1 and FasterViT-0 batch-256
eval), `strong` (global batch 256 split over the N GPUs), `latency_n1_ms` / `small_batch` (CUDA-graph replay),
`sustained` (>= 2 s of steps), `fp32` (the strict path).  `--model` makes any served model the headline of its own line.

One "step" = one pass of the hot path (`logits = model(images)`, orchestrator.py:587-592, plus the
softmax/argmax pack) over one synthetic batch of `--batch` images per GPU.  Prints ONE JSON line:

  value      whole-job images/s with the input batch already resident in HBM (CUDA events, max over ranks)
  e2e        the same metric through the public nn.Module call with HOST (pinned) fp32 NCHW batches:
             H2D copy of every batch and D2H of its probs/preds inside the timed region
  roofline   dominant kernel class: algorithmic bytes / CUDA-event time vs MEASURED_PEAKS.json
  cpu_baseline  the oracle port (oracle/effnet_b0.py, fp32 eager) on the host cores, bounded sample

Only the `cpu_baseline` leg and `--impl reference` touch oracle/ (as the thing compared against).
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

METRIC = "efficientnet_b0_224_eval_images_per_s"
UNIT = "images/s"
# Algorithmic HBM bytes per image of the bf16 engine (DESIGN.md section 4 / SURVEY.md 8d): every
# activation written once and read once by its consumer, fp32 NCHW network input.
ALGO_MB_PER_IMG = 27.87
WEIGHT_MB = 8.0

# served model -> (oracle module, oracle class factory) for the CPU legs, and how BASELINE.json refers to it
MODELS = {
    "efficientnet_b0": "BASELINE.json configs[1]",
    "efficientnet_b3": "reference registry default, efficientnet_pytorch layout; SURVEY 8f-2",
    "efficientformerv2_s1": "BASELINE.json configs[3]",
    "faster_vit_0_224": "BASELINE.json configs[4]",
    "faster_vit_2_224": "reference registry default `faster_vit` entry; SURVEY 8f-2",
}


def oracle_for(arch: str):
    """The fp32 CPU restatement of `arch` (cpu_baseline / reference arm only: the checker, never the product)."""
    if arch == "efficientnet_b3":
        from oracle.effnet_b3 import EfficientNetB3Oracle
        return EfficientNetB3Oracle(2)
    if arch == "efficientformerv2_s1":
        from oracle.efficientformerv2 import EfficientFormerV2Oracle
        return EfficientFormerV2Oracle(2)
    if arch.startswith("faster_vit"):
        from oracle.fastervit import FasterViTOracle
        return FasterViTOracle(2, arch)
    from oracle.effnet_b0 import EfficientNetB0Oracle
    return EfficientNetB0Oracle(2)


def oracle_file(arch: str) -> str:
    return {"efficientnet_b0": "effnet_b0", "efficientnet_b3": "effnet_b3", "efficientformerv2_s1": "efficientformerv2"}.get(arch, "fastervit")


def parse_args() -> argparse.Namespace:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=256, help="images per GPU per step")
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--workload", choices=["eval", "cam"], default="eval")
    ap.add_argument("--precision", choices=["bf16", "fp32"], default="bf16")
    ap.add_argument("--model", choices=sorted(MODELS), default="efficientnet_b0",
                    help="efficientnet_b0 = BASELINE.json's headline (timm layout, configs[1]); efficientformerv2_s1 / "
                         "faster_vit_0_224 = configs[3] / [4]; efficientnet_b3 / faster_vit_2_224 = the reference "
                         "registry's own entries (SURVEY 8f-2)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the extra keys (cam, models, strong, latency, ...)")
    return ap.parse_args()


# ---- clocks ------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int) -> None:
        self.index = index
        self.rows: list[list[str]] = []
        self.proc = None

    def start(self) -> None:
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index), "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self) -> None:
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def mark(self) -> int:
        return len(self.rows)

    def stop(self) -> None:
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self, lo: int = 0, hi: int | None = None) -> dict:
        rows = [r for r in self.rows[lo:hi] if len(r) >= 7] or [r for r in self.rows if len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for j, nm in enumerate(names) if any(r[3 + j].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": reasons,
                "power_w_max": max(float(r[2]) for r in rows), "samples": len(rows)}


# ---- synthetic model / data (SURVEY.md 8d recipe; no oracle import here) ------------------------------
def synth_state(model, seed: int = 0) -> None:
    """Module default init (timm's scheme) + randomised BN stats/affine so folding is exercised."""
    import torch
    from torch import nn

    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for m in model.modules():
            if isinstance(m, nn.BatchNorm2d):
                c = m.num_features
                m.running_mean.copy_(torch.randn(c, generator=g) * 0.1)
                m.running_var.copy_(torch.rand(c, generator=g) + 0.5)
                m.weight.copy_(torch.rand(c, generator=g) + 0.5)
                m.bias.copy_(torch.randn(c, generator=g) * 0.1)
        for name, p in model.named_parameters():   # LayerScale 1e-5 / zero attention biases would hide whole branches
            if name.endswith(".gamma"):
                p.copy_(torch.rand(p.shape, generator=g) * 0.4 + 0.1)
            elif name.endswith("attention_biases"):
                p.copy_(torch.randn(p.shape, generator=g))


_CLASS_KERNEL = {"dw": "dwconv", "expand": "pwconv_tc_kernel<0, 1, 0>", "project": "pwconv_tc_kernel<1", "stem": "stem_tc",
                 "se": "se_kernel", "conv_head": "pwconv_tc_kernel<0, 0, 0>", "head_pool": "head_pool", "classifier": "classifier"}


def ncu_traffic_per_launch(cls: str):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch (bytes) of a kernel class, averaged over its
    launches in the newest committed `ncu --set full` capture of one step (profiles/*_step_metrics.txt)."""
    files = sorted((ROOT / "profiles").glob("*_step_metrics.txt"), key=lambda f: f.name)
    files = [f for f in files if not f.name.endswith(("_b3.txt", "_efv2.txt", "_fv0.txt"))]   # B0 captures only
    key = _CLASS_KERNEL.get(cls)
    if not files or key is None:
        return None, None
    # the capture must describe the kernels that are being timed: its header carries the source hash it profiled
    # (tools/ncu_step.py writes `# kernels-sha: <sha16 of csrc>`); a capture of other sources is not reported
    head = files[-1].read_text().splitlines()[:3]
    stamp = next((ln.split(":", 1)[1].strip() for ln in head if ln.startswith("# kernels-sha:")), None)
    if stamp != kernels_sha():
        return None, f"{files[-1].name} is stale (kernels-sha {stamp} != {kernels_sha()})"
    tot, cnt = 0.0, 0
    for line in files[-1].read_text().splitlines():
        if line.startswith("#") or key not in line:
            continue
        rest = line.split(key, 1)[1]
        rest = rest.split(">", 1)[1] if ">" in rest[:12] else rest
        nums = [t for t in rest.split() if t.replace(".", "", 1).isdigit()]
        if len(nums) >= 3:
            tot += float(nums[1]) + float(nums[2])
            cnt += 1
    return (tot / cnt * 1e6, files[-1].name) if cnt else (None, None)


def kernels_sha() -> str:
    """sha16 over the device sources of the EfficientNet path: identifies the kernels an ncu capture belongs to."""
    import hashlib

    h = hashlib.sha256()
    for name in ("act.h", "common.cuh", "tc.cuh", "stem.cu", "stem_tc.cu", "dwconv.cu", "dwconv_tma.cu", "se.cu", "pwconv_simt.cu",
                 "pwconv_tc.cu", "tail.cu"):
        f = ROOT / "deepfakedetection_b200" / "csrc" / name
        h.update(name.encode())
        h.update(f.read_bytes())
    return h.hexdigest()[:16]


def peaks() -> tuple[dict, str]:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return json.loads(p.read_text()), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def cpu_eval_images_per_s(state: dict, batch: int, steps: int, warmup: int, seed: int = 0,
                          arch: str = "efficientnet_b0", channels_last: bool = False):
    """The reference eval-loop body (orchestrator.py:587-592) on the fp32 oracle port, host cores."""
    import torch

    torch.set_num_threads(os.cpu_count() or 1)   # torch.distributed.run exports OMP_NUM_THREADS=1: undo it for the CPU arm
    model = oracle_for(arch).eval()
    model.load_state_dict(state, strict=False)
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(batch, 3, 224, 224, generator=g)
    if channels_last:   # what the reference's trainers use (trainers/efficientnet.py:250-252); second reported row
        model = model.to(memory_format=torch.channels_last)
        x = x.contiguous(memory_format=torch.channels_last)
    times = []
    with torch.inference_mode():
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            logits = model(x)
            probs = torch.softmax(logits, dim=1)
            preds = torch.argmax(probs, dim=1)
            _ = probs.cpu(), preds.cpu()
            if i >= warmup:
                times.append(time.perf_counter() - t0)
    total = sum(times)
    return batch * len(times) / total, total / len(times) * 1e3, torch.get_num_threads()


def run_reference(args) -> None:
    """--impl reference: the reference's own CPU path for this metric.  The reference's arithmetic
    lives in timm / efficientnet_pytorch / fastervit (absent, not installable here), so this is the oracle port
    (kind "port"), on ALL host cores whatever launcher started us (rank 0 only under torchrun)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from deepfakedetection_b200 import registry

    m = registry.get_model_spec(args.model + ".b200_fp32").builder(args.model + ".b200_fp32", 2)
    synth_state(m)
    sample_batch = 64 if args.model.startswith("efficientnet") else 32
    ips, ms, cores = cpu_eval_images_per_s(m.state_dict(), sample_batch, args.steps, min(args.warmup, 3),
                                           arch=args.model)
    line = {
        "impl": "reference", "metric": METRIC.replace("efficientnet_b0", args.model), "value": round(ips, 2), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": min(args.warmup, 3), "ms_per_step": round(ms, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.model} eval forward 224x224, 2-class head, fp32 eager on host cores",
                   "batch_per_step": sample_batch, "torch_threads": cores},
        "cpu_baseline": {"value": round(ips, 2), "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{args.steps} steps x {sample_batch} images, oracle/{oracle_file(args.model)}.py fp32 eager "
                                   f"(third-party model code not installable; os.cpu_count={os.cpu_count()})"},
        "e2e": {"value": round(ips, 2), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def bind_to_gpu_numa_node(local: int) -> None:
    """Pin this rank's host threads to the CPUs next to its GPU (NVML's ideal-CPU mask) BEFORE any pinned buffer is
    allocated, so first-touch puts the staging memory on the GPU's NUMA node and H2D copies do not cross sockets.
    Host plumbing only; silently skipped where NVML or the cpuset does not allow it."""
    try:
        import pynvml

        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = int(vis.split(",")[local]) if vis and all(t.strip().isdigit() for t in vis.split(",")) else local
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64 + 4)
        cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1}
        allowed = os.sched_getaffinity(0)
        want = cpus & allowed
        if want and want != allowed:
            os.sched_setaffinity(0, want)
        print(f"[bench] rank-local GPU {idx}: ideal CPUs {len(cpus)}, allowed {len(allowed)}, bound to {len(want) or len(allowed)}",
              file=sys.stderr)
    except Exception as ex:  # noqa: BLE001 - optional
        print(f"[bench] NUMA binding skipped: {ex}", file=sys.stderr)


# ---- our arm ----------------------------------------------------------------------------------------------
def run_b200(args) -> None:
    import torch
    import torch.distributed as dist

    from deepfakedetection_b200 import _lib, registry
    from deepfakedetection_b200.evaluate import gather_records, pack_records

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or _lib.lib().b200df_device_count() == 0:
        raise SystemExit("bench.py: no sm_100 device visible -- the b200 backend has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        bind_to_gpu_numa_node(local)   # N = 1 keeps every host core for the cpu_baseline leg
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL prints its version banner to STDOUT (fd 1, from C) when the communicator comes up, whatever
        # NCCL_DEBUG says on this build: park the real stdout and point fd 1 at stderr until the JSON line is written
        global _JSON_FD
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)

    name = args.model if args.precision == "bf16" else args.model + ".b200_fp32"
    model = registry.get_model_spec(name).builder(name, 2)   # the reference's builder(name, num_classes) call
    synth_state(model)
    model.to(dev).eval()
    B, K, W = args.batch, args.steps, max(args.warmup, 3)
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    images = torch.randn(B, 3, 224, 224, device=dev, generator=g)          # 154 MB at B=256 (> 126 MB L2)
    labels = torch.randint(0, 2, (B,), device=dev, generator=g)
    cam_mode = args.workload == "cam"

    def step(x):
        if cam_mode:
            logits, low, cam = model.forward_cam(x, None)
            return logits
        return model(x)

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier(device_ids=[local])
            torch.cuda.synchronize(dev)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    with torch.inference_mode():
        for _ in range(W):
            rec = pack_records(step(images), labels)
        # the collective is part of the timed region: warm it (communicator set-up, first-use buffers) like the kernels
        for _ in range(2):
            gather_records(torch.cat([rec] * K, 0), equal_shards=True)
        sync_all()
        # ---- timed region 1: device-resident inputs -> `value` ----
        _lib.reset_launch_count()
        ev0, ev1, evc = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        lo = sampler.mark()
        ev0.record()
        recs = []
        for _ in range(K):
            recs.append(pack_records(step(images), labels))
        evc.record()
        # the one collective: all-gather of probs/preds/labels (every rank holds K*B records: no size exchange)
        allrec = gather_records(torch.cat(recs, 0), equal_shards=True)
        ev1.record()
        sync_all()
        hi = sampler.mark()
        launches = _lib.launch_count() + (1 if world > 1 else 0)
        ms = torch.tensor([ev0.elapsed_time(ev1), ev0.elapsed_time(evc), evc.elapsed_time(ev1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms_total, ms_compute, ms_gather = (float(v) for v in ms.tolist())
        assert allrec.shape[0] == world * K * B
        # keep the clock sampler honest for very short runs
        while rank == 0 and sampler.proc is not None and sampler.mark() - lo < 3 and hi - lo < 3:
            _ = step(images)
            torch.cuda.synchronize(dev)
            hi = sampler.mark()

        # ---- timed region 2: host (pinned) inputs through the public module call -> `e2e` ----
        e2e = e2e_u8 = None
        if not args.no_e2e:
            e2e = run_e2e(model, step, B, K, W, dev, rank, world, local)
            if not cam_mode:
                e2e_u8 = run_e2e(model, step, B, K, W, dev, rank, world, local, u8=True)

        # ---- profiling pass (events around every launch; outside the timed regions) ----
        eng = model._engine
        per_class: dict[str, list[float]] = {}
        reps = 5
        eng.set_profile(True)
        for _ in range(reps):
            step(images)
        torch.cuda.synchronize(dev)
        for nm, t, by, fl in eng.profile_records():
            cls = nm.split(".")[0]
            a = per_class.setdefault(cls, [0.0, 0.0, 0.0, 0])
            a[0] += t; a[1] += by; a[2] += fl; a[3] += 1
        eng.set_profile(False)

        extras = {}
        if not args.no_extras and args.model == "efficientnet_b0" and args.workload == "eval" and args.precision == "bf16":
            extras = run_extras(args, model, images, labels, dev, rank, world, local, sync_all, sampler)
    sampler.stop()
    if world > 1:
        dist.barrier(device_ids=[local])

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    finish_line(args, world, B, K, W, ms_total, launches, e2e, per_class, reps, sampler.summary(lo, hi), model, e2e_u8,
                ms_compute=ms_compute, ms_gather=ms_gather, extras=extras)
    if world > 1:
        dist.destroy_process_group()


def timed_steps(fn, k, dev, world, local, warm: int = 3) -> float:
    """ms for k calls of fn() (CUDA events on the current stream, barrier + sync on both sides, max over ranks)."""
    import torch
    import torch.distributed as dist

    for _ in range(warm):
        fn()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier(device_ids=[local])
        torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def class_table(eng, fn, dev, reps: int = 3):
    """Per-kernel-class CUDA-event times / algorithmic bytes / flops of one call of fn() (profiling pass)."""
    import torch

    per = {}
    eng.set_profile(True)
    for _ in range(reps):
        fn()
    torch.cuda.synchronize(dev)
    for nm, t, by, fl in eng.profile_records():
        a = per.setdefault(nm.split(".")[0], [0.0, 0.0, 0.0, 0])
        a[0] += t; a[1] += by; a[2] += fl; a[3] += 1
    eng.set_profile(False)
    return {c: {"ms": round(t / reps, 4), "launches": cnt // reps,
                "GBps": round(by / t / 1e6, 1) if t > 0 else None, "TFLOPs": round(fl / t / 1e9, 2) if t > 0 else None,
                "bytes": by / reps, "flops": fl / reps}
            for c, (t, by, fl, cnt) in per.items()}


def roofline_of(table: dict, pk: dict) -> dict:
    """Dominant class of a per-class table against whichever measured roof (HBM copy rate / dense bf16 rate) it is nearer to."""
    dom = max(table, key=lambda c: table[c]["ms"])
    r = table[dom]
    t = r["ms"] * 1e-3
    hbm = r["bytes"] / t / 1e9 / float(pk["hbm_gbs"])
    tens = r["flops"] / t / 1e12 / float(pk.get("bf16_tflops_sustained", pk["bf16_tflops"]))
    tot = sum(v["ms"] for v in table.values())
    if tens > hbm:
        return {"kernel": dom, "bound": "tensor", "achieved": round(r["flops"] / t / 1e12, 1), "unit": "TFLOP/s",
                "peak": float(pk.get("bf16_tflops_sustained", pk["bf16_tflops"])), "frac": round(tens, 4),
                "frac_of_hbm": round(hbm, 4), "share_of_step": round(r["ms"] / tot, 3), "launches_per_step": r["launches"]}
    return {"kernel": dom, "bound": "hbm", "achieved": round(r["bytes"] / t / 1e9, 1), "unit": "GB/s", "peak": float(pk["hbm_gbs"]),
            "frac": round(hbm, 4), "frac_of_tensor": round(tens, 4), "share_of_step": round(r["ms"] / tot, 3),
            "launches_per_step": r["launches"]}


def run_extras(args, model, images, labels, dev, rank, world, local, sync_all, sampler) -> dict:
    """Extra keys of the default line (see the module docstring).  Every number: CUDA events, >= 3 warm-ups, max over ranks."""
    import torch

    from deepfakedetection_b200 import registry
    from deepfakedetection_b200.evaluate import pack_records

    pk, _ = peaks()
    out: dict = {}
    B = images.shape[0]

    # strong scaling: BASELINE configs[1]'s global batch 256 split over the N GPUs; CUDA-graph replay below 65
    bs = max(1, 256 // world)
    xs = images[:bs]
    ms = timed_steps(lambda: pack_records(model(xs), labels[:bs]), 50, dev, world, local)
    out["strong"] = {"global_batch": bs * world, "batch_per_gpu": bs, "value": round(bs * world * 50 / (ms * 1e-3), 1), "unit": UNIT,
                     "ms_per_step": round(ms / 50, 4), "cuda_graph": bs <= model._engine.graph_max_n}
    # the UI's regime: one image (web_ui.py:251-268), and 32 images (8-GPU strong scaling's per-GPU share)
    x1, x32 = images[:1], images[:32]
    ms1 = timed_steps(lambda: model(x1), 200, dev, 1, local)
    out["latency_n1_ms"] = round(ms1 / 200, 4)
    ms32 = timed_steps(lambda: model(x32), 100, dev, 1, local)
    out["small_batch"] = {"batch": 32, "value": round(32 * 100 / (ms32 * 1e-3), 1), "unit": UNIT, "ms_per_step": round(ms32 / 100, 4)}
    keep = model._engine.graph_max_n
    model._engine.graph_max_n = 0
    ms1d = timed_steps(lambda: model(x1), 100, dev, 1, local)
    ms32d = timed_steps(lambda: model(x32), 50, dev, 1, local)
    model._engine.graph_max_n = keep
    out["latency_n1_ms_no_graph"] = round(ms1d / 100, 4)
    out["small_batch"]["ms_per_step_no_graph"] = round(ms32d / 50, 4)

    # >= 2 s of back-to-back steps: the rate once the burst clock is gone
    lo = sampler.mark()
    k = max(100, int(2200.0 / max(timed_steps(lambda: model(images), 10, dev, world, local) / 10, 1e-3)))
    ms = timed_steps(lambda: pack_records(model(images), labels), k, dev, world, local, warm=0)
    hi = sampler.mark()
    out["sustained"] = {"steps": k, "seconds": round(ms * 1e-3, 2), "value": round(world * B * k / (ms * 1e-3), 1), "unit": UNIT,
                        "clocks": sampler.summary(lo, hi) if rank == 0 else None}

    # configs[2]: Grad-CAM, batch 64, one GPU's worth per rank (forward + fused CAM reduction + 224^2 map)
    x64 = images[:64]
    ms = timed_steps(lambda: model.forward_cam(x64, None), 50, dev, world, local)
    cam = {"metric": "efficientnet_b0_gradcam_maps_per_s", "value": round(world * 64 * 50 / (ms * 1e-3), 1), "unit": "maps/s",
           "batch_per_gpu": 64, "ms_per_step": round(ms / 50, 4), "config": "BASELINE.json configs[2]: forward + fused Grad-CAM, batch 64"}
    # the UI's regime (web_ui.py:241-306): one image, forward + Grad-CAM
    cam["latency_n1_ms"] = round(timed_steps(lambda: model.forward_cam(x1, None), 100, dev, 1, local) / 100, 4)
    tab = class_table(model._engine, lambda: model.forward_cam(x64, None), dev)
    cam["kernels"] = {c: {k2: v[k2] for k2 in ("ms", "launches", "GBps")} for c, v in tab.items() if c in ("cam", "cam_upsample", "conv_head", "dw")}
    out["cam"] = cam

    # strict fp32 path throughput (the <= 1e-3 logit path), batch 256
    m32 = registry.get_model_spec("efficientnet_b0.b200_fp32").builder("efficientnet_b0.b200_fp32", 2)
    synth_state(m32)
    m32.to(dev).eval()
    ms = timed_steps(lambda: m32(images), 5, dev, world, local)
    out["fp32"] = {"value": round(world * B * 5 / (ms * 1e-3), 1), "unit": UNIT, "ms_per_step": round(ms / 5, 3)}
    del m32
    torch.cuda.empty_cache()

    # configs[3], configs[4]: the other two backbones, bf16 batch 256 per GPU
    models = {}
    for name in ("efficientformerv2_s1", "faster_vit_0_224"):
        m = registry.get_model_spec(name).builder(name, 2)
        synth_state(m)
        m.to(dev).eval()
        ms = timed_steps(lambda: pack_records(m(images), labels), 10, dev, world, local)
        tab = class_table(m._engine, lambda: m(images), dev)
        algo_gb = sum(v["bytes"] for v in tab.values()) / 1e9
        gflop = sum(v["flops"] for v in tab.values()) / 1e9
        models[name] = {
            "metric": f"{name}_224_eval_images_per_s", "value": round(world * B * 10 / (ms * 1e-3), 1), "unit": UNIT,
            "ms_per_step": round(ms / 10, 4), "batch_per_gpu": B, "dtype": "bf16", "config": MODELS[name],
            "roofline": roofline_of(tab, pk),
            "net": {"algorithmic_GB_per_step": round(algo_gb, 3), "GFLOP_per_step": round(gflop, 1),
                    "frac_of_hbm_peak": round(algo_gb / (ms / 10 * 1e-3) / float(pk["hbm_gbs"]), 4),
                    "frac_of_tensor_peak": round(gflop / 1e3 / (ms / 10 * 1e-3) / float(pk.get("bf16_tflops_sustained", pk["bf16_tflops"])), 4)},
            "kernels": {c: {k2: v[k2] for k2 in ("ms", "launches", "GBps", "TFLOPs")} for c, v in sorted(tab.items(), key=lambda kv: -kv[1]["ms"])},
        }
        # Grad-CAM on this backbone at configs[2]'s batch (64): EfficientFormerV2's closed form / FasterViT's backward program
        msc = timed_steps(lambda: m.forward_cam(x64, None), 20, dev, world, local)
        tabc = class_table(m._engine, lambda: m.forward_cam(x64, None), dev)
        models[name]["cam"] = {
            "metric": f"{name}_gradcam_maps_per_s", "value": round(world * 64 * 20 / (msc * 1e-3), 1), "unit": "maps/s",
            "batch_per_gpu": 64, "ms_per_step": round(msc / 20, 4),
            "latency_n1_ms": round(timed_steps(lambda: m.forward_cam(x1, None), 50, dev, 1, local) / 50, 4),
            "kernels": {c: {k2: v[k2] for k2 in ("ms", "launches", "GBps", "TFLOPs")}
                        for c, v in sorted(tabc.items(), key=lambda kv: -kv[1]["ms"]) if c.startswith("cam")}}
        del m
        torch.cuda.empty_cache()
    out["models"] = models
    return out


def run_e2e(model, step, B, K, W, dev, rank, world, local, u8: bool = False):
    """Host-resident fp32 NCHW batches (what the reference's DataLoader hands over, pin_memory=True,
    orchestrator.py:386-395,588): every step copies its batch H2D (the package's HostBatchFeeder: two halves on
    two copy streams, double-buffered against the compute stream -- `.to(device, non_blocking=True)` with both
    copy engines busy), runs the module, and reads the packed
    probs/preds back D2H.  Timed with CUDA events on the compute stream, max over ranks."""
    import torch
    import torch.distributed as dist

    from deepfakedetection_b200.evaluate import pack_records

    if u8:  # GPU-side input path: uint8 HWC crops, ToTensor + Normalize fused into the stem
        host = [torch.randint(0, 256, (B, 224, 224, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
        step = model.forward_u8
    else:
        host = [torch.randn(B, 3, 224, 224).pin_memory() for _ in range(2)]
    in_bytes = host[0].numel() * host[0].element_size()
    out_host = torch.empty(K + W, B, 4, dtype=torch.float32).pin_memory()
    from deepfakedetection_b200.evaluate import HostBatchFeeder

    feeder = HostBatchFeeder(host[0], dev, parts=2)   # the package's own H2D path: two halves on two copy streams
    comp_s = torch.cuda.current_stream(dev)

    def loop(n, base):
        for i in range(n):
            x = feeder.feed(i, host[i & 1])
            rec = pack_records(step(x), None)
            feeder.release(i)
            out_host[base + i].copy_(rec, non_blocking=True)

    loop(W, 0)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier(device_ids=[local])
        torch.cuda.synchronize(dev)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(comp_s)
    feeder.wait_event_on_copy_streams(ev0)
    loop(K, W)
    ev1.record(comp_s)
    torch.cuda.synchronize(dev)
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    assert bool(torch.isfinite(out_host[W:]).all())
    return {"value": round(world * B * K / (ms * 1e-3), 1), "unit": UNIT,
            "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": B * 4 * 4,
            "ms_per_step": round(ms / K, 4),
            "note": ("pinned uint8 HWC host batches (forward_u8)" if u8 else "pinned fp32 NCHW host batches") +
                    ", H2D in two halves on two copy streams (evaluate.HostBatchFeeder), double-buffered with compute; h2d GB/s per GPU = %.1f"
                    % (in_bytes * K / (ms * 1e-3) / 1e9)}


_JSON_FD = None  # the real stdout while fd 1 is parked on stderr (multi-GPU runs)


def emit(line: dict) -> None:
    text = json.dumps(line) + "\n"
    if _JSON_FD is None:
        sys.stdout.write(text)
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_JSON_FD, text.encode())


def finish_line(args, world, B, K, W, ms_total, launches, e2e, per_class, reps, clocks, model, e2e_u8=None,
                ms_compute=None, ms_gather=None, extras=None) -> None:
    import torch

    pk, pk_kind = peaks()
    ms_step = ms_total / K
    value = world * B * K / (ms_total * 1e-3)
    # per-class table from the profiling pass (per step)
    table = {}
    for cls, (t, by, fl, cnt) in per_class.items():
        table[cls] = {"ms": round(t / reps, 4), "launches": cnt // reps,
                      "GBps": round(by / reps / (t / reps * 1e-3) / 1e9, 1) if t > 0 else None,
                      "TFLOPs": round(fl / reps / (t / reps * 1e-3) / 1e12, 2) if t > 0 else None}
    prof_ms = sum(v["ms"] for v in table.values())
    dom = max(per_class, key=lambda c: per_class[c][0])
    t, by, fl, cnt = per_class[dom]
    t_launch = t / cnt * 1e-3
    hbm_peak = float(pk["hbm_gbs"])
    achieved = by / cnt / t_launch / 1e9
    traffic, traffic_src = ncu_traffic_per_launch(dom) if args.model == "efficientnet_b0" else (None, None)  # ncu capture is B0's
    tens_peak = float(pk.get("bf16_tflops_sustained", pk["bf16_tflops"]))
    tens_frac = fl / cnt / t_launch / 1e12 / tens_peak
    roofline = {"kernel": dom, "bound": "hbm", "achieved": round(achieved, 1), "peak": hbm_peak, "unit": "GB/s",
                "frac": round(achieved / hbm_peak, 4), "traffic": None if traffic is None else round(traffic),
                "traffic_source": traffic_src, "algorithmic_bytes_per_launch": round(by / cnt),
                "peak_source": f"{pk_kind} (MEASURED_PEAKS.json hbm_gbs)",
                "share_of_step": round(t / reps / prof_ms, 3), "launches_per_step": cnt // reps,
                "avg_launch_ms": round(t_launch * 1e3, 4), "frac_of_tensor": round(tens_frac, 4)}
    if tens_frac > achieved / hbm_peak:   # a tensor-pipe-bound dominant kernel (FasterViT's dense 3x3 convs / MLP GEMMs)
        roofline.update({"bound": "tensor", "achieved": round(fl / cnt / t_launch / 1e12, 1), "peak": tens_peak, "unit": "TFLOP/s",
                         "frac": round(tens_frac, 4), "frac_of_hbm": round(achieved / hbm_peak, 4),
                         "peak_source": f"{pk_kind} (MEASURED_PEAKS.json bf16_tflops_sustained)"})
    es = 2 if args.precision == "bf16" else 4
    algo_gb = (ALGO_MB_PER_IMG * (es / 2) * B + WEIGHT_MB * (es / 2)) / 1e3
    if args.model != "efficientnet_b0":  # sum of the per-launch algorithmic bytes the engine tags its launches with
        algo_gb = sum(v[1] for v in per_class.values()) / reps / 1e9
    net = {"algorithmic_GB_per_step": round(algo_gb, 3), "achieved_GBps": round(algo_gb / (ms_step * 1e-3), 1),
           "frac_of_hbm_peak": round(algo_gb / (ms_step * 1e-3) / hbm_peak, 4),
           "profiled_ms_per_step": round(prof_ms, 4)}
    line = {
        "metric": METRIC.replace("efficientnet_b0", args.model), "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": args.precision, "data": "synthetic",
        "config": {"workload": (f"{args.model} eval forward" if args.workload == "eval" else f"{args.model} forward + fused Grad-CAM")
                   + " 224x224, 2-class head (" + MODELS[args.model] + ")", "batch_per_gpu": B, "global_batch": B * world,
                   "input": "float32 NCHW [B,3,224,224] resident in HBM", "l2": "inputs (154 MB/batch) and per-step "
                   "activations (>3 GB) exceed the 126 MB L2; no explicit flush",
                   "collective": "one all_gather of packed probs/preds/labels after the K steps (inside the timed region, warmed)",
                   "weights": "random init + randomised BN (synthetic)"},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "roofline_net": net, "kernels": table,
    }
    if ms_compute is not None:
        line["compute_ms_per_step"] = round(ms_compute / K, 4)
        line["gather_ms"] = round(ms_gather, 4)
    if e2e is not None:
        line["e2e"] = e2e
    if e2e_u8 is not None:  # extra, not the headline: same metric fed with uint8 crops (4x fewer PCIe bytes)
        line["e2e_u8"] = e2e_u8
    if not args.no_cpu_baseline and world == 1:  # rank 0 at N=1 only
        steps_cpu = 6 if args.model.startswith("efficientnet") else 3
        bs_cpu = 64 if args.model.startswith("efficientnet") else 32
        state = {k: v.cpu() for k, v in model.state_dict().items()}
        ips, ms, cores = cpu_eval_images_per_s(state, bs_cpu, steps_cpu, 2, arch=args.model)
        line["cpu_baseline"] = {"value": round(ips, 2), "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{steps_cpu} timed steps x {bs_cpu} images (+2 warm-up), oracle/{oracle_file(args.model)}.py fp32 eager, "
                                          f"torch threads={cores}, os.cpu_count={os.cpu_count()}"}
        if args.model == "efficientnet_b0":  # BASELINE.md section 2's second row
            ips_cl, _, _ = cpu_eval_images_per_s(state, bs_cpu, 4, 1, arch=args.model, channels_last=True)
            line["cpu_baseline"]["channels_last"] = round(ips_cl, 2)
    for k2, v in (extras or {}).items():
        line[k2] = v
    emit(line)


def main() -> None:
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
