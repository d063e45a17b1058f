#!/usr/bin/env python
"""bench.py -- EfficientNet-B0 224x224 eval images/s on N B200s (BASELINE.json metric, configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch 256] [--workload eval|cam]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the reference's CPU path (oracle port) on the host cores

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


def parse_args() -> argparse.Namespace:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=256, help="images per GPU per step")
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--workload", choices=["eval", "cam"], default="eval")
    ap.add_argument("--precision", choices=["bf16", "fp32"], default="bf16")
    ap.add_argument("--model", choices=["efficientnet_b0", "efficientnet_b3"], default="efficientnet_b0",
                    help="efficientnet_b0 = BASELINE.json's headline (timm layout); efficientnet_b3 = the reference "
                         "registry's default entry (efficientnet_pytorch layout, SURVEY 8f-2), an extra line")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
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


_CLASS_KERNEL = {"dw": "dwconv", "expand": "pwconv_tc_kernel<0, 1, 0>", "project": "pwconv_tc_kernel<1", "stem": "stem_tc",
                 "se": "se_kernel", "conv_head": "pwconv_tc_kernel<0, 0, 0>", "head_pool": "head_pool", "classifier": "classifier"}


def ncu_traffic_per_launch(cls: str):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch (bytes) of a kernel class, averaged over its
    launches in the newest committed `ncu --set full` capture of one step (profiles/*_step_metrics.txt)."""
    files = sorted((ROOT / "profiles").glob("*_step_metrics.txt"))
    key = _CLASS_KERNEL.get(cls)
    if not files or key is None:
        return None, None
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


def peaks() -> tuple[dict, str]:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return json.loads(p.read_text()), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def cpu_eval_images_per_s(state: dict, batch: int, steps: int, warmup: int, seed: int = 0,
                          arch: str = "efficientnet_b0"):
    """The reference eval-loop body (orchestrator.py:587-592) on the fp32 oracle port, host cores."""
    import torch
    from oracle.effnet_b0 import EfficientNetB0Oracle  # the checker: cpu_baseline / reference arm only
    from oracle.effnet_b3 import EfficientNetB3Oracle

    model = (EfficientNetB3Oracle(2) if arch == "efficientnet_b3" else EfficientNetB0Oracle(2)).eval()
    model.load_state_dict(state, strict=False)
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(batch, 3, 224, 224, generator=g)
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
    lives in timm (absent, not installable here), so this is the oracle port (kind "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch

    from deepfakedetection_b200.module import B200EfficientNetB0, B200EfficientNetB3

    m = (B200EfficientNetB3 if args.model == "efficientnet_b3" else B200EfficientNetB0)(2, "fp32")
    synth_state(m)
    sample_batch = 64
    ips, ms, cores = cpu_eval_images_per_s(m.state_dict(), sample_batch, args.steps, min(args.warmup, 3),
                                           arch=args.model)
    line = {
        "impl": "reference", "metric": METRIC.replace("efficientnet_b0", args.model), "value": round(ips, 2), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": min(args.warmup, 3), "ms_per_step": round(ms, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.model} eval forward 224x224, 2-class head, fp32 eager on host cores",
                   "batch_per_step": sample_batch, "torch_threads": cores},
        "cpu_baseline": {"value": round(ips, 2), "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{args.steps} steps x {sample_batch} images, oracle/{args.model.replace('efficientnet', 'effnet')}.py fp32 eager "
                                   f"(timm not installable; os.cpu_count={os.cpu_count()})"},
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
        sync_all()
        # ---- timed region 1: device-resident inputs -> `value` ----
        _lib.reset_launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        lo = sampler.mark()
        ev0.record()
        recs = []
        for _ in range(K):
            recs.append(pack_records(step(images), labels))
        allrec = gather_records(torch.cat(recs, 0))    # the one collective: all-gather of probs/preds/labels
        ev1.record()
        sync_all()
        hi = sampler.mark()
        launches = _lib.launch_count() + (2 if world > 1 else 0)
        ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms_total = float(ms.item())
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
    sampler.stop()
    if world > 1:
        dist.barrier(device_ids=[local])

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    finish_line(args, world, B, K, W, ms_total, launches, e2e, per_class, reps, sampler.summary(lo, hi), model, e2e_u8)
    if world > 1:
        dist.destroy_process_group()


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


def finish_line(args, world, B, K, W, ms_total, launches, e2e, per_class, reps, clocks, model, e2e_u8=None) -> None:
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
    roofline = {"kernel": dom, "bound": "hbm", "achieved": round(achieved, 1), "peak": hbm_peak, "unit": "GB/s",
                "frac": round(achieved / hbm_peak, 4), "traffic": None if traffic is None else round(traffic),
                "traffic_source": traffic_src, "algorithmic_bytes_per_launch": round(by / cnt),
                "peak_source": f"{pk_kind} (MEASURED_PEAKS.json hbm_gbs)",
                "share_of_step": round(t / reps / prof_ms, 3), "launches_per_step": cnt // reps,
                "avg_launch_ms": round(t_launch * 1e3, 4)}
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
                   + " 224x224, 2-class head" + (" (BASELINE.json configs[1])" if args.model == "efficientnet_b0" else
                                                 " (reference registry default, efficientnet_pytorch layout; SURVEY 8f-2)"), "batch_per_gpu": B, "global_batch": B * world,
                   "input": "float32 NCHW [B,3,224,224] resident in HBM", "l2": "inputs (154 MB/batch) and per-step "
                   "activations (>3 GB) exceed the 126 MB L2; no explicit flush",
                   "collective": "one all_gather of packed probs/preds/labels after the K steps (inside the timed region)",
                   "weights": "random init + randomised BN (synthetic)"},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "roofline_net": net, "kernels": table,
    }
    if e2e is not None:
        line["e2e"] = e2e
    if e2e_u8 is not None:  # extra, not the headline: same metric fed with uint8 crops (4x fewer PCIe bytes)
        line["e2e_u8"] = e2e_u8
    if not args.no_cpu_baseline and world == 1:  # rank 0 at N=1 only
        steps_cpu = 6
        ips, ms, cores = cpu_eval_images_per_s({k: v.cpu() for k, v in model.state_dict().items()}, 64, steps_cpu, 2,
                                               arch=args.model)
        line["cpu_baseline"] = {"value": round(ips, 2), "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{steps_cpu} timed steps x 64 images (+2 warm-up), oracle/{args.model.replace('efficientnet', 'effnet')}.py fp32 eager, "
                                          f"torch threads={cores}, os.cpu_count={os.cpu_count()}"}
    emit(line)


def main() -> None:
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
