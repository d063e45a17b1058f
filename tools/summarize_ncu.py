"""Turn ncu CSV exports into the small per-kernel tables committed under profiles/ (dev tool, no GPU needed).

  python tools/summarize_ncu.py launches gpurun_out/X_launches.csv > profiles/X_launch_shares.txt
  python tools/summarize_ncu.py step     gpurun_out/X_step_raw.csv > profiles/X_step_metrics.txt
"""
import csv
import re
import sys
from collections import OrderedDict


def short(name: str) -> str:
    name = re.sub(r"^void\s+", "", name)
    name = name.replace("b200df::", "")
    name = re.sub(r"(\(anonymous namespace\)|<?unnamed>)::", "", name)   # kernels in an unnamed namespace (attn_mma.cu, cam_bwd.cu)
    m = re.match(r"([\w:]+)(<[^>]*>)?", name)
    return (m.group(1) + (m.group(2) or "")) if m else name[:40]


def stamp() -> None:
    """Which build the capture belongs to (bench.py refuses traffic numbers from a capture of other kernel sources)."""
    import subprocess
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    sys.path.insert(0, str(root))
    from bench import kernels_sha

    commit = subprocess.run(["git", "-C", str(root), "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    print(f"# kernels-sha: {kernels_sha()}")
    print(f"# summarised at commit {commit} (capture taken on the tree with that kernels-sha)")


def launches(path: str) -> None:
    rows = [r for r in csv.reader(open(path, errors="ignore")) if len(r) > 10]
    hdr = rows[0]
    i_name, i_val, i_metric = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name")
    agg: "OrderedDict[str, list]" = OrderedDict()
    for r in rows[1:]:
        if r[i_metric] != "gpu__time_duration.sum":
            continue
        k = short(r[i_name])
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += float(r[i_val].replace(",", ""))
    tot = sum(v[1] for v in agg.values())
    print(f"# per-kernel share of GPU time over the captured launches (ncu gpu__time_duration.sum, cold-cache, serialised)")
    print(f"{'kernel':58s} {'launches':>8s} {'total_us':>12s} {'share':>7s} {'avg_us':>9s}")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k[:58]:58s} {n:8d} {t / 1e3:12.1f} {100 * t / tot:6.1f}% {t / n / 1e3:9.2f}")
    print(f"{'TOTAL':58s} {sum(v[0] for v in agg.values()):8d} {tot / 1e3:12.1f}")


def step(path: str) -> None:
    rows = list(csv.reader(open(path, errors="ignore")))
    h0 = next(i for i, r in enumerate(rows) if r and r[0] == "ID")   # ncu's own messages may precede the table
    hdr, units, data = rows[h0], rows[h0 + 1], rows[h0 + 2:]
    hdr = [h.split(".TriageCompute.")[-1] if ".TriageCompute." in h else h for h in hdr]
    idx = {h: i for i, h in enumerate(hdr)}
    stamp()
    cols = [("gpu__time_duration.sum", "us", 8), ("dram__bytes_read.sum", "rdMB", 8), ("dram__bytes_write.sum", "wrMB", 8),
            ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%", 6),
            ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%", 6),
            ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%", 6),
            ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%", 6),
            ("sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active", "tens%", 6),
            ("launch__registers_per_thread", "regs", 5), ("launch__grid_size", "grid", 6), ("launch__block_size", "blk", 5)]
    # the tensor-pipe column: whichever sm__pipe_tensor* / sm__inst_executed_pipe_tensor* percentage this ncu version exports
    tens = [h for h in hdr if "pipe_tensor" in h and "pct_of_peak" in h]
    tens = [h for h in tens if "cycles_active" in h] or tens
    # tcgen05.mma runs on the `tc` pipe: sm__pipe_tensor* (the mma.sync / HMMA pipe) reads 1-15 % on kernels whose
    # sm__pipe_tc_cycles_active is 40-60 % -- prefer the tc pipe when this ncu version exports it
    tc = [h for h in hdr if h.startswith("sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_elapsed")]
    if tc:
        tens = tc
    if tens:
        cols = [(tens[0], c[1], c[2]) if c[1] == "tens%" else c for c in cols]
    print("# one forward step (N=256, bf16), ncu --set full --clock-control none; dram bytes are per launch"
          + (f"; tens% = {tens[0]}" if tens else ""))
    print(f"{'#':>3s} {'kernel':44s} " + " ".join(f"{c[1]:>{c[2]}s}" for c in cols))
    for n, r in enumerate(data):
        if len(r) < len(hdr):
            continue
        vals = []
        for key, label, w in cols:
            v = r[idx[key]] if key in idx else "-"
            try:
                if label in ("us", "rdMB", "wrMB"):   # normalise whatever unit this ncu version chose
                    u = units[idx[key]].lower()
                    scale = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "byte": 1e-6,
                             "kbyte": 1e-3, "mbyte": 1.0, "gbyte": 1e3}.get(u, 1.0)
                    v = f"{float(v.replace(',', '')) * scale:.1f}"
                else:
                    v = f"{float(v):.1f}" if "." in v else v
            except ValueError:
                pass
            vals.append(f"{v:>{w}s}")
        print(f"{n:3d} {short(r[idx['Kernel Name']])[:44]:44s} " + " ".join(vals))


def details(path: str) -> None:
    """Long-form CSV (ncu --set full --csv without --page raw: one metric per row) -> one line per launch."""
    rows = [r for r in csv.reader(open(path, errors="ignore")) if len(r) > 14]
    hdr = rows[0]
    idx = {h: i for i, h in enumerate(hdr)}
    want = [("Duration", "us", 8), ("DRAM Throughput", "dram%", 6), ("Compute (SM) Throughput", "sm%", 6),
            ("Issue Slots Busy", "issue%", 6), ("Achieved Occupancy", "occ%", 6), ("Registers Per Thread", "regs", 5),
            ("Grid Size", "grid", 6), ("Block Size", "blk", 5)]
    per: dict[int, dict] = {}
    for r in rows[1:]:
        try:
            k = int(r[idx["ID"]])
        except ValueError:
            continue
        d = per.setdefault(k, {"name": r[idx["Kernel Name"]]})
        nm = r[idx["Metric Name"]]
        if nm not in d:
            val, unit = r[idx["Metric Value"]].replace(",", ""), r[idx["Metric Unit"]]
            if nm == "Duration":
                try:
                    val = f"{float(val) * {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 'msecond': 1e3, 'usecond': 1.0, 'nsecond': 1e-3}.get(unit, 1.0):.1f}"
                except ValueError:
                    pass
            d[nm] = val
    stamp()
    print("# one forward step (N=256, bf16), ncu --set full --clock-control none (details page, one line per launch)")
    print(f"{'#':>3s} {'kernel':44s} " + " ".join(f"{c[1]:>{c[2]}s}" for c in want))
    tot = 0.0
    for k in sorted(per):
        d = per[k]
        vals = []
        for key, _, w in want:
            v = d.get(key, "-")
            try:
                v = f"{float(v):.1f}" if "." in v else v
            except ValueError:
                pass
            vals.append(f"{v:>{w}s}")
        try:
            tot += float(d.get("Duration", 0))
        except ValueError:
            pass
        print(f"{k:3d} {short(d['name'])[:44]:44s} " + " ".join(vals))
    print(f"# sum of durations: {tot / 1e3:.3f} ms over {len(per)} launches")


if __name__ == "__main__":
    {"launches": launches, "step": step, "details": details}[sys.argv[1]](sys.argv[2])
