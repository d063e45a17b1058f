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
    m = re.match(r"([\w:]+)(<[^>]*>)?", name)
    return (m.group(1) + (m.group(2) or "")) if m else name[:40]


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
    hdr, data = rows[0], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    cols = [("gpu__time_duration.sum", "us", 8), ("dram__bytes_read.sum", "rdMB", 8), ("dram__bytes_write.sum", "wrMB", 8),
            ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%", 6),
            ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%", 6),
            ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%", 6),
            ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%", 6),
            ("sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active", "tens%", 6),
            ("launch__registers_per_thread", "regs", 5), ("launch__grid_size", "grid", 6), ("launch__block_size", "blk", 5)]
    print("# one forward step (N=256, bf16), ncu --set full --clock-control none; dram bytes are per launch")
    print(f"{'#':>3s} {'kernel':44s} " + " ".join(f"{c[1]:>{c[2]}s}" for c in cols))
    for n, r in enumerate(data):
        if len(r) < len(hdr):
            continue
        vals = []
        for key, _, w in cols:
            v = r[idx[key]] if key in idx else "-"
            try:
                v = f"{float(v):.1f}" if "." in v else v
            except ValueError:
                pass
            vals.append(f"{v:>{w}s}")
        print(f"{n:3d} {short(r[idx['Kernel Name']])[:44]:44s} " + " ".join(vals))


if __name__ == "__main__":
    {"launches": launches, "step": step}[sys.argv[1]](sys.argv[2])
