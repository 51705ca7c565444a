#!/usr/bin/env python
"""Turn ncu output (read on the CPU box) into the small CSV summaries committed under profiles/.

  python profiles/summarize_ncu.py launches gpurun_out/X_launches.csv  > profiles/X_launches_summary.csv
  python profiles/summarize_ncu.py full     gpurun_out/X_full_raw.csv  > profiles/X_full_summary.csv

`launches` input: `ncu --metrics gpu__time_duration.sum --clock-control none --csv` log.
`full` input:     `ncu -i X.ncu-rep --page raw --csv` of an `ncu --set full` capture.
"""
import csv
import re
import sys
from collections import OrderedDict


def short(name: str) -> str:
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("xg::<unnamed>::", "").replace("<unnamed>::", "").replace("void ", "").replace("xg::", "")
    return name.strip()


def launches(path: str, skip_prefixes=("synth_", "part_", "at::", "at_cuda_detail", "ncclDev", "native::", "cuda::kernelHistogram", "cub::")):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    k, v = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = OrderedDict()
    for r in rows[1:]:
        n = short(r[k])
        if n.startswith(skip_prefixes) or "elementwise" in n or "at::native" in r[k]:
            continue
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += float(r[v].replace(",", "")) / 1e3
    tot = sum(a[1] for a in agg.values())
    w = csv.writer(sys.stdout)
    w.writerow(["kernel", "launches", "total_us", "share", "avg_us"])
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        w.writerow([n, c, f"{t:.1f}", f"{t / tot:.3f}", f"{t / c:.1f}"])


FULL = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"]


def full(path: str):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    cols = [(m, hdr.index(m)) for m in FULL if m in hdr]
    k = hdr.index("Kernel Name")
    w = csv.writer(sys.stdout)
    w.writerow(["kernel"] + [f"{m}[{units[i]}]" for m, i in cols])
    for r in rows[2:]:
        w.writerow([short(r[k])] + [r[i] for _, i in cols])


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
