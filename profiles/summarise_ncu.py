#!/usr/bin/env python
"""Turn ncu output into the markdown tables kept under profiles/.

  launches  <launches.csv>              per-kernel totals and shares of an
                                        `ncu --metrics gpu__time_duration.sum --csv` launch list
  full      <raw.csv> [kernel-regex]    selected metrics of `ncu -i X.ncu-rep --page raw --csv`
"""
import csv
import re
import sys
from collections import OrderedDict

FULL_METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum.per_second",
    "dram__bytes_write.sum.per_second", "lts__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warp_latency_issue_stalled_barrier.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
]


def rows_of(path):
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    return list(csv.reader(lines))


def launches(path):
    """Per-kernel launches, time share, average duration and (when the pass collected them) DRAM
    bytes read / written per launch."""
    rows = rows_of(path)
    hdr = rows[0]
    ik, iv, im, iu = (hdr.index(k) for k in ("Kernel Name", "Metric Value", "Metric Name", "Metric Unit"))
    scale = {"ns": 1.0, "us": 1e3, "ms": 1e6, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    agg = OrderedDict()
    for r in rows[1:]:
        a = agg.setdefault(r[ik], {"n": 0, "t": 0.0, "rd": 0.0, "wr": 0.0})
        v = float(r[iv].replace(",", "")) * scale.get(r[iu], 1.0)
        if r[im] == "gpu__time_duration.sum":
            a["n"] += 1
            a["t"] += v
        elif r[im] == "dram__bytes_read.sum":
            a["rd"] += v
        elif r[im] == "dram__bytes_write.sum":
            a["wr"] += v
    tot = sum(a["t"] for a in agg.values())
    print("| kernel | launches | total (ns) | share | avg (ns) | DRAM read / launch (MB) | DRAM write / launch (MB) |")
    print("|---|---:|---:|---:|---:|---:|---:|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["t"]):
        n = a["n"]
        print(f"| `{k}` | {n} | {a['t']:.0f} | {100 * a['t'] / tot:.2f} % | {a['t'] / n:.1f} | "
              f"{a['rd'] / n / 1e6:.1f} | {a['wr'] / n / 1e6:.1f} |")
    print(f"| **all** | {sum(a['n'] for a in agg.values())} | {tot:.0f} | 100 % | | | |")


def full(path, pattern=".*"):
    rows = rows_of(path)
    hdr, units, data = rows[0], rows[1], rows[2:]
    ik = hdr.index("Kernel Name")
    data = [r for r in data if re.search(pattern, r[ik])]
    print("| metric | unit | " + " | ".join(f"launch {i + 1}" for i in range(len(data))) + " |")
    print("|---|---|" + "---:|" * len(data))
    for m in FULL_METRICS:
        if m in hdr:
            i = hdr.index(m)
            print(f"| {m} | {units[i]} | " + " | ".join(r[i] for r in data) + " |")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        full(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else ".*")
