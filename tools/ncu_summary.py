#!/usr/bin/env python
"""Summarise ncu output brought back in gpurun_out/ into markdown for profiles/.

    python tools/ncu_summary.py launches <launches.csv>            # per-kernel totals and shares
    python tools/ncu_summary.py raw <file.ncu-rep> [metric ...]    # selected raw metrics per launch
"""
import collections
import csv
import subprocess
import sys

DEFAULT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
           "launch__block_size"]


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr = rows[start]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    tot, cnt = collections.defaultdict(float), collections.Counter()
    for r in rows[start + 1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        name = r[ki].split("(")[0][:70]
        tot[name] += v
        cnt[name] += 1
    T = sum(tot.values())
    print("| kernel | launches | total us | share |\n|---|---|---|---|")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:14]:
        print(f"| `{k}` | {cnt[k]} | {v / 1e3:.1f} | {100 * v / T:.1f}% |")
    print(f"\nTotal {T / 1e3:.0f} us over {sum(cnt.values())} launches.")


def raw(path, metrics):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h = rows[0]
    n = len(rows) - 2
    ki = h.index("Kernel Name")
    print("| metric | unit | " + " | ".join(rows[2 + j][ki].split("(")[0][:28] for j in range(n)) + " |")
    print("|---|---|" + "---|" * n)
    for w in metrics or DEFAULT:
        if w in h:
            i = h.index(w)
            print(f"| {w} | {rows[1][i]} | " + " | ".join(rows[2 + j][i] for j in range(n)) + " |")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        raw(sys.argv[2], sys.argv[3:])
