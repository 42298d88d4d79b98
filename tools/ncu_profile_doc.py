#!/usr/bin/env python
"""profiles/r2_ncu_full_n1m.md and profiles/r2_ncu_summary.json from the two --set full captures of tools/final_run.sh
(gpurun_out/r2_tiles_n1m.ncu-rep, gpurun_out/r2_peel_n1m.ncu-rep).  bench.py reads the JSON for roofline.traffic and
issue_roofline."""
import csv, json, subprocess, sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
           "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_active",
           "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
           "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_shared_mem",
           "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
           "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"]
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}


def load(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units, vals = rows[0], rows[1], rows[2]
    return {k: (vals[i], units[i]) for i, k in enumerate(h)}


def num(d, k):
    v, u = d[k]
    return float(v.replace(",", "")) * SCALE.get(u, 1.0)


def main(n=1_000_000):
    t, p = load("gpurun_out/r2_tiles_n1m.ncu-rep"), load("gpurun_out/r2_peel_n1m.ncu-rep")
    alg = n * n / 16
    lines = ["# Round 2: ncu --set full --clock-control none, N = 1 000 000, M = 3, one B200 (kernels at HEAD)", "",
             "Command: `python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras --no-nsga2` (plain run first, then under",
             "ncu: `tools/final_run.sh`; one launch of each kernel captured after three warm-up launches).  Per-launch values.", "",
             f"| metric | unit | `{t['Kernel Name'][0].split('(')[0][:28]}` | `{p['Kernel Name'][0].split('(')[0][:28]}` |", "|---|---|---|---|"]
    for m in METRICS:
        if m in t:
            pv = p[m][0] if p[m][1] == t[m][1] else f"{p[m][0]} {p[m][1]}"
            lines.append(f"| {m} | {t[m][1]} | {t[m][0]} | {pv} |")
    tb = num(t, "dram__bytes_read.sum") + num(t, "dram__bytes_write.sum")
    pb = num(p, "dram__bytes_read.sum") + num(p, "dram__bytes_write.sum")
    ti = num(t, "smsp__inst_executed.sum")
    per = ti / (n * n / 2 / (64 * 32))
    issue = num(t, "smsp__issue_active.avg.pct_of_peak_sustained_active")
    lines += ["", f"Tile kernel: DRAM traffic {tb / 1e9:.2f} GB per launch against N^2/16 = {alg / 1e9:.2f} GB algorithmic (x{tb / alg:.3f}); "
              f"{ti / 1e9:.2f} G warp instructions = {per:.1f} per 64 x 32 tested pairs (round 1: 69); issue slots {issue:.1f} % busy.",
              f"Peel kernel: DRAM traffic {pb / 1e9:.2f} GB per launch against {alg / 1e9:.2f} GB (x{pb / alg:.3f}): every stored byte is read once.", ""]
    open("profiles/r2_ncu_full_n1m.md", "w").write("\n".join(lines))
    json.dump({"source": "profiles/r2_ncu_full_n1m.md (ncu --set full --clock-control none, N=1M, M=3, kernels at HEAD; "
                         "dram__bytes_read.sum + dram__bytes_write.sum per launch)",
               "tiles_dram_bytes_per_launch": tb, "peel_dram_bytes_per_launch": pb,
               "tiles_issue": {"bound": "issue-slots", "issue_active_pct": issue, "warp_instructions_per_launch": ti,
                               "warp_instructions_per_64x32_tested_pairs": per,
                               "round1_warp_instructions_per_64x32_tested_pairs": 69.0}},
              open("profiles/r2_ncu_summary.json", "w"), indent=1)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
