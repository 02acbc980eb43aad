#!/usr/bin/env python
"""Prints the handful of ncu raw-page metrics the profiles/ summaries quote:
    python tools/ncu_summary.py rep_or_raw_csv [...]      (.ncu-rep files, or `ncu -i rep --page raw --csv` exports)"""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "l1tex__t_sector_hit_rate.pct",
        "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second", "launch__registers_per_thread",
        "launch__grid_size", "l1tex__m_xbar2l1tex_read_bytes.sum", "lts__t_sectors_srcunit_tex_op_read.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "launch__block_size",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__warps_eligible.avg.per_cycle_active"]
for rep in sys.argv[1:]:
    if rep.endswith(".csv"):
        out = open(rep).read()
    else:
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    print("#", rep, "|", vals[hdr.index("Kernel Name")][:90] if "Kernel Name" in hdr else "")
    for h, u, v in zip(hdr, units, vals):
        if h in WANT:
            print(f"  {h} = {v} {u}")
    stalls = sorted(((float(v), h.split("issue_stalled_")[1].split("_per_issue")[0]) for h, v in zip(hdr, vals)
                     if "issue_stalled" in h and h.endswith("per_issue_active.ratio") and v not in ("", "n/a")), reverse=True)
    print("  stalls per issue: " + ", ".join(f"{n} {x:.2f}" for x, n in stalls if x >= 0.05))
