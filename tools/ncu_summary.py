#!/usr/bin/env python
"""Condense `ncu -i X.ncu-rep --page raw --csv` into the handful of columns DESIGN.md / profiles/ quote.
usage: ncu -i gpurun_out/prof.ncu-rep --page raw --csv > raw.csv; python tools/ncu_summary.py raw.csv
"""
import csv
import sys

KEYS = [
    ("time_us", "gpu__time_duration.sum"),
    ("dram_rd_MB", "dram__bytes_read.sum"),
    ("dram_wr_MB", "dram__bytes_write.sum"),
    ("regs", "launch__registers_per_thread"),
    ("occ_lim_regs", "launch__occupancy_limit_registers"),
    ("occ_lim_smem", "launch__occupancy_limit_shared_mem"),
    ("warps_active_%", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("issue_active_%", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    ("inst_executed", "smsp__inst_executed.sum"),
    ("pipe_fma_%", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
    ("pipe_fmaheavy_%", "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active"),
    ("pipe_alu_%", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
    ("pipe_fp64_%", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
    ("pipe_lsu_%", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
    ("sm_throughput_%", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("dram_throughput_%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("l2_hit_%", "lts__t_sector_hit_rate.pct"),
    ("smem_bank_conflicts", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"),
]
STALL = "smsp__average_warps_issue_stalled_"


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, body = rows[0], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    names = [r[col["Kernel Name"]].split("(")[0].replace("void ", "")[:44] for r in body]
    print("%-22s" % "metric", *["%-28s" % n[:28] for n in names])
    for label, key in KEYS:
        if key in col:
            print("%-22s" % label, *["%-28s" % r[col[key]][:14] for r in body])
    stalls = [h for h in hdr if h.startswith(STALL) and h.endswith("_per_issue_active.ratio")]
    table = []
    for h in stalls:
        vals = [float(r[col[h]] or 0) for r in body]
        table.append((max(vals), h[len(STALL):-len("_per_issue_active.ratio")], vals))
    for _, name, vals in sorted(table, reverse=True)[:8]:
        print("%-22s" % ("stall:" + name)[:22], *["%-28.2f" % v for v in vals])


if __name__ == "__main__":
    main(sys.argv[1])
