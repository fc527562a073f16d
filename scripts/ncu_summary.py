#!/usr/bin/env python
"""Condense `ncu -i X.ncu-rep --page raw --csv` into the few columns DESIGN.md and bench.py cite.
usage: ncu -i rep.ncu-rep --page raw --csv | python scripts/ncu_summary.py > profiles/<name>.tsv"""
import csv
import sys

COLS = ["launch__grid_size", "launch__block_size", "launch__registers_per_thread", "gpu__time_duration.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"]
rows = list(csv.reader(sys.stdin))
head, units = rows[0], rows[1]
idx = {n: i for i, n in enumerate(head)}
name_i = idx.get("Kernel Name")
print("\t".join(["kernel"] + [f"{c} [{units[idx[c]]}]" for c in COLS if c in idx]))
for r in rows[2:]:
    print("\t".join([r[name_i]] + [r[idx[c]] for c in COLS if c in idx]))
