#!/usr/bin/env python
"""Per-kernel summary of an `ncu --page raw --csv` export: duration, DRAM bytes, DRAM %, issue / pipe utilisation,
registers, occupancy (last launch of each kernel name).  python tools/ncu_summary.py raw.csv > summary.txt"""
import csv
import sys

KEYS = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram %"), ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "issue %"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu %"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma %"),
        ("launch__registers_per_thread", "regs"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
        ("launch__grid_size", "grid"), ("launch__block_size", "block"), ("smsp__inst_executed.sum", "warp inst")]
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
name_i = hdr.index("Kernel Name")
last = {}
for r in data:
    last[r[name_i]] = r
for name, r in last.items():
    print(name[:150])
    for k, label in KEYS:
        if k in hdr:
            i = hdr.index(k)
            print("    %-16s %s %s" % (label, r[i], units[i]))
