#!/usr/bin/env python
"""One line per kernel from `ncu -i X.ncu-rep --page raw --csv`."""
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
want = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rdMB"),
        ("dram__bytes_write.sum", "wrMB"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "bankconf"),
        ("smsp__inst_executed.sum", "winst"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%")]
idx = [(n, hdr.index(k)) for k, n in want if k in hdr]
print(" | ".join(n for n, _ in idx))
for r in rows[2:]:
    print(" | ".join(r[i][:30].replace("void residual_kernel", "rk") for _, i in idx))
