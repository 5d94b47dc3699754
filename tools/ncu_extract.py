"""Extracts the columns used in profiles/*.csv from `ncu -i report.ncu-rep --page raw --csv` output.
usage: python tools/ncu_extract.py raw.csv out.csv [--append]"""
import csv
import sys

COLS = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
units = rows[1]
idx = [hdr.index(c) if c in hdr else -1 for c in COLS]
mode = "a" if "--append" in sys.argv else "w"
with open(sys.argv[2], mode, newline="") as f:
    w = csv.writer(f)
    if mode == "w":
        w.writerow(COLS)
        w.writerow([units[i] if i >= 0 else "" for i in idx])
    for r in rows[2:]:
        w.writerow([r[i] if i >= 0 else "" for i in idx])
