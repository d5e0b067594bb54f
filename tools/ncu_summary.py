"""Turn `ncu -i X.ncu-rep --page raw --csv` output into the short per-kernel summary kept under profiles/.
usage: python tools/ncu_summary.py raw.csv [raw2.csv ...] > profiles/rNN_ncu_full_*.csv"""
import csv
import sys

COLS = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__waves_per_multiprocessor",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]

w = csv.writer(sys.stdout)
first = True
for path in sys.argv[1:]:
    rows = list(csv.reader(open(path)))
    h, units = rows[0], rows[1]
    idx = [h.index(c) if c in h else -1 for c in COLS]
    if first:
        w.writerow(COLS)
        w.writerow([units[i] if i >= 0 else "" for i in idx])
        first = False
    for r in rows[2:]:
        w.writerow([(r[i][:80] if i >= 0 else "") for i in idx])
