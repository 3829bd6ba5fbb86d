"""Print selected raw metrics of an .ncu-rep (first profiled launch).  usage: ncu_keys.py report.ncu-rep [substr ...]"""
import csv
import subprocess
import sys

keys = sys.argv[2:] or ["gpu__time_duration.sum", "dram__bytes_read.sum ", "dram__bytes_write.sum ", "Grid Size", "launch__grid_size",
                        "sm__warps_active.avg.pct", "launch__registers_per_thread ", "launch__occupancy_limit",
                        "sm__inst_executed_pipe_fp64.avg.pct", "smsp__issue_active.avg.pct", "smsp__average_warps_issue_stalled",
                        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum ", "lts__t_sector_hit_rate",
                        "sm__inst_executed_pipe_lsu", "sm__throughput.avg.pct", "l1tex__throughput.avg.pct", "lts__throughput.avg.pct"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u, v = rows[0], rows[1], rows[2]
for a, b, c in zip(h, u, v):
    if any(k.strip() in a for k in keys):
        print("%-90s %-12s %s" % (a, b, c))
