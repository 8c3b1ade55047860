"""Selects the roofline-relevant columns of an `ncu --page raw --csv` export (one row per profiled launch).
    ncu -i rep.ncu-rep --page raw --csv | python tools/ncu_select.py > profiles/<name>.csv"""
import csv
import sys

COLS = ["ID", "Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__inst_executed.sum",
        "sm__cycles_elapsed.max"]
rows = list(csv.reader(sys.stdin))
head = rows[0]
idx = [head.index(c) for c in COLS if c in head]
w = csv.writer(sys.stdout)
for r in rows:
    if len(r) >= len(head):
        w.writerow([r[i] for i in idx])
