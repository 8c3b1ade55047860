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
# pipe-utilisation counters that decide which unit binds a gather kernel: every raw-page column
# whose name starts with one of these is kept, whatever sub-metric suffixes this ncu version exports
PREFIXES = ["l1tex__throughput", "l1tex__lsu_writeback_active", "l1tex__data_pipe_lsu_wavefronts",
            "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld", "lts__throughput", "lts__t_sectors_srcunit_tex_op_read.sum",
            "l1tex__m_xbar2l1tex_read_bytes", "l1tex__m_l1tex2xbar_throughput", "dram__throughput",
            "smsp__inst_executed_op_shfl", "sm__inst_executed_pipe_lsu", "smsp__inst_executed_pipe_lsu"]
rows = list(csv.reader(sys.stdin))
head = rows[0]
idx = [head.index(c) for c in COLS if c in head]
idx += [i for i, c in enumerate(head) if i not in idx and any(c.startswith(p) for p in PREFIXES)]
w = csv.writer(sys.stdout)
for r in rows:
    if len(r) >= len(head):
        w.writerow([r[i] for i in idx])
