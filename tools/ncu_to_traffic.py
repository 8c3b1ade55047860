"""Turns the selected ncu CSVs of one hop launch per shape (tools/ncu_select.py output) into
profiles/roofline_traffic.json, the per-launch hardware counters bench.py's `roofline` block reads.

    python tools/ncu_to_traffic.py <sha256 of the profiled .so> reddit=profiles/r03_ncu_full_spmm_hop_reddit.csv products=... > profiles/roofline_traffic.json

Every figure is per launch of the hop kernel as captured by `ncu --set full --clock-control none`; `build_sha256` is
the sha256 of the libsgc_b200.so that was profiled (a record: nvcc output is not byte-reproducible), `source_sha256`
that of the kernel sources it was built from (bench.kernel_source_sha256), which bench.py compares with the
sources of the library it times."""
import csv
import datetime
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import kernel_source_sha256  # noqa: E402

UNIT_SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ms": 1, "us": 1e-3, "s": 1e3,
              "msecond": 1, "usecond": 1e-3, "second": 1e3}


def load(path):
    rows = list(csv.reader(open(path)))
    head, units, vals = rows[0], rows[1], rows[-1]
    out = {}
    for h, u, v in zip(head, units, vals):
        try:
            out[h] = float(v.replace(",", "")) * UNIT_SCALE.get(u, 1)
        except ValueError:
            out[h] = v
    return out


def entry(path):
    c = load(path)
    pct = lambda k: round(c[k], 2)
    dur_ms = c["gpu__time_duration.sum"]
    shared_wf = c["l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]
    # the data pipe retires one wavefront per cycle per SM: the peak over the launch follows from any of its
    # sub-counters and that counter's own percentage
    pipe_peak_wf = shared_wf / (c["l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"] / 100.0)
    pipes = {
        "l1tex_lsu_data_pipe": pct("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
        "l1tex_lsu_writeback": pct("l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed"),
        "lts": pct("lts__throughput.avg.pct_of_peak_sustained_elapsed"),
        "xbar_to_l1tex_read": pct("l1tex__m_xbar2l1tex_read_bytes.sum.pct_of_peak_sustained_elapsed"),
        "hbm": pct("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        "sm_issue": pct("sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    }
    return {
        "dram_bytes": int(c["dram__bytes_read.sum"] + c["dram__bytes_write.sum"]),
        "dram_read_bytes": int(c["dram__bytes_read.sum"]), "dram_write_bytes": int(c["dram__bytes_write.sum"]),
        "l2_to_sm_bytes": int(c["l1tex__m_xbar2l1tex_read_bytes.sum"]),
        "l2_hit_rate_pct": pct("lts__t_sector_hit_rate.pct"), "l1_hit_rate_pct": pct("l1tex__t_sector_hit_rate.pct"),
        "duration_ms_under_ncu": round(dur_ms, 4),
        "registers": int(c["launch__registers_per_thread"]), "ctas_per_sm": int(c["launch__occupancy_limit_registers"]),
        "kernel": c["Kernel Name"].replace("void ", "").split("(")[0],
        "pipes_pct_of_peak_sustained_elapsed": pipes,
        "busiest_unit": max(pipes, key=pipes.get),
        "lts_busiest_slice_pct": pct("lts__throughput.max.pct_of_peak_sustained_elapsed"),
        "l1tex_busiest_sm_pct": pct("l1tex__data_pipe_lsu_wavefronts.max.pct_of_peak_sustained_elapsed"),
        "lsu_data_pipe_wavefronts": int(pipe_peak_wf * pipes["l1tex_lsu_data_pipe"] / 100.0),
        "lsu_data_pipe_peak_wavefronts_in_launch": int(pipe_peak_wf),
        "shfl_wavefronts": int(shared_wf),      # mem_shared wavefronts of a kernel without shared memory = SHFL broadcasts
        "global_load_wavefronts": int(c["l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum"]),
        "capture": path, "date": datetime.date.today().isoformat(),
    }


sha = sys.argv[1]
out = {"_note": "per-launch counters of the hop kernel from `ncu --set full --clock-control none` captures "
                "(tools/gpu_r3_evidence.sh, 1x B200); bench.py reads dram_bytes into roofline.traffic / roofline.dram, "
                "the pipe percentages into roofline.hw_counters / roofline.bound, and compares source_sha256 with the "
                "kernel sources of the library it times",
       "build_sha256": sha, "source_sha256": kernel_source_sha256()}
for arg in sys.argv[2:]:
    name, path = arg.split("=", 1)
    out[name] = entry(path)
json.dump(out, sys.stdout, indent=1)
print()
