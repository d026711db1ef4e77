#!/usr/bin/env python
"""Summarise an ncu report (first kernel in it) into the counters DESIGN.md / bench.py quote: duration, DRAM traffic, issue
utilisation, warp execution efficiency, L1/L2 hit rates, occupancy, stall reasons.  Prints text; --json also writes a dict."""
import csv
import json
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.per_cycle_active", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__shared_mem_per_block", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__grid_size",
        "launch__block_size", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld_lookup_miss.sum", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tc.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
        "smsp__sass_average_branch_targets_threads_uniform.pct"]


def main():
    rep = sys.argv[1]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    h, u, v = rows[0], rows[1], rows[2]
    d = {}
    for a, b, c in zip(h, u, v):
        if a in ("Kernel Name",):
            d["kernel"] = c
        if a in KEYS or (a.startswith("smsp__average_warps_issue_stalled") and a.endswith("per_issue_active.ratio")):
            try:
                d[a] = (float(c.replace(",", "")), b)
            except ValueError:
                d[a] = (c, b)
    print("report:", rep)
    print("kernel:", d.pop("kernel", "?"))
    for k, (val, unit) in d.items():
        print("  %-86s %s %s" % (k, ("%.6g" % val) if isinstance(val, float) else val, unit))
    if "--json" in sys.argv:
        out = sys.argv[sys.argv.index("--json") + 1]
        mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        rd = d["dram__bytes_read.sum"][0] * mult[d["dram__bytes_read.sum"][1]]
        wr = d["dram__bytes_write.sum"][0] * mult[d["dram__bytes_write.sum"][1]]
        json.dump({"report": rep, "dram_bytes_read": rd, "dram_bytes_write": wr, "traffic_bytes_per_launch": rd + wr,
                   "metrics": {k: v[0] for k, v in d.items()}}, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
