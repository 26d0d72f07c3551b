#!/usr/bin/env python
"""misc/r2/ncu_stalls.py REPORT [kernel-regex] — headline metrics + warp-state (stall) breakdown of each captured launch."""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__cycles_elapsed.avg", "smsp__cycles_active.avg",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "lts__t_sectors_srcunit_tex_op_read.sum"]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print("==", d["Kernel Name"][:70], d.get("Grid Size"))
    for k in want:
        if k in d: print(f"   {k:70s} {d[k]:>16s} {units[hdr.index(k)]}")
    st = [(k, float(d[k])) for k in hdr if k.startswith("smsp__average_warp") and k.endswith("_per_issue_active.ratio") or (k.startswith("smsp__average_warps_issue_stalled") and k.endswith(".ratio")) if d.get(k, "") not in ("", "n/a")]
    for k, v in sorted(st, key=lambda kv: -kv[1])[:9]:
        print(f"      {k.replace('smsp__average_warps_issue_stalled_', '').replace('smsp__average_warp_latency_issue_stalled_', ''):60s} {v:8.3f}")
