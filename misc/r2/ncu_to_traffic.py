#!/usr/bin/env python
"""misc/r2/ncu_to_traffic.py — profiles/ncu_traffic.json (what bench.py's roofline.traffic / roofline.ncu quote) from the committed
full-set summaries of the trace pass, one entry per workload."""
import csv, json
SRC = [("c2", "profiles/r2_w_c2_frame_ncu_full_summary.csv"), ("c3", "profiles/r2_w_c3_trace_ncu_full_summary.csv"),
       ("c4", "profiles/r2_w_c4_trace_ncu_full_summary.csv")]
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "msecond": 1e3, "usecond": 1, "nsecond": 1e-3, "s": 1e6}
out = {}
for w, p in SRC:
    rows = list(csv.reader(open(p)))
    hdr, units = rows[1], rows[2]
    r = next(r for r in rows[3:] if "k_trace" in r[0])
    d, u = dict(zip(hdr, r)), dict(zip(hdr, units))
    v = lambda k: float(d[k]) * SCALE.get(u[k], 1)
    rd, wr = v("dram__bytes_read.sum"), v("dram__bytes_write.sum")
    out[w] = {"kernel": d["Kernel Name"] + ")", "dram_bytes_per_launch": rd + wr, "dram_read_bytes": rd, "dram_write_bytes": wr,
              "issue_active_pct": v("smsp__issue_active.avg.pct_of_peak_sustained_active"),
              "lanes_per_instruction": v("smsp__thread_inst_executed_per_inst_executed.ratio"),
              "warp_instructions": v("smsp__inst_executed.sum"), "l2_hit_pct": v("lts__t_sector_hit_rate.pct"),
              "launch_us_under_ncu": v("gpu__time_duration.sum"),
              "source": p + " (ncu --set full --clock-control none, one launch, L2 flushed before the frame; misc/r2/run_w.sh)"}
json.dump(out, open("profiles/ncu_traffic.json", "w"), indent=1)
print({w: (round(o["dram_bytes_per_launch"] / 1e6, 1), o["launch_us_under_ncu"]) for w, o in out.items()})
