#!/usr/bin/env python
"""misc/ncu_summary.py REPORT.ncu-rep OUT.csv "comment" — the per-launch metrics DESIGN.md quotes, from a --set full capture.
Also rewrites profiles/ncu_traffic.json (dram bytes of the first k_trace_fast launch) when --traffic is given."""
import csv, io, json, subprocess, sys
METRICS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
           "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed",
           "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum",
           "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_lsu.sum", "smsp__cycles_active.avg", "sm__cycles_elapsed.avg"]
rep, out, comment = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
cols = [hdr.index("Kernel Name")] + [hdr.index(m) for m in METRICS if m in hdr]
with open(out, "w", newline="") as f:
    f.write('"# %s"\n' % comment.replace('"', "'"))
    w = csv.writer(f)
    w.writerow([hdr[c] for c in cols]); w.writerow([units[c] for c in cols])
    for r in rows[2:]:
        w.writerow([r[c][:60] for c in cols])
if "--traffic" in sys.argv:
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        if "k_trace_fast" in d["Kernel Name"]:
            def b(k):
                v, u = float(d[k]), units[hdr.index(k)]
                return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
            rd, wr = b("dram__bytes_read.sum"), b("dram__bytes_write.sum")
            json.dump({"k_trace_fast": {"dram_bytes_per_launch": rd + wr, "dram_read_bytes": rd, "dram_write_bytes": wr,
                                        "source": "%s (ncu --set full, C2 workload, one launch, L2 flushed before the frame)" % out}},
                      open("profiles/ncu_traffic.json", "w"), indent=1)
            break
