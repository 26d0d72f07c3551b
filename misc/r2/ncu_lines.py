#!/usr/bin/env python
"""misc/r2/ncu_lines.py REPORT KERNEL-REGEX [N] — executed warp instructions and stall samples per CUDA source line of one captured
launch (ncu --import-source on), from `ncu --page source --print-source cuda,sass --csv`; the N heaviest lines."""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 60
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout
cur, hdr, out, seen_fn = None, None, [], None
for r in csv.reader(io.StringIO(raw)):
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        if seen_fn is None: seen_fn = r[1]
        elif r[1] != seen_fn: cur = None        # a second captured launch: stop
        continue
    if r[0] == "Line No": hdr = r; continue
    if cur and hdr and r[0] not in ("", "-"):
        try: n = int(r[hdr.index("Instructions Executed")]); smp = int(r[hdr.index("# Samples")])
        except ValueError: continue
        thr = int(r[hdr.index("Thread Instructions Executed")])
        out.append((n, smp, thr, cur, r[0], r[1].strip()[:120]))
tot = sum(o[0] for o in out); ts = sum(o[1] for o in out)
print(seen_fn); print("warp instructions", tot, "samples", ts)
for n, smp, thr, f, l, s in sorted(out, reverse=True)[:top]:
    print(f"{n:10d} {100*n/tot:5.1f}%  smp {100*smp/max(ts,1):5.1f}%  lanes {thr/max(n,1):5.1f}  {f}:{l}  {s}")
