#!/usr/bin/env python
"""misc/r2/sass_loop.py [LIB] [INSTANCE-SUBSTRING] — static look at k_trace_fast's SASS: registers, total instructions and the
length of every backward-branch loop (the march loop is the one with FMNMX3 + LDG.E.U8)."""
import re, subprocess, sys, tempfile, os
lib = sys.argv[1] if len(sys.argv) > 1 else "raytrace_b200/libvxrt.so"
want = sys.argv[2] if len(sys.argv) > 2 else "ILi2ELb1ELb0E"
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", sys.argv[3] if len(sys.argv) > 3 and not sys.argv[3].startswith("-") else "trace_fast", os.path.abspath(lib)], cwd=d, check=True, stdout=subprocess.DEVNULL)
cub = max([f for f in os.listdir(d) if f.endswith(".cubin")], key=lambda f: os.path.getsize(os.path.join(d, f)))
sass = subprocess.run(["nvdisasm", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout.split("\n")
res = subprocess.run(["cuobjdump", "-res-usage", os.path.join(d, cub)], capture_output=True, text=True).stdout
for m in re.finditer(r"Function (\S+):\n\s*(REG:\d+).*?(SHARED:\d+)", res):
    if want in m.group(1): print(m.group(1)[-40:], m.group(2), m.group(3))
start = [i for i, l in enumerate(sass) if l.startswith(".text.") and want in l][0]
end = next(i for i in range(start + 1, len(sass)) if sass[i].startswith("//-----"))
seg = sass[start:end]
labels, ins = {}, []
for l in seg:
    m = re.match(r"(\.L_x_\d+):", l)
    if m: labels[m.group(1)] = len(ins)
    m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
    if m: ins.append(m.group(2).strip())
print("instructions:", len(ins))
for i, t in enumerate(ins):
    m = re.search(r"BRA.*`\((\.L_x_\d+)\)", t)
    if m and m.group(1) in labels and labels[m.group(1)] <= i:
        body = ins[labels[m.group(1)]:i + 1]
        tag = "MARCH" if any("FMNMX3" in b for b in body) and any("LDG.E.U8" in b for b in body) else ""
        print(f"loop of {len(body)} instructions {tag}")
        if tag and "-v" in sys.argv:
            print("\n".join("    " + b for b in body))
