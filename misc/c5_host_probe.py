"""Where the host waits in the C5 end-to-end loop: wall time of each call of the per-frame sequence."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytrace_b200 as R
from raytrace_b200 import _native as N, scenes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from oracle import orc          # scene inputs only (edit lists, frame blocks), as bench.py does

dims, kind, seed = scenes.WORLDS["c2"]
W, H = 1920, 1080
r = R.Renderer(W, H, dims, N.MODE_FULL, N.FLAG_DENOISE | N.FLAG_FINALIZE)
r.upload_blue_noise(orc.blue_noise()); r.generate_world(kind, seed)
blocks = bench.frame_blocks(orc, dims, 64, 40)
edits = [np.ascontiguousarray(orc.generate_edits(10000, *dims, 7, i), np.int32) for i in range(40)]
out = [np.empty((H, W, 4), np.uint8) for _ in range(2)]
for b in out: r.host_register(b)
for i in range(4):
    r.apply_edits(edits[i]); r.set_frame(blocks[i]); r.render_frame()
r.sync()
names = ["apply_edits", "set_frame", "render_frame", "read_async", "wait_prev"]
acc = dict.fromkeys(names, 0.0)
prev = None
t0 = time.perf_counter()
for k in range(4, 36):
    a = time.perf_counter(); r.apply_edits(edits[k])
    b = time.perf_counter(); r.set_frame(blocks[k])
    c = time.perf_counter(); r.render_frame()
    d = time.perf_counter(); t = r.read_output_async(out[k & 1])
    e = time.perf_counter()
    if prev is not None: r.wait_output(prev)
    f = time.perf_counter(); prev = t
    for n, v in zip(names, (b - a, c - b, d - c, e - d, f - e)): acc[n] += v
r.wait_output(prev); r.sync()
tot = time.perf_counter() - t0
print("per frame: total %.3f ms; " % (tot / 32 * 1e3) + ", ".join("%s %.3f" % (n, acc[n] / 32 * 1e3) for n in names))
