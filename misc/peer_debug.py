import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytrace_b200 as R
from raytrace_b200 import _native as N, scenes
from oracle import orc
dims, kind, seed = scenes.WORLDS["c1"]
W, H = 160, 400
flags = N.FLAG_DENOISE | N.FLAG_FINALIZE
noise = orc.blue_noise()
cam = scenes.camera(dims)
blk = orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], 5)
with R.Renderer(W, H, dims, 2, flags) as full:
    full.upload_blue_noise(noise); full.generate_world(kind, seed); full.set_frame(blk); full.render_frame(); want = full.read_output().copy()
for root in (0, 1, 2):
    bands = [(0, 136), (136, 264), (264, 400)]
    rs = [R.Renderer(W, H, dims, 2, flags, rows=b) for b in bands]
    for r in rs: r.upload_blue_noise(noise); r.generate_world(kind, seed)
    blobs = [r.peer_export() for r in rs]
    for k, r in enumerate(rs): r.peer_attach(blobs, k, root=root)
    t0 = time.perf_counter()
    for r in rs: r.set_frame(blk); r.render_frame()
    t1 = time.perf_counter()
    got = rs[root].read_output().copy()
    t2 = time.perf_counter()
    for r in rs: r.sync()
    got2 = rs[root].read_output().copy()
    bad = np.nonzero((got != want).any(axis=(1, 2)))[0]; bad2 = np.nonzero((got2 != want).any(axis=(1, 2)))[0]
    errs = []
    for r in rs:
        try: r.frame_stats(); errs.append("ok")
        except Exception as e: errs.append(str(e)[:60])
    print(f"root={root}: enqueue {t1-t0:.3f}s read {t2-t1:.3f}s bad rows first read {bad.size} {bad[:3]}..{bad[-3:]} second read {bad2.size}; stats {errs}", flush=True)
    for r in rs: r.close()
