"""Probe: frames of a camera path rendered on S handles (S CUDA streams) of ONE GPU, submitted round-robin.
Prints frames/s for S = 1, 2, 3 (no L2 flush: the 640 MiB world is larger than L2)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import raytrace_b200 as R
from raytrace_b200 import _native as N, scenes
from oracle import orc
import bench

wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
scene, W, H, mode, n_path, accumulate, desc = bench.WORKLOADS[wl]
dims, kind, seed = scenes.WORLDS[scene]
noise = R.world.default_blue_noise()
K = 128
blocks = bench.frame_blocks(orc, dims, n_path, K)
for S in (1, 2, 3, 4):
    rs = []
    for s in range(S):
        r = R.Renderer(W, H, dims, mode, N.FLAG_DENOISE | N.FLAG_FINALIZE)
        r.upload_blue_noise(noise); r.generate_world(kind, seed); rs.append(r)
    for i in range(8):
        rs[i % S].set_frame(blocks[i]); rs[i % S].render_frame()
    for r in rs: r.sync()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(K):
        r = rs[i % S]; r.set_frame(blocks[i]); r.render_frame()
    for r in rs: r.sync()
    dt = time.perf_counter() - t0
    print(f"{wl} S={S}: {K/dt:.0f} frames/s ({dt/K*1e3:.4f} ms/frame)")
    for r in rs: r.close()
