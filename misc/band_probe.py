"""Trace / denoise time of a band as a function of its height (1 GPU): where does strong scaling lose?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import raytrace_b200 as R
from raytrace_b200 import _native as N, scenes
from oracle import orc
dims, kind, seed = scenes.WORLDS["c2"]
W, H = 1920, 1080
noise = R.world.default_blue_noise()
cam = scenes.camera(dims, 0, 64)
blk = orc.fill_frame(cam["origin"], cam["heading"], cam["pitch"], cam["sun_angle"], 1)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for mode in (1, 2):
  for flags, name in ((N.FLAG_FINALIZE, "trace-only"), (N.FLAG_FINALIZE | N.FLAG_DENOISE, "with-denoise(halo traced)")):
    for rows in [(0, 64), (0, 136), (0, 208), (0, 392), (208, 376), (600, 1080), (0, 1080)]:
        with R.Renderer(W, H, dims, mode, flags, rows=rows) as r:
            r.upload_blue_noise(noise); r.generate_world(kind, seed); r.set_frame(blk)
            for fl in (True, False):
                ts = []
                for i in range(8):
                    if fl: flush.fill_(i); torch.cuda.synchronize()
                    r.render_frame(); s = r.frame_stats()
                    if i >= 3: ts.append((s.trace_ms, s.denoise_ms, s.finalize_ms))
                t = np.median(np.array(ts), axis=0)
                print(f"mode={mode} {name:28s} rows={rows} flush={fl}: trace {t[0]:.4f} denoise {t[1]:.4f} finalize {t[2]:.4f}  rays={s.rays}")
