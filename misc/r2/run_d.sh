#!/bin/bash
mkdir -p gpurun_out
for v in "A=1"; do
echo "== $v"
env $v timeout 600 python - <<'PY' 2>&1 | tail -4
import numpy as np, sys
sys.path.insert(0,'.')
import raytrace_b200 as R
from raytrace_b200 import _native as N, scenes
from oracle import orc
dims,kind,seed=scenes.WORLDS["c1"]
m=orc.generate_world(*dims,kind,seed); m,f=orc.build_minefield(m)
noise=orc.blue_noise()
cam=scenes.camera(dims)
blk=orc.fill_frame(cam["origin"],cam["heading"],cam["pitch"],cam["sun_angle"],7)
for mode in (0,1,2):
    ref=orc.render(m,f,noise,blk,mode,256,256)
    with R.Renderer(256,256,dims,mode,N.FLAG_HIT_BUFFERS) as r:
        r.upload_blue_noise(noise); r.upload_world(m,f); r.set_frame(blk); r.render_frame()
        st=r.read_buffer(N.BUF_STEPS); ry=r.read_buffer(N.BUF_RAYS)
        s=r.frame_stats()
        print("mode",mode,"steps mismatch px",(st!=ref.steps).sum(),"rays mismatch px",(ry!=ref.rays).sum(),"dev rays/steps/hits",s.rays,s.steps,s.hits,"ref",ref.stats.rays,ref.stats.steps,ref.stats.hits)
PY
done
