#!/bin/bash
mkdir -p gpurun_out
CUDA_LAUNCH_BLOCKING=1 timeout 300 python - <<'PY' 2>&1 | tail -30
import sys, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import raytrace_b200 as R
from raytrace_b200 import _native as N, scenes
from oracle import orc
import test_gpu_parity as T
orc.build()
noise = orc.blue_noise()
dims, kind, seed = scenes.WORLDS["c1"]
m, f = orc.build_minefield(orc.generate_world(*dims, kind, seed))
blk = T._frame_block(orc, dims, seed=7)
for flags in (0, N.FLAG_HIT_BUFFERS, N.FLAG_HIT_BUFFERS | N.FLAG_DENOISE | N.FLAG_FINALIZE):
    for mode in (0, 1, 2):
        try:
            with R.Renderer(256, 256, dims, mode, flags) as r:
                r.upload_blue_noise(noise); r.upload_world(m, f); r.set_frame(blk)
                r.render_frame()
                r.read_buffer(N.BUF_LIGHTING)
            print("flags", flags, "mode", mode, "ok", flush=True)
        except Exception as e:
            print("flags", flags, "mode", mode, "FAIL", str(e)[:160], flush=True)
            break
PY
timeout 300 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio --clock-control none -k regex:k_trace_fast -c 24 --csv --log-file gpurun_out/r1_c_c3_trace.csv python bench.py --no-cpu-baseline --workload c3 --steps 4 --warmup 3 > /dev/null 2>&1
tail -12 gpurun_out/r1_c_c3_trace.csv | cut -d, -f5,12-15
