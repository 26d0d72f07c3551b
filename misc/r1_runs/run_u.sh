#!/bin/bash
# vxrt_apply_edits without the stream synchronize: GPU tests, then C5 (device and end-to-end frames/s)
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 150 python bench.py --no-cpu-baseline --workload c5 --steps 32 > gpurun_out/u_c5.json 2>gpurun_out/u_c5.err || tail -3 gpurun_out/u_c5.err
python - <<PY
import json
d=json.load(open("gpurun_out/u_c5.json")); k=d["kernel_ms_per_step_rank0"]
print("c5 fps=%.0f ms=%.4f e2e fps=%.0f | trace %.4f denoise %.4f"%(d["frames_per_s"],d["ms_per_step"],d["e2e"]["frames_per_s"],k["trace"],k["denoise"]))
PY
