#!/bin/bash
# (E16) second-bounce hand-off: GPU tests (bounded), then C3 / C4 / C5 with and without it
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/s2_pytest.txt; tail -6 gpurun_out/s2_pytest.txt
for v in hand nohand; do
  if [ $v = nohand ]; then export VXRT_NO_HANDOFF=1; else unset VXRT_NO_HANDOFF; fi
  for w in c3 c4 c5; do
    timeout 120 python bench.py --no-cpu-baseline --workload $w --steps 16 > gpurun_out/s2_${w}_$v.json 2>gpurun_out/s2_${w}_$v.err || tail -3 gpurun_out/s2_${w}_$v.err
    python - <<PY
import json
try:
    d=json.load(open("gpurun_out/s2_${w}_$v.json")); k=d["kernel_ms_per_step_rank0"]
    print("$v $w fps=%.0f ms=%.4f | trace %.4f denoise %.4f | rays %.0f steps/ray %.3f"%(d["frames_per_s"],d["ms_per_step"],k["trace"],k["denoise"],d["rays_per_frame"],d["mean_steps_per_ray"]))
except Exception as e: print("$v $w failed",e)
PY
  done
done
