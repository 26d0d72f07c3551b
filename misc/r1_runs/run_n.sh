#!/bin/bash
# trace march (E14)-(E15) + packed counter reduction: A/B against the previous build on one box, then the GPU tests
mkdir -p gpurun_out
for v in new base new base; do
  if [ $v = base ]; then export VXRT_LIB=$PWD/raytrace_b200/libvxrt_base.so; else unset VXRT_LIB; fi
  for w in c2 c3 c4; do
    timeout 300 python bench.py --no-cpu-baseline --workload $w --steps 32 > gpurun_out/n_${w}_$v.json 2>gpurun_out/n_${w}_$v.err || tail -5 gpurun_out/n_${w}_$v.err
    python - <<PY
import json
try:
    d=json.load(open("gpurun_out/n_${w}_$v.json")); o=d["one_frame_in_flight"]; k=d["kernel_ms_per_step_rank0"]
    print("$v","$w","S=%d fps=%.0f ms=%.4f | serial ms=%.4f | trace %.4f denoise %.4f fin %.4f | e2e fps=%.0f"%(d["config"]["frames_in_flight"],d["frames_per_s"],d["ms_per_step"],o["ms_per_step"],k["trace"],k["denoise"],k["finalize"],d["e2e"]["frames_per_s"]))
except Exception as e: print("$v","$w","failed",e)
PY
  done
done
unset VXRT_LIB
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/n_pytest.txt; tail -5 gpurun_out/n_pytest.txt
