#!/bin/bash
# launch priority of the passes after the trace pass, with two frames in flight: C2 and C3, on / off
mkdir -p gpurun_out
for pr in 1 0 1 0; do
  export VXRT_POST_PRIORITY=$pr
  for w in c2 c3; do
  timeout 300 python bench.py --no-cpu-baseline --workload $w --steps 32 --in-flight 2 > gpurun_out/r_${w}_p$pr.json 2>gpurun_out/r_${w}_p$pr.err || tail -5 gpurun_out/r_${w}_p$pr.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r_${w}_p$pr.json")); o=d["one_frame_in_flight"]; k=d["kernel_ms_per_step_rank0"]
    print("prio=$pr $w S=%d fps=%.0f ms=%.4f | serial ms=%.4f | trace %.4f denoise %.4f fin %.4f | e2e fps=%.0f"%(d["config"]["frames_in_flight"],d["frames_per_s"],d["ms_per_step"],o["ms_per_step"],k["trace"],k["denoise"],k["finalize"],d["e2e"]["frames_per_s"]))
except Exception as e: print("$w prio=$pr failed",e)
PY
  done
done
