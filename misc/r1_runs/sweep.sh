#!/bin/bash
# usage: sweep.sh tag "<nvcc extra>" ...   (pairs)
mkdir -p gpurun_out
while [ $# -gt 0 ]; do
  tag=$1; extra=$2; shift 2
  VXRT_NVCC_EXTRA="$extra" python raytrace_b200/build.py --force > /dev/null 2>gpurun_out/build_$tag.err || { echo "build $tag failed"; continue; }
  timeout 200 python bench.py --no-cpu-baseline --steps 32 > gpurun_out/sweep_c2_$tag.json 2>gpurun_out/sweep_c2_$tag.err
  timeout 200 python bench.py --no-cpu-baseline --workload c3 --steps 16 > gpurun_out/sweep_c3_$tag.json 2>gpurun_out/sweep_c3_$tag.err
  python - <<PY
import json
for w in ("c2","c3"):
    try:
        d=json.load(open("gpurun_out/sweep_%s_$tag.json"%w))
        print("$tag",w,"trace_ms=%.4f total_ms=%.4f Mrays/s=%.0f"%(d["kernel_ms_per_step_rank0"]["trace"],d["ms_per_step"],d["value"]))
    except Exception as e: print("$tag",w,"failed",e)
PY
done
