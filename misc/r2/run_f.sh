#!/bin/bash
# A/B of library variants: run_f.sh TAG "lib1 lib2 ..." "workloads" ; first the GPU suite on each, then bench lines
tag=$1; libs=$2; wl=${3:-"c2 c3 c4"}
mkdir -p gpurun_out
for l in $libs; do
  name=$(basename $l .so)
  ( VXRT_LIB=$PWD/$l timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_${name}_pytest.txt 2>&1; echo "$name pytest rc=$? $(tail -1 gpurun_out/${tag}_${name}_pytest.txt)"
  for w in $wl; do
    VXRT_LIB=$PWD/$l timeout 600 python bench.py --workload $w --steps 32 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_${name}_$w.json 2> gpurun_out/${tag}_${name}_$w.err
    python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${tag}_${name}_$w.json")); print("$name $w ms/step %.4f 1fif %.4f"%(d["ms_per_step"], d["one_frame_in_flight"]["ms_per_step"]), {k:round(v,4) for k,v in d["kernel_ms_per_step_rank0"].items()})
except Exception as e: print("$name $w ERR", e)
PY
  done
done
