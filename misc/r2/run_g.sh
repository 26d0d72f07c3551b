#!/bin/bash
# layout A/B: occupancy field through LDG (linear), TLD (integer coords, block-linear array), TEX (normalised + wrap)
tag=${1:-r2_g}
mkdir -p gpurun_out
for f in lin tex texn; do
  export VXRT_FIELD=$f
  ( timeout 900 python -m pytest tests/test_gpu_parity_sizes.py -m gpu -x -q -k "c2_at_size or c3_at_size or c4_at_size" ) > gpurun_out/${tag}_${f}_pytest.txt 2>&1; echo "$f pytest rc=$? $(tail -1 gpurun_out/${tag}_${f}_pytest.txt)"
  for w in c2 c3 c4; do
    timeout 600 python bench.py --workload $w --steps 32 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_${f}_$w.json 2> gpurun_out/${tag}_${f}_$w.err
    python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${tag}_${f}_$w.json")); print("$f $w ms/step %.4f 1fif %.4f"%(d["ms_per_step"], d["one_frame_in_flight"]["ms_per_step"]), {k:round(v,4) for k,v in d["kernel_ms_per_step_rank0"].items()})
except Exception as e: print("$f $w ERR", e)
PY
  done
done
