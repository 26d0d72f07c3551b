#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -40 > gpurun_out/run_e2_pytest.txt; tail -3 gpurun_out/run_e2_pytest.txt
for w in c2 c3 c4; do
timeout 300 python bench.py --no-cpu-baseline --workload $w --steps 16 > gpurun_out/sweep_${w}_e2.json 2>gpurun_out/sweep_${w}_e2.err
done
python - <<PY
import json
for w in ("c2","c3","c4"):
    d=json.load(open("gpurun_out/sweep_%s_e2.json"%w)); print(w,"trace_ms=%.4f total=%.4f"%(d["kernel_ms_per_step_rank0"]["trace"],d["ms_per_step"]))
PY
