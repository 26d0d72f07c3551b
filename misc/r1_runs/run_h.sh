#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -40 > gpurun_out/run_h_pytest.txt; tail -5 gpurun_out/run_h_pytest.txt
for w in c2 c3; do
timeout 300 python bench.py --no-cpu-baseline --workload $w --steps 16 > gpurun_out/sweep_${w}_h.json 2>gpurun_out/sweep_${w}_h.err
done
python - <<PY
import json
for w in ("c2","c3"):
    d=json.load(open("gpurun_out/sweep_%s_h.json"%w)); print(w,"trace_ms=%.4f total=%.4f"%(d["kernel_ms_per_step_rank0"]["trace"],d["ms_per_step"]))
PY
