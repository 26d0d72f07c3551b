#!/bin/bash
# two-launch trace (second bounce level from a queue): parity first, then trace time on C2/C3/C4, then per-stage ncu times
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -40 > gpurun_out/run_c_pytest.txt; tail -3 gpurun_out/run_c_pytest.txt
timeout 200 python bench.py --no-cpu-baseline --steps 32 > gpurun_out/sweep_c2_split.json 2>gpurun_out/sweep_c2_split.err
timeout 200 python bench.py --no-cpu-baseline --workload c3 --steps 16 > gpurun_out/sweep_c3_split.json 2>gpurun_out/sweep_c3_split.err
timeout 300 python bench.py --no-cpu-baseline --workload c4 --steps 8 > gpurun_out/sweep_c4_split.json 2>gpurun_out/sweep_c4_split.err
python - <<PY
import json
for w in ("c2","c3","c4"):
    d=json.load(open("gpurun_out/sweep_%s_split.json"%w)); print(w,"split trace_ms=%.4f total=%.4f"%(d["kernel_ms_per_step_rank0"]["trace"],d["ms_per_step"]))
PY
timeout 300 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio --clock-control none -k regex:k_trace_fast -c 16 --csv --log-file gpurun_out/r1_c_c3_trace.csv python bench.py --no-cpu-baseline --workload c3 --steps 4 --warmup 3 > /dev/null 2>&1
tail -4 gpurun_out/r1_c_c3_trace.csv | python -c "
import csv,sys
for r in csv.reader(sys.stdin): print(r[4][:40], r[-4:])"
