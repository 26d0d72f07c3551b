#!/bin/bash
# C1 (trace pass out-of-loop work): tests, bench A/B against run-time dims, ncu instruction counts of the C2 trace pass
tag=${1:-r2_c}
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_pytest.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest.txt
for w in c2 c3 c4; do
  timeout 600 python bench.py --workload $w --steps 32 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_bench_$w.json 2> gpurun_out/${tag}_bench_$w.err; echo "bench $w rc=$?"
  VXRT_NO_DIMS=1 timeout 600 python bench.py --workload $w --steps 32 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_bench_${w}_nodims.json 2> gpurun_out/${tag}_bench_${w}_nodims.err; echo "bench $w nodims rc=$?"
done
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:k_trace -s 3 -c 2 --csv --log-file gpurun_out/${tag}_ncu_c2_trace.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_ncu_c2.log 2>&1; echo "ncu c2 rc=$?"
python - <<PY
import json
for w in ("c2","c3","c4"):
    for v in ("","_nodims"):
        try:
            d=json.load(open(f"gpurun_out/${tag}_bench_{w}{v}.json")); print(w,v,"ms/step %.4f"%d["ms_per_step"], "1fif %.4f"%d["one_frame_in_flight"]["ms_per_step"], d["kernel_ms_per_step_rank0"])
        except Exception as e: print(w,v,"ERR",e)
PY
grep -v "^==" gpurun_out/${tag}_ncu_c2_trace.csv | tail -9
