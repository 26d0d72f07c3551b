#!/bin/bash
tag=${1:-r2_k}
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_pytest.txt 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/${tag}_pytest.txt)"
for w in c3 c4 c5; do for g in pool nopool; do
  if [ $g = nopool ]; then export VXRT_NO_POOL=1; else unset VXRT_NO_POOL; fi
  timeout 600 python bench.py --workload $w --steps 32 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_${w}_$g.json 2> gpurun_out/${tag}_${w}_$g.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${tag}_${w}_$g.json")); print("$w $g ms/step %.4f fps %.0f"%(d["ms_per_step"], d["frames_per_s"]), {k:round(v,4) for k,v in d["kernel_ms_per_step_rank0"].items()}, "rays", d["rays_per_frame"])
except Exception as e: print("$w $g ERR", e); print(open("gpurun_out/${tag}_${w}_$g.err").read()[-800:])
PY
done; done
unset VXRT_NO_POOL
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct --clock-control none -k regex:k_trace -s 3 -c 1 --csv --log-file gpurun_out/${tag}_ncu_c3_trace.csv \
    python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_ncu_c3.log 2>&1; echo "ncu c3 rc=$?"
grep -v "^==" gpurun_out/${tag}_ncu_c3_trace.csv | cut -d, -f5,13,15 | tail -7
