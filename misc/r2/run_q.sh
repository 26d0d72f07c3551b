#!/bin/bash
# wavefront form of the FULL-lighting trace pass (trace_wave.cu) and the bricked occupancy field: parity first, then C3/C4/C5 against
# the sequential form on the same box
tag=${1:-r2_q}
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "wavefront or ray_pool" ) > gpurun_out/${tag}_pytest_wave.txt 2>&1; echo "wave pytest rc=$? $(tail -1 gpurun_out/${tag}_pytest_wave.txt)"
( timeout 900 python -m pytest tests/test_gpu_parity_sizes.py -m gpu -x -q -k "wavefront or c4" ) > gpurun_out/${tag}_pytest_sizes.txt 2>&1; echo "sizes pytest rc=$? $(tail -1 gpurun_out/${tag}_pytest_sizes.txt)"
for w in c3 c4 c5; do for g in 1 0; do
  VXRT_WAVE=$g timeout 600 python bench.py --workload $w --steps 24 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_${w}_wave$g.json 2> gpurun_out/${tag}_${w}_wave$g.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${tag}_${w}_wave$g.json")); print("$w wave=$g ms/step %.4f fps %.0f"%(d["ms_per_step"], d["frames_per_s"]), {k:round(v,4) for k,v in d["kernel_ms_per_step_rank0"].items()}, "launches", d["gpu_launches"])
except Exception as e: print("$w wave=$g ERR", e); print(open("gpurun_out/${tag}_${w}_wave$g.err").read()[-1500:])
PY
done; done
for f in; do
  VXRT_FIELD=$f timeout 600 python bench.py --workload c2 --steps 24 --no-cpu-baseline --sustained-seconds 0 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c2 field=$f trace', round(d['kernel_ms_per_step_rank0']['trace'],4))"
  VXRT_WAVE=0 VXRT_FIELD=$f timeout 600 python bench.py --workload c4 --steps 12 --no-cpu-baseline --sustained-seconds 0 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c4 (sequential) field=$f trace', round(d['kernel_ms_per_step_rank0']['trace'],4))"
done
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'k_trace|k_wave' -s 15 -c 5 --csv --log-file gpurun_out/${tag}_ncu_c3_wave.csv \
    python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_ncu_c3.log 2>&1; echo "ncu c3 rc=$?"
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/${tag}_ncu_c3_wave.csv")) if r and not r[0].startswith("==")]
h=rows[0]
cur=None
for r in rows[1:]:
    k=(r[h.index("ID")], r[h.index("Kernel Name")][:50])
    if k!=cur: print(); print(k, end=" "); cur=k
    print(r[h.index("Metric Name")].split(".")[0][-22:], r[h.index("Metric Value")], end=" | ")
print()
PY
