#!/bin/bash
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_i_pytest.txt 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/r2_i_pytest.txt)"
for w in c1 c2 c3; do for g in graph nograph; do
  if [ $g = nograph ]; then export VXRT_NO_GRAPH=1; else unset VXRT_NO_GRAPH; fi
  timeout 600 python bench.py --workload $w --steps 64 --no-cpu-baseline --sustained-seconds 1 > gpurun_out/r2_i_${w}_$g.json 2> gpurun_out/r2_i_${w}_$g.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2_i_${w}_$g.json")); print("$w $g ms/step %.4f fps %.0f e2e fps %.0f sustained fps %.0f 1fif %.4f"%(d["ms_per_step"], d["frames_per_s"], d["e2e"]["frames_per_s"], d["sustained"]["frames_per_s"], d["one_frame_in_flight"]["ms_per_step"]), {k:round(v,4) for k,v in d["kernel_ms_per_step_rank0"].items()})
except Exception as e: print("$w $g ERR", e); print(open("gpurun_out/r2_i_${w}_$g.err").read()[-800:])
PY
done; done
