#!/bin/bash
# programmatic dependent launch of denoise passes 2..6: A/B by env on one box
mkdir -p gpurun_out
for v in pdl nopdl; do
  if [ $v = nopdl ]; then export VXRT_NO_PDL=1; else unset VXRT_NO_PDL; fi
  for w in c2 c3; do
    timeout 300 python bench.py --no-cpu-baseline --workload $w --steps 32 > gpurun_out/l_${w}_$v.json 2>gpurun_out/l_${w}_$v.err || tail -5 gpurun_out/l_${w}_$v.err
  done
done
unset VXRT_NO_PDL
python - <<PY
import json
for v in ("pdl","nopdl"):
  for w in ("c2","c3"):
    try:
        d=json.load(open("gpurun_out/l_%s_%s.json"%(w,v))); o=d["one_frame_in_flight"]; k=d["kernel_ms_per_step_rank0"]
        print(v,w,"S=%d fps=%.0f ms=%.4f | serial ms=%.4f | trace %.4f denoise %.4f fin %.4f | e2e fps=%.0f"%(d["config"]["frames_in_flight"],d["frames_per_s"],d["ms_per_step"],o["ms_per_step"],k["trace"],k["denoise"],k["finalize"],d["e2e"]["frames_per_s"]))
    except Exception as e: print(v,w,"failed",e)
PY
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/l_pytest.txt; tail -5 gpurun_out/l_pytest.txt
