#!/bin/bash
# shared accumulation chain (C3 with two frames in flight): tests, then C3 at S = 1 / 2 / 3
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/q_pytest.txt; tail -5 gpurun_out/q_pytest.txt
for S in 1 2 3; do
  timeout 300 python bench.py --no-cpu-baseline --workload c3 --steps 32 --in-flight $S > gpurun_out/q_c3_s$S.json 2>gpurun_out/q_c3_s$S.err || tail -5 gpurun_out/q_c3_s$S.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/q_c3_s$S.json")); o=d["one_frame_in_flight"]; k=d["kernel_ms_per_step_rank0"]
    print("S=%d fps=%.0f ms=%.4f | serial ms=%.4f | trace %.4f denoise %.4f fin %.4f | e2e fps=%.0f"%(d["config"]["frames_in_flight"],d["frames_per_s"],d["ms_per_step"],o["ms_per_step"],k["trace"],k["denoise"],k["finalize"],d["e2e"]["frames_per_s"]))
except Exception as e: print("S=$S failed",e)
PY
done
