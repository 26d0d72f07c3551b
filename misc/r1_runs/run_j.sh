#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -40 > gpurun_out/run_j_pytest.txt; tail -5 gpurun_out/run_j_pytest.txt
for w in c2 c4 c1 c3; do
timeout 300 python bench.py --no-cpu-baseline --workload $w --steps 32 > gpurun_out/sweep_${w}_j.json 2>gpurun_out/sweep_${w}_j.err || tail -5 gpurun_out/sweep_${w}_j.err
done
timeout 300 python bench.py --no-cpu-baseline --in-flight 3 --steps 32 > gpurun_out/sweep_c2_j3.json 2>gpurun_out/sweep_c2_j3.err
python - <<PY
import json
for w in ("c2","c4","c1","c3","c2_j3"):
    try:
        f = "gpurun_out/sweep_%s_j.json"%w if not w.endswith("j3") else "gpurun_out/sweep_c2_j3.json"
        d=json.load(open(f)); o=d["one_frame_in_flight"]
        print(w,"S=%d value=%.0f fps=%.0f ms=%.4f | serial fps=%.0f ms=%.4f | e2e fps=%.0f"%(d["config"]["frames_in_flight"],d["value"],d["frames_per_s"],d["ms_per_step"],o["frames_per_s"],o["ms_per_step"],d["e2e"]["frames_per_s"]))
    except Exception as e: print(w,"failed",e)
PY
