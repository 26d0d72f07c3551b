#!/bin/bash
# round-end check of the committed state: smoke, GPU tests, the default bench line and the reference arm
mkdir -p gpurun_out
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 200 python bench.py > gpurun_out/final_bench_c2.json 2> gpurun_out/final_bench_c2.err; echo "bench rc=$?"
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/final_bench_ref.json 2> gpurun_out/final_bench_ref.err; echo "ref rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/final_bench_c2.json")); k=d["kernel_ms_per_step_rank0"]
print("c2 value=%.0f fps=%.0f ms=%.4f serial=%.4f e2e=%.0f trace %.4f denoise %.4f roof %.3f cpu %.2f launches %s clocks %s"%(d["value"],d["frames_per_s"],d["ms_per_step"],d["one_frame_in_flight"]["ms_per_step"],d["e2e"]["frames_per_s"],k["trace"],k["denoise"],d["roofline"]["frac"],d["cpu_baseline"]["value"],d["gpu_launches"],d["clocks"]))
r=json.load(open("gpurun_out/final_bench_ref.json")); print("ref", r["value"], r["unit"], r["cpu_baseline"])
PY
