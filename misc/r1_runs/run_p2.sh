#!/bin/bash
# final-state multi-GPU check: the multi-process tests, then the scaling bench lines (frames dealt to GPUs + row bands)
N=$1; tag=r1_scale_t
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -5
for wl in c2 c3 c4 c5; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload $wl --steps 32 --no-cpu-baseline > gpurun_out/${tag}_n${N}_${wl}.json 2> gpurun_out/${tag}_n${N}_${wl}.err; echo "bench N=$N $wl rc=$?"
done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/${tag}_n${N}_*.json")):
    try:
        d=json.load(open(f)); print(f, d["scaling"], "Mrays/s %.0f ms %.4f fps %.0f e2e fps %.0f"%(d["value"], d["ms_per_step"], d["frames_per_s"], d["e2e"]["frames_per_s"]), "bands:", (d.get("image_bands") or {}).get("ms_per_step"))
    except Exception as e: print(f, "failed", e)
PY
