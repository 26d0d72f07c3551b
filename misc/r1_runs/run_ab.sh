N=$1; tag=$2
for v in first whole; do
  if [ $v = whole ]; then export VXRT_PEER_TRACE_WHOLE=1; else unset VXRT_PEER_TRACE_WHOLE; fi
  for wl in c2 c3; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --comm peer --workload $wl --steps 32 --no-cpu-baseline > gpurun_out/${tag}_n${N}_${v}_${wl}.json 2> gpurun_out/${tag}_n${N}_${v}_${wl}.err; echo "bench $v rc=$?"
  done
done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/${tag}_n*.json")):
    try:
        d=json.load(open(f)); print(f, "ms %.4f e2e fps %.0f"%(d["ms_per_step"], d["e2e"]["frames_per_s"])); [print("   ", x) for x in d["kernel_ms_per_step_by_rank"]]
    except Exception as e: print(f, "failed", e)
PY
