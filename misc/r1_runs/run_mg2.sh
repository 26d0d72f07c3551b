# usage: misc/run_mg2.sh TAG N [N ...]  — peer-mode bench at each N (c2, c3), per-rank breakdown
tag=$1; shift
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for N in "$@"; do
  for wl in c2 c3; do
    if [ $N = 1 ]; then timeout 200 python bench.py --steps 32 --no-cpu-baseline --workload $wl > gpurun_out/${tag}_n1_peer_${wl}.json 2>gpurun_out/${tag}_n1_peer_${wl}.err
    else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --comm peer --workload $wl --steps 32 --no-cpu-baseline > gpurun_out/${tag}_n${N}_peer_${wl}.json 2> gpurun_out/${tag}_n${N}_peer_${wl}.err; fi
    echo "bench N=$N $wl rc=$?"
  done
done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/${tag}_n*.json")):
    try:
        d=json.load(open(f)); print(f, "Mrays/s %.0f ms %.4f fps %.0f e2e fps %.0f"%(d["value"], d["ms_per_step"], d["frames_per_s"], d["e2e"]["frames_per_s"])); [print("   ", x) for x in d["kernel_ms_per_step_by_rank"]]
    except Exception as e: print(f, "failed", e)
PY
