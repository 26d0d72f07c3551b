# usage: misc/run_mg3.sh TAG "N N .." "wl wl .."  — peer-mode bench per N and workload, per-rank breakdown
tag=$1; Ns=$2; WLs=$3
mkdir -p gpurun_out
for N in $Ns; do
  for wl in $WLs; do
    if [ $N = 1 ]; then timeout 300 python bench.py --steps 16 --no-cpu-baseline --workload $wl > gpurun_out/${tag}_n1_${wl}.json 2>gpurun_out/${tag}_n1_${wl}.err
    else timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload $wl --steps 16 --no-cpu-baseline > gpurun_out/${tag}_n${N}_${wl}.json 2> gpurun_out/${tag}_n${N}_${wl}.err; fi
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
