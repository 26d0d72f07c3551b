# usage: misc/run_mg.sh N TAG
N=$1; tag=$2
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/${tag}_topo.txt 2>&1
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/${tag}_pytest_mg.log 2>&1; echo "pytest mg rc=$?"; tail -5 gpurun_out/${tag}_pytest_mg.log
for comm in peer nccl; do
  for wl in c2 c3; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --comm $comm --workload $wl --steps 32 --no-cpu-baseline > gpurun_out/${tag}_n${N}_${comm}_${wl}.json 2> gpurun_out/${tag}_n${N}_${comm}_${wl}.err; echo "bench $comm $wl rc=$?"
  done
done
timeout 200 python bench.py --steps 32 --no-cpu-baseline > gpurun_out/${tag}_n1_c2.json 2>gpurun_out/${tag}_n1_c2.err
timeout 200 python bench.py --steps 32 --no-cpu-baseline --workload c3 > gpurun_out/${tag}_n1_c3.json 2>gpurun_out/${tag}_n1_c3.err
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/${tag}_n*.json")):
    try:
        d=json.load(open(f)); print(f, "Mrays/s %.0f fps %.0f ms %.4f e2e fps %.0f"%(d["value"], d["frames_per_s"], d["ms_per_step"], d["e2e"]["frames_per_s"]), d["kernel_ms_per_step_rank0"])
    except Exception as e: print(f, "failed", e)
PY
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/${tag}_n*_peer_*.json")):
    try:
        d=json.load(open(f)); print(f); [print("   ", x) for x in d["kernel_ms_per_step_by_rank"]]
    except Exception as e: print(f, "failed", e)
PY
