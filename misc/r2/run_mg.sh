#!/bin/bash
# multi-GPU: the 2-process tests (peer IPC frame, NCCL gather + edit broadcast) and the bench at N GPUs with the band check
N=${1:-2}; tag=${2:-r2_mg$N}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${tag}_gpus.txt
( timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -v ) > gpurun_out/${tag}_pytest_multi_gpu.txt 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/${tag}_pytest_multi_gpu.txt
for w in c2 c3 c4 c5; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --workload $w --steps 32 --sustained-seconds 0 > gpurun_out/${tag}_bench_$w.json 2> gpurun_out/${tag}_bench_$w.err; echo "bench $w rc=$?"
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${tag}_bench_$w.json"))
    print("$w N=$N", d["scaling"], "ms/step %.4f fps %.0f e2e fps %.0f"%(d["ms_per_step"], d["frames_per_s"], d["e2e"]["frames_per_s"]), "match", d.get("matches_single_gpu"))
    b=d.get("image_bands")
    if b: print("   bands ms %.4f match %s rows_differing %s"%(b["ms_per_step"], b.get("matches_single_gpu"), b.get("rows_differing")), [ (round(k["trace"],3),round(k["denoise"],3),round(k["finalize"],3)) for k in b["kernel_ms_per_step_by_rank"]])
    else: print("   by rank", [ (round(k["trace"],3),round(k["denoise"],3),round(k["finalize"],3)) for k in d["kernel_ms_per_step_by_rank"]])
except Exception as e: print("$w ERR", e); print(open("gpurun_out/${tag}_bench_$w.err").read()[-1500:])
PY
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/${tag}_bench_ref.json
