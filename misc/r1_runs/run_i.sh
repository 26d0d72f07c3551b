#!/bin/bash
# usage: run_i.sh N
N=$1
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 32 --warmup 4 > gpurun_out/scale_i_n${N}_c2.json 2> gpurun_out/scale_i_n${N}_c2.err || tail -20 gpurun_out/scale_i_n${N}_c2.err
python - <<PY
import json
d=json.load(open("gpurun_out/scale_i_n${N}_c2.json"))
print("frames: value=%.0f fps=%.0f ms/step=%.4f e2e fps=%.0f scaling=%s"%(d["value"],d["frames_per_s"],d["ms_per_step"],d["e2e"]["frames_per_s"],d["scaling"]))
b=d.get("image_bands")
if b: print("bands: value=%.0f fps=%.0f ms/step=%.4f e2e fps=%.0f"%(b["value"],b["frames_per_s"],b["ms_per_step"],b["e2e"]["frames_per_s"]))
PY
