mkdir -p gpurun_out
bash misc/sweep.sh u_base "" u_g4_mb5 "-DVXRT_DN_MINB=5" u_g4_mb6 "-DVXRT_DN_MINB=6" u_g2_mb6 "-DVXRT_DN_G=2 -DVXRT_DN_MINB=6" u_g2_mb5 "-DVXRT_DN_G=2 -DVXRT_DN_MINB=5" u_g2 "-DVXRT_DN_G=2"
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/sweep_c2_u_*.json")):
    try:
        d=json.load(open(f)); print(f, d["kernel_ms_per_step_rank0"])
    except Exception as e: print(f, e)
PY
