mkdir -p gpurun_out
python raytrace_b200/build.py --force > /dev/null
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
bash misc/sweep.sh g_ty2_16_8 "" g_ty1_8_8 "-DVXRT_DN_TY=1 -DVXRT_DN_L0=8 -DVXRT_DN_L1=8" g_ty1_16_8 "-DVXRT_DN_TY=1 -DVXRT_DN_L0=16 -DVXRT_DN_L1=8" g_ty4_16_16 "-DVXRT_DN_TY=4 -DVXRT_DN_L0=16 -DVXRT_DN_L1=16" g_ty2_8_8 "-DVXRT_DN_TY=2 -DVXRT_DN_L0=8 -DVXRT_DN_L1=8" g_ty1_12_8 "-DVXRT_DN_TY=1 -DVXRT_DN_L0=12 -DVXRT_DN_L1=8"
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/sweep_c2_g_*.json")):
    try:
        d=json.load(open(f)); print(f, d["kernel_ms_per_step_rank0"])
    except Exception as e: print(f, e)
PY
