mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_gpu.log
bash misc/sweep.sh e_l16_8 "" e_l8_8 "-DVXRT_DN_L0=8 -DVXRT_DN_L1=8" e_l32_8 "-DVXRT_DN_L0=32 -DVXRT_DN_L1=8" e_l32_16 "-DVXRT_DN_L0=32 -DVXRT_DN_L1=16" e_l24_8 "-DVXRT_DN_L0=24 -DVXRT_DN_L1=8"
