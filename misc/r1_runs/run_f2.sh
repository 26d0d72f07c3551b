#!/bin/bash
mkdir -p gpurun_out
bash misc/sweep.sh skyinl "" skycall "-DVXRT_SKY_INLINE=0"
python raytrace_b200/build.py --force > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_trace' -s 3 -c 1 -f -o gpurun_out/r1_h_trace python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1_h_ncu.log 2>&1; echo "ncu rc=$?"
