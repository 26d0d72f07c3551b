#!/bin/bash
# where the FULL-lighting trace pass (C3, C4) stalls: one full-set capture of k_trace_fast per workload
mkdir -p gpurun_out
for w in c3 c4; do
ncu --set full --clock-control none --import-source on -k regex:'k_trace' -s 3 -c 1 -f -o gpurun_out/r1_o_${w}_trace \
    python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1_o_${w}_ncu.log 2>&1; echo "ncu $w rc=$?"
done
