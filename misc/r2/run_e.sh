#!/bin/bash
# full-set ncu captures (with source counters) of the trace pass on C2, C3, C4 and of the denoise passes on C2
tag=${1:-r2_e}
mkdir -p gpurun_out
for w in c2 c3 c4; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_trace' -s 3 -c 1 -f -o gpurun_out/${tag}_${w}_trace \
    python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_${w}_ncu.log 2>&1; echo "ncu $w rc=$?"
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_denoise|k_finalize' -s 21 -c 7 -f -o gpurun_out/${tag}_c2_post \
    python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_c2_post_ncu.log 2>&1; echo "ncu post rc=$?"
ls -la gpurun_out/*.ncu-rep
