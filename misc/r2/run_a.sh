#!/bin/bash
# round-2 baseline: the whole GPU suite (with the at-size parity tests) + bench lines of every workload on this round's box
tag=r2_a
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${tag}_smi.txt
nproc >> gpurun_out/${tag}_smi.txt; free -g >> gpurun_out/${tag}_smi.txt
( time timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 ) > gpurun_out/${tag}_pytest.txt 2>&1; echo "pytest rc=$?"
timeout 600 python bench.py > gpurun_out/${tag}_bench_c2.json 2> gpurun_out/${tag}_bench_c2.err; echo "bench c2 rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; echo "bench ref rc=$?"
for w in c3 c4 c5 c1; do timeout 600 python bench.py --workload $w --steps 32 > gpurun_out/${tag}_bench_$w.json 2> gpurun_out/${tag}_bench_$w.err; echo "bench $w rc=$?"; done
tail -5 gpurun_out/${tag}_pytest.txt
