#!/bin/bash
# final-state evidence of round 2 on one box (k_trace_fast at 32 registers): the whole GPU suite, bench lines of every workload + the reference arm,
# ncu launch list of the bench command, full-set captures (source counters) of one C2 frame and of the C3 / C4 trace pass
tag=${1:-r2_w}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${tag}_smi.txt
( time timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 ) > gpurun_out/${tag}_pytest.txt 2>&1; echo "pytest rc=$? $(grep -E 'passed|failed' gpurun_out/${tag}_pytest.txt | tail -1)"
timeout 600 python bench.py > gpurun_out/${tag}_bench_c2.json 2> gpurun_out/${tag}_bench_c2.err; echo "bench c2 rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; echo "bench ref rc=$?"
for w in c3 c4 c5 c1; do timeout 600 python bench.py --workload $w --steps 32 > gpurun_out/${tag}_bench_$w.json 2> gpurun_out/${tag}_bench_$w.err; echo "bench $w rc=$?"; done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_trace|k_denoise|k_finalize' -s 24 -c 8 -f -o gpurun_out/${tag}_c2_full \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full c2 rc=$?"
for w in c3 c4; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_trace' -s 3 -c 1 -f -o gpurun_out/${tag}_${w}_trace \
    python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_${w}_ncu.log 2>&1; echo "ncu $w rc=$?"
done
ls -la gpurun_out/*.ncu-rep
