#!/bin/bash
# usage: misc/prof.sh TAG   — tests, bench lines, ncu launch list and one full-set capture (1 GPU)
tag=$1
mkdir -p gpurun_out
python raytrace_b200/build.py > /dev/null
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${tag}_pytest_gpu.log
python bench.py > gpurun_out/${tag}_bench_c2.json 2> gpurun_out/${tag}_bench_c2.err; echo "bench c2 rc=$?"
python bench.py --workload c3 --steps 32 --no-cpu-baseline > gpurun_out/${tag}_bench_c3.json 2> gpurun_out/${tag}_bench_c3.err; echo "bench c3 rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; echo "bench ref rc=$?"
# launch list of the bench command (cold-cache, serialised: shares must agree, not absolutes)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu_launches.log 2>&1; echo "ncu list rc=$?"
# one full-set capture of the step's kernels (skip warm-up launches: 3 warm-up frames x 8 kernels + setup)
ncu --set full --clock-control none --import-source on -k regex:'k_trace|k_denoise|k_finalize' -s 24 -c 8 -f -o gpurun_out/${tag}_full \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
