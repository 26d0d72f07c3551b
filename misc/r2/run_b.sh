#!/bin/bash
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_parity_sizes.py -m gpu -q --durations=15 ) > gpurun_out/r2_b_pytest.txt 2>&1; echo "pytest rc=$?"
tail -40 gpurun_out/r2_b_pytest.txt
