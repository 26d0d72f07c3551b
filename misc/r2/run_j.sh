#!/bin/bash
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "ray_pool" ) > gpurun_out/r2_j_pytest_pool.txt 2>&1; echo "pool pytest rc=$?"; tail -25 gpurun_out/r2_j_pytest_pool.txt
