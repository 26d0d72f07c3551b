#!/bin/bash
# the whole GPU suite once, then the C4 test (1 GiB occupancy field: the allocation that used to cross a 4 GiB line in some processes) five times,
# and C4's trace time in five fresh processes
( timeout 1500 python -m pytest tests -m gpu -x -q ) 2>&1 | tail -2
for i in 1 2 3 4 5; do timeout 600 python -m pytest tests/test_gpu_parity_sizes.py -m gpu -q -k "c4" 2>&1 | grep -E "^E  |FAILED|passed|failed" | head -4; done
for i in 1 2 3 4 5; do timeout 300 python bench.py --workload c4 --steps 12 --no-cpu-baseline --sustained-seconds 0 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c4 trace', round(d['kernel_ms_per_step_rank0']['trace'],4))"; done
