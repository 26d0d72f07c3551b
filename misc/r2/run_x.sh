#!/bin/bash
# second-bounce hand-off (E16) with the window dims baked in (and through the texture unit on the C4 window), 32 registers, against the sequential default
mkdir -p gpurun_out
show() { python -c "import sys,json; d=json.loads(sys.stdin.read()); k=d['kernel_ms_per_step_rank0']; print('$1', 'trace', round(k['trace'],4), 'ms/step', round(d['ms_per_step'],4))"; }
B="--steps 24 --no-cpu-baseline --sustained-seconds 0"
for w in c3 c5 c4; do
  timeout 300 python bench.py --workload $w $B | show "$w sequential"
  VXRT_HANDOFF=1 timeout 300 python bench.py --workload $w $B | show "$w handoff"
done
VXRT_HANDOFF=1 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_sizes.py -m gpu -x -q -k "handoff or c3 or c4 or accumul" 2>&1 | tail -1
