#!/bin/bash
# AUTO hand-off policy: its test, C5 / C3 / C4 with and without it, then the whole suite
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "auto_handoff" 2>&1 | grep -E "^E  |passed|failed" | head -6
show() { python -c "import sys,json; d=json.loads(sys.stdin.read()); k=d['kernel_ms_per_step_rank0']; print('$1', 'trace', round(k['trace'],4), 'ms/step', round(d['ms_per_step'],4), 'fps', round(d['frames_per_s']), 'e2e', round(d['e2e']['frames_per_s']))"; }
B="--steps 32 --no-cpu-baseline --sustained-seconds 0"
for w in c5 c3 c4; do
  timeout 300 python bench.py --workload $w $B | show "$w auto"
  VXRT_NO_AUTO_HANDOFF=1 timeout 300 python bench.py --workload $w $B | show "$w no-auto"
done
( timeout 1500 python -m pytest tests -m gpu -x -q ) 2>&1 | tail -2
