#!/bin/bash
# second-bounce hand-off (E16) at 32 registers against the sequential default; the bench line's new e2e.d2h_probe
mkdir -p gpurun_out
show() { python -c "import sys,json; d=json.loads(sys.stdin.read()); k=d['kernel_ms_per_step_rank0']; print('$1', 'trace', round(k['trace'],4), 'ms/step', round(d['ms_per_step'],4), 'probe', d['e2e'].get('d2h_probe'))"; }
B="--steps 24 --no-cpu-baseline --sustained-seconds 0"
for w in c3 c5 c4; do
  timeout 300 python bench.py --workload $w $B | show "$w sequential"
  VXRT_HANDOFF=1 timeout 300 python bench.py --workload $w $B | show "$w handoff"
done
timeout 300 python bench.py --workload c2 $B | show "c2"
