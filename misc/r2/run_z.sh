#!/bin/bash
# frames in flight per GPU: 1 / 2 / 3 / 4 on C2, C1 and C4 (frames of a path are independent there)
show() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', 'ms/step', round(d['ms_per_step'],4), 'fps', round(d['frames_per_s']), 'e2e fps', round(d['e2e']['frames_per_s']))"; }
B="--steps 48 --no-cpu-baseline --sustained-seconds 0"
for w in c2 c1 c4; do for s in 1 2 3 4; do timeout 300 python bench.py --workload $w --in-flight $s $B | show "$w in-flight $s"; done; done
