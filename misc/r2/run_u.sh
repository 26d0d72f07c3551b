#!/bin/bash
# one-flag variants on one box: L2 fetch granularity of the occupancy load, denoise residency / staging unroll, wavefront at 32 registers
mkdir -p gpurun_out
show() { python -c "import sys,json; d=json.loads(sys.stdin.read()); k=d['kernel_ms_per_step_rank0']; print('$1', 'trace', round(k['trace'],4), 'denoise', round(k['denoise'],4), 'ms/step', round(d['ms_per_step'],4), 'one-in-flight', round(d['one_frame_in_flight']['ms_per_step'],4))"; }
B="--steps 24 --no-cpu-baseline --sustained-seconds 0"
for w in c2 c3 c5; do
  for l in libvxrt libvxrt_l2_128 libvxrt_l2_256; do VXRT_LIB=$PWD/raytrace_b200/$l.so timeout 300 python bench.py --workload $w $B | show "$w $l"; done
done
for l in libvxrt libvxrt_l2_128 libvxrt_l2_256; do VXRT_FIELD=lin VXRT_LIB=$PWD/raytrace_b200/$l.so timeout 300 python bench.py --workload c4 $B | show "c4 lin $l"; done
timeout 300 python bench.py --workload c4 $B | show "c4 tex libvxrt"
for w in c3 c5; do VXRT_WAVE=1 timeout 300 python bench.py --workload $w $B | show "$w wavefront"; done
for l in libvxrt_dn_minb6 libvxrt_dn_unroll4 libvxrt_dn_unroll1; do VXRT_LIB=$PWD/raytrace_b200/$l.so timeout 300 python bench.py --workload c2 $B | show "c2 $l"; done
