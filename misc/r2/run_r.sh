#!/bin/bash
# wavefront variants (built by `python raytrace_b200/build.py --variant ...`): trace time of C3 / C4 / C5 per variant, sequential beside it
mkdir -p gpurun_out
show() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', round(d['kernel_ms_per_step_rank0']['trace'],4), 'ms trace,', round(d['ms_per_step'],4), 'ms/step')"; }
for w in c3 c5 c4; do
  VXRT_WAVE=0 timeout 300 python bench.py --workload $w --steps 16 --no-cpu-baseline --sustained-seconds 0 | show "$w sequential"
  for l in raytrace_b200/libvxrt.so raytrace_b200/libvxrt_w_*.so; do
    VXRT_LIB=$PWD/$l timeout 300 python bench.py --workload $w --steps 16 --no-cpu-baseline --sustained-seconds 0 | show "$w $(basename $l .so)"
  done
done
