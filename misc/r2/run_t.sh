#!/bin/bash
# k_trace_fast at 32 registers (16 CTAs of 4 warps per SM = 64 warps instead of 48; the march loop grows from 36 to 39 slots): trace time per workload
mkdir -p gpurun_out
show() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', round(d['kernel_ms_per_step_rank0']['trace'],4), 'ms trace,', round(d['ms_per_step'],4), 'ms/step', round(d['frames_per_s']), 'fps')"; }
for w in c2 c3 c5 c4 c1; do
  for l in raytrace_b200/libvxrt.so raytrace_b200/libvxrt_t_*.so; do
    VXRT_LIB=$PWD/$l timeout 300 python bench.py --workload $w --steps 24 --no-cpu-baseline --sustained-seconds 0 | show "$w $(basename $l .so)"
  done
done
for w in c3 c5; do VXRT_FIELD=nib VXRT_LIB=$PWD/raytrace_b200/libvxrt_t_minb16.so timeout 300 python bench.py --workload $w --steps 24 --no-cpu-baseline --sustained-seconds 0 | show "$w minb16+nib"; done
VXRT_LIB=$PWD/raytrace_b200/libvxrt_t_minb16.so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1
