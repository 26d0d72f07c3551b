#!/bin/bash
export VXRT_POOL=1
for l in raytrace_b200/libvxrt_p_*.so; do
  n=$(basename $l .so)
  VXRT_LIB=$PWD/$l timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "ray_pool" 2>&1 | tail -1
  for w in c3 c4; do VXRT_LIB=$PWD/$l timeout 300 python bench.py --workload $w --steps 24 --no-cpu-baseline --sustained-seconds 0 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(\"$n $w trace\", round(d[\"kernel_ms_per_step_rank0\"][\"trace\"],4))"; done
done
