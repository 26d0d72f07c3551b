for v in "" "-DVXRT_MARCH_PACKED=0" "-DVXRT_DN_UNROLL=1" "-DVXRT_MARCH_PACKED=0 -DVXRT_DN_UNROLL=1"; do
  VXRT_NVCC_EXTRA="$v" python raytrace_b200/build.py --force > /dev/null 2>&1
  echo "== variant [$v]"
  timeout 100 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "peer_memory" 2>&1 | grep -E "passed|failed|differ" | head -3
done
