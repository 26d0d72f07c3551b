mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
bash misc/sweep.sh s_packed "" s_scalar "-DVXRT_MARCH_PACKED=0" s_packed_mb10 "-DVXRT_TRACE_MINB=10" 
