mkdir -p gpurun_out
for wl in c3 c4; do
ncu --set full --clock-control none -k regex:k_trace_fast -s 3 -c 1 -f -o gpurun_out/r1_z_$wl python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1_z_$wl.log 2>&1; echo "ncu $wl rc=$?"
done
