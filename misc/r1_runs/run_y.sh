mkdir -p gpurun_out
python raytrace_b200/build.py > /dev/null
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "frame_matches_oracle or peer_memory or edits_rebuild or accumulation or slice_upload" > gpurun_out/r1_y_memcheck.log 2>&1; echo "memcheck rc=$?"; tail -4 gpurun_out/r1_y_memcheck.log
timeout 600 compute-sanitizer --tool racecheck --error-exitcode 7 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "frame_matches_oracle and c1" > gpurun_out/r1_y_racecheck.log 2>&1; echo "racecheck rc=$?"; tail -4 gpurun_out/r1_y_racecheck.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r1_y_c5_launches.csv python bench.py --workload c5 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1_y_c5_ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv
rows=[r for r in csv.reader(open("gpurun_out/r1_y_c5_launches.csv")) if len(r)>5]
hdr=rows[0]; ik=hdr.index("Kernel Name"); iv=hdr.index("Metric Value")
for r in rows[-24:]: print(r[ik][:60], r[iv])
PY
