#!/bin/bash
M=gpu__time_duration.sum,smsp__inst_executed.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed
for l in libvxrt libvxrt_b8 libvxrt_b6; do
  VXRT_LIB=$PWD/raytrace_b200/$l.so timeout 300 python bench.py --workload c5 --steps 32 --no-cpu-baseline --sustained-seconds 0 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c5 $l ms/step', round(d['ms_per_step'],4), 'fps', round(d['frames_per_s']))"
  VXRT_LIB=$PWD/raytrace_b200/$l.so timeout 300 ncu --metrics $M --clock-control none -k regex:k_build_minefield -s 4 -c 1 --csv --log-file gpurun_out/r2_z_build_$l.csv python bench.py --workload c5 --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > /dev/null 2>&1
  grep -v "^==" gpurun_out/r2_z_build_$l.csv | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]
print('   ncu', {r[h.index('Metric Name')].split('.')[0][-22:]: r[h.index('Metric Value')] for r in rows[1:]})"
done
