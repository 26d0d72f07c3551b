#!/bin/bash
# whole GPU suite (wavefront + layout tests included), then the occupancy-field layouts (lin / nib / brick / tex) on C2, C3, C5, C4:
# trace time per variant from bench.py, and one ncu launch of the trace pass per variant with hit rates and DRAM bytes
tag=${1:-r2_s}
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_pytest.txt 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/${tag}_pytest.txt)"
show() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', round(d['kernel_ms_per_step_rank0']['trace'],4), 'ms trace,', round(d['ms_per_step'],4), 'ms/step')"; }
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum
for w in c2 c3 c5 c4; do for f in lin nib brick tex; do
  VXRT_FIELD=$f timeout 300 python bench.py --workload $w --steps 16 --no-cpu-baseline --sustained-seconds 0 > gpurun_out/${tag}_${w}_${f}.json 2>/dev/null; show "$w $f" < gpurun_out/${tag}_${w}_${f}.json
  if [ $f = nib -o $f = brick -o $w = c3 ]; then
  VXRT_FIELD=$f timeout 300 ncu --metrics $M --clock-control none -k regex:'k_trace' -s 3 -c 1 --csv --log-file gpurun_out/${tag}_layout_${w}_$f.csv \
      python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline --sustained-seconds 0 > /dev/null 2>&1
  grep -v "^==" gpurun_out/${tag}_layout_${w}_$f.csv | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]
print('   ncu', {r[h.index('Metric Name')].split('.')[0][-24:]: r[h.index('Metric Value')] for r in rows[1:]})"
  fi
done; done
