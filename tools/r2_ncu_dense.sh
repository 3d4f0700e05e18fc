#!/bin/bash
# ncu --set full of the streaming kernel on the bench shape at the given hit densities (consensus scores from $pos),
# with the SASS source page turned into a per-segment stall table (tools/src_regions.py).
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
for pos in "$@"; do
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:ssv_f16x2 -s 1 -c 1 -f -o gpurun_out/ncu_pos$pos python tools/quick_bench.py 4e7 35804 2 $pos > gpurun_out/ncu_pos$pos.log 2>&1; echo "pos $pos rc=$?"
  ncu -i gpurun_out/ncu_pos$pos.ncu-rep --page source --csv --print-source sass 2>/dev/null > gpurun_out/src_pos$pos.csv
  ncu -i gpurun_out/ncu_pos$pos.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/ncu_pos$pos.txt
  python tools/src_regions.py gpurun_out/src_pos$pos.csv --min 0.3 | tee gpurun_out/regions_pos$pos.txt
  grep -E "sm__inst_executed.avg.per_cycle_active|issue_inst0.avg|gpu__time_duration.sum|smsp__thread_inst_executed_per_inst_executed.ratio" gpurun_out/ncu_pos$pos.txt
done
