#!/bin/bash
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
for rows in 72 200; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:ssv_f16x2 -s 1 -c 1 -f -o gpurun_out/ncu_rows$rows python tools/quick_bench.py 2e8 $rows 2 > gpurun_out/ncu_rows$rows.log 2>&1
  ncu -i gpurun_out/ncu_rows$rows.ncu-rep --page source --csv --print-source sass 2>/dev/null > gpurun_out/src_rows$rows.csv
  ncu -i gpurun_out/ncu_rows$rows.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/ncu_rows$rows.txt
  grep -E "sm__inst_executed.avg.per_cycle_active|issue_inst0.avg|gpu__time_duration.sum" gpurun_out/ncu_rows$rows.txt
done
