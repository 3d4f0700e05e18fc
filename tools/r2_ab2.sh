#!/bin/bash
# hit-density sweep + C2 shape for the main library and the named variants; optional ncu of main
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
for v in "" "$@"; do
  if [ -z "$v" ]; then unset HAVAC_CUDA_LIB; else export HAVAC_CUDA_LIB=tools/variants/libhavac_cuda_$v.so; fi
  echo "== variant ${v:-main}"
  bash tools/sb_sweep.sh 2>&1 | grep -E "==|run"
done 2>&1 | tee gpurun_out/ab2_variants.log
unset HAVAC_CUDA_LIB
if [ -n "$NCU_MAIN" ]; then
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:ssv_f16x2 -s 1 -c 1 -f -o gpurun_out/ncu_main2 python tools/quick_bench.py 4e7 35804 2 > gpurun_out/ncu_main2.log 2>&1
  ncu -i gpurun_out/ncu_main2.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/ncu_main2.txt
  ncu -i gpurun_out/ncu_main2.ncu-rep --page source --csv --print-source sass 2>/dev/null > gpurun_out/src_main2.csv
  grep -E "sm__inst_executed.avg.per_cycle_active|issue_inst0.avg" gpurun_out/ncu_main2.txt
fi
