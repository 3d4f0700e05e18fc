#!/bin/bash
# ncu --set full of the fast kernel for the main library and for the named variants (same workload).
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
for v in "" "$@"; do
  if [ -z "$v" ]; then unset HAVAC_CUDA_LIB; tag=main; else export HAVAC_CUDA_LIB=tools/variants/libhavac_cuda_$v.so; tag=$v; fi
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:ssv_f16x2 -s 1 -c 1 -f -o gpurun_out/ncu_$tag python tools/quick_bench.py 4e7 35804 2 > gpurun_out/ncu_$tag.log 2>&1
  echo "$tag rc=$?"
  ncu -i gpurun_out/ncu_$tag.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/ncu_$tag.txt
  grep -E "issue_stalled.*per_issue_active|sm__inst_executed.avg.per_cycle_active|pipe_alu_cycles_active.avg.pct_of_peak_sustained_active|pipe_fmaheavy|issue_inst0|registers_per_thread|no_instruction" gpurun_out/ncu_$tag.txt | head -40
done
