#!/bin/bash
# Throughput against the number of table rows, 1 Gbp so that the launch lasts milliseconds and the clocks are up.
# One JSON line per case into gpurun_out/rows_sweep.jsonl: {"rows", "kernel", "gcups", "ms"}.
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
out=gpurun_out/rows_sweep.jsonl; : > $out
for kernel in f16x2 f16x2-streaming; do
  for rows in 64 72 200 512 513 1000 2040; do
    if [ $kernel == f16x2-streaming ] && [ $rows -gt 512 ]; then continue; fi
    line=$(HAVAC_KERNEL=$kernel timeout 120 python tools/quick_bench.py 1e9 $rows 3 2>&1 | tail -1)
    echo "$kernel rows=$rows: $line"
    ms=$(echo "$line" | sed -n 's/.*device=\([0-9.]*\) ms.*/\1/p'); g=$(echo "$line" | sed -n 's/.*GCUPS=\([0-9.]*\).*/\1/p')
    echo "{\"rows\": $rows, \"kernel\": \"$kernel\", \"symbols\": 1000000000, \"device_ms\": ${ms:-null}, \"gcups\": ${g:-null}}" >> $out
  done
done
