#!/bin/bash
# A/B on one GPU: main library against tools/variants/libhavac_cuda_<name>.so, parity suite first, then the bench
# shape at four hit densities (100 Mbp so that the clocks are up) and two mid-size tables.
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/ab_parity.log 2>&1; echo "parity rc=$?"; tail -4 gpurun_out/ab_parity.log
for v in "" "$@"; do
  if [ -z "$v" ]; then unset HAVAC_CUDA_LIB; else export HAVAC_CUDA_LIB=tools/variants/libhavac_cuda_$v.so; fi
  echo "== variant ${v:-main}"
  for pos in 14 18 20 22; do
    echo "-- consensus scores from $pos"; timeout 200 python tools/quick_bench.py 1e8 35804 3 $pos 2>&1 | tail -2
  done
  echo "-- 1000 rows"; timeout 200 python tools/quick_bench.py 1e9 1000 3 2>&1 | tail -1
  echo "-- 5000 rows"; timeout 200 python tools/quick_bench.py 4e8 5000 3 2>&1 | tail -1
done 2>&1 | tee gpurun_out/ab_variants.log
