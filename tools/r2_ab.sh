#!/bin/bash
# A/B on one GPU: the re-scheduled block against ptxas's own schedule (same source), parity first.
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_havac.py -x -q -m gpu > gpurun_out/ab_parity.log 2>&1; echo "parity rc=$?"; tail -4 gpurun_out/ab_parity.log
for v in "" "$@"; do
  if [ -z "$v" ]; then unset HAVAC_CUDA_LIB; else export HAVAC_CUDA_LIB=tools/variants/libhavac_cuda_$v.so; fi
  echo "== variant ${v:-main}"
  timeout 200 python tools/quick_bench.py 1e8 35804 4 2>&1 | tail -2
  timeout 200 python tools/quick_bench.py 2e8 200 3 2>&1 | tail -1
  timeout 200 python tools/quick_bench.py 2e8 72 3 2>&1 | tail -1
done 2>&1 | tee gpurun_out/ab_variants.log
