#!/bin/bash
# short-table kernel check: parity tests that run it, then the short end of the rows sweep, main library against a variant
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_havac.py tests/test_gpu_hits.py -x -q -m gpu 2>&1 | tail -2
for v in "" "$@"; do
  if [ -z "$v" ]; then unset HAVAC_CUDA_LIB; else export HAVAC_CUDA_LIB=tools/variants/libhavac_cuda_$v.so; fi
  echo "== variant ${v:-main}"
  for rows in 64 72 200 512; do echo -n "rows $rows: "; timeout 120 python tools/quick_bench.py 1e9 $rows 3 2>&1 | tail -1; done
  echo -n "1 Mbp x 200: "; python tools/quick_bench.py 1e6 200 3 2>&1 | tail -1
done
