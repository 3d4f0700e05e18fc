#!/bin/bash
# partial last block: parity, then throughput against the previous build (tools/variants/libhavac_cuda_base.so)
mkdir -p gpurun_out
( timeout 100 python -m pytest tests/test_gpu_parity.py tests/test_gpu_havac.py -m gpu -x -q ) > gpurun_out/tail_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tail_pytest.log
tail -3 gpurun_out/tail_pytest.log
for lib in "" tools/variants/libhavac_cuda_base.so; do
  if [ -z "$lib" ]; then unset HAVAC_CUDA_LIB; else export HAVAC_CUDA_LIB=$lib; fi
  echo "== lib ${lib:-new}"
  for rows in 72 200 1000 2040; do echo -n "rows $rows: "; timeout 30 python tools/quick_bench.py 3e8 $rows 3 2>&1 | tail -1; done
  echo -n "rows 35804: "; timeout 30 python tools/quick_bench.py 1e8 35804 3 2>&1 | tail -1
done 2>&1 | tee gpurun_out/tail_sweep.log
