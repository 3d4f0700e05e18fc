#!/bin/bash
# Development aid: throughput against hit density (consensus scores from 14 are the benchmark's;
# each step of 2 multiplies the hits by roughly 10). Set HAVAC_CUDA_LIB to compare a variant.
for pos in 14 18 20 22; do
  echo "== consensus scores from $pos"
  timeout 60 python tools/quick_bench.py 2e7 35804 2 $pos 2>&1 | tail -1
done
python tools/quick_bench.py 1e8 35804 3 2>&1 | tail -2
