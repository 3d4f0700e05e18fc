#!/bin/bash
# Development aid: throughput against hit density on the bench shape (consensus scores from 14 are the benchmark's;
# each step of 2 multiplies the hits by roughly 4-10). 100 Mbp and three runs so that the clocks are up; the last run
# counts. Set HAVAC_CUDA_LIB to compare a variant.
for pos in 14 18 20 22; do
  echo "== consensus scores from $pos"
  timeout 120 python tools/quick_bench.py 1e8 35804 3 $pos 2>&1 | tail -1
done
