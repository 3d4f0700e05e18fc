#!/bin/bash
# The no-allocation entry point's test alone in a fresh process, then the whole GPU suite (which ends with it).
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/pytest_noalloc_alone.log
tail -4 gpurun_out/pytest_noalloc_alone.log
timeout 1200 python -m pytest tests -x -q -m gpu --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -16 gpurun_out/pytest_gpu.log
