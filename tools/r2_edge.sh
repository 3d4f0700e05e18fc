#!/bin/bash
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
for k in f16x2 f16x2-streaming; do HAVAC_KERNEL=$k python tools/quick_bench.py 1e6 200 3 2>&1 | tail -1; done
python tools/quick_bench.py 1e8 35804 4 2>&1 | tail -2
python tools/quick_bench.py 1e9 72 3 2>&1 | tail -1
python tools/quick_bench.py 1e9 200 3 2>&1 | tail -1
python tools/quick_bench.py 4e8 5000 3 2>&1 | tail -1
ncu --metrics gpu__time_duration.sum,sm__cycles_elapsed.max,sm__cycles_active.avg,sm__cycles_active.max,sm__cycles_active.min --clock-control none -k regex:ssv -s 1 -c 1 python tools/quick_bench.py 1e8 35804 2 2>&1 | grep -E "ssv_|duration|cycles" | tail -6
