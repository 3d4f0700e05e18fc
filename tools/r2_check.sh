#!/bin/bash
# Round-2 check on a 1-GPU box: GPU suite, smoke, bench (N=1), then configs[3] and configs[4] at full size.
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 120 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 400 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench_n1.json
if [ "$1" == "full" ]; then
  for cfg in c4 c5:0.02 c5:0.05 c5:0.1; do
    timeout 400 python tests/run_config_full.py --config $cfg > gpurun_out/full_${cfg}.log 2>&1; echo "$cfg rc=$?"; tail -c 600 gpurun_out/full_${cfg}.log
  done
fi
