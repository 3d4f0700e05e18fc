#!/bin/bash
# 2-GPU box: the tests a 1-GPU box skips, then bench.py at N=2 (weak leg, e2e gather on rank 0, strong leg)
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu -k "all_gpus or cuda_all or two_gpus or multi" > gpurun_out/pytest_gpu_n2.log 2>&1; tail -3 gpurun_out/pytest_gpu_n2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/bench_n2.json; grep -E "e2e wall|Error|error" gpurun_out/bench_n2.err | tail -5
