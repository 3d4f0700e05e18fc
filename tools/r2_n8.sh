#!/bin/bash
# 8-GPU box: configs[3] and configs[4] at full size sharded over 8 ranks, then bench.py at N=8
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
N=${1:-8}
for cfg in c4 c5:0.02 c5:0.05 c5:0.1; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 tests/run_config_full.py --config $cfg > gpurun_out/full${N}_${cfg}.log 2>&1; echo "$cfg rc=$?"; tail -c 400 gpurun_out/full${N}_${cfg}.log
done
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench_n$N.json
