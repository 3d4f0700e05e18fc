#!/bin/bash
# 8-GPU box: bench.py at N=8 only (weak leg, e2e with the gather on rank 0, strong leg)
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
N=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_n$N.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['strong']['value'], d['strong']['ms_per_step'], d['strong']['hits'], d['strong']['hits_on_rank0_after_gather'], d.get('device_open_s'))"
grep -E "Error|error" gpurun_out/bench_n$N.err | tail -3
