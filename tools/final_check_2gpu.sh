#!/bin/bash
# final round-1 verification on two GPUs: the multi-GPU tests the 1-GPU box skips, and the N=2 bench line
mkdir -p gpurun_out
( time timeout 300 python -m pytest tests -m gpu -q -k "multi or two_gpu or peer or cuda_all or all_gpus or shard" ) > gpurun_out/pytest_gpu_n2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_n2.log
( time timeout 300 python -m pytest tests/test_gpu_hits.py tests/test_gpu_ingest.py tests/test_gpu_parity.py tests/test_strand.py -m gpu -q -rs ) > gpurun_out/pytest_gpu_n2_files.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_n2_files.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2_v8.json 2> gpurun_out/bench_n2_v8.err; echo "bench rc=$?" >> gpurun_out/bench_n2_v8.err
tail -4 gpurun_out/pytest_gpu_n2_files.log; cat gpurun_out/bench_n2_v8.json
