#!/bin/bash
# Repeats the parity tests that lean on the shared-memory protocols of the streaming kernels (chunk hand-over, hit
# stage, abort) to flush out anything timing-dependent: N rounds of the selected tests, then M rounds of C2 at full
# size (fast == plain on the whole matrix).
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
N=${1:-15}; M=${2:-4}; fail=0
for i in $(seq 1 $N); do
  timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_havac.py -x -q -m gpu -k "uneven or stage_overflow or long_model or hit_dense or abort or extreme or tail_length or short_table" > gpurun_out/stress_$i.log 2>&1 || { fail=$((fail+1)); echo "round $i FAILED"; tail -20 gpurun_out/stress_$i.log; }
done
for i in $(seq 1 $M); do
  timeout 300 python -m pytest tests/test_gpu_havac.py -x -q -m gpu -k "full_size_properties_c2" > gpurun_out/stress_c2_$i.log 2>&1 || { fail=$((fail+1)); echo "c2 round $i FAILED"; tail -20 gpurun_out/stress_c2_$i.log; }
done
tail -1 gpurun_out/stress_1.log; tail -1 gpurun_out/stress_c2_1.log
echo "stress: $N + $M rounds, $fail failed"
