#!/bin/bash
# hit tests, then bench.py twice without the strong leg (is a slow e2e step a recurring thing?)
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_hits.py tests/test_zz_no_alloc.py tests/test_gpu_havac.py -x -q -m gpu 2>&1 | tail -3
for i in 1 2; do
  timeout 300 python bench.py --no-strong > gpurun_out/bench_try$i.json 2> gpurun_out/bench_try$i.err; echo "bench $i rc=$?"
  grep "e2e wall" gpurun_out/bench_try$i.err | cut -c1-600
  python -c "
import json; d=json.load(open('gpurun_out/bench_try$i.json')); print(d['value'], d['e2e']['value'], d['e2e']['phases_ms'])"
done
