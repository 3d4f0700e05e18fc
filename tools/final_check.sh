#!/bin/bash
# final round-1 verification: GPU suite, bench lines, ncu launch list and full capture of the final kernel
mkdir -p gpurun_out
( time timeout 480 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
timeout 240 python bench.py > gpurun_out/bench_v8.json 2> gpurun_out/bench_v8.err; echo "bench rc=$?" >> gpurun_out/bench_v8.err
timeout 240 python bench.py --impl reference > gpurun_out/bench_ref_v8.json 2> gpurun_out/bench_ref_v8.err
timeout 120 bash tools/sb_sweep.sh > gpurun_out/sb_sweep_v8.log 2>&1
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v6.csv python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_launch.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:ssv_f16x2 -s 1 -c 1 -f -o gpurun_out/full_v6 python tools/quick_bench.py 1e8 35804 2 > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_v8.json
