#!/bin/bash
# Round-2 evidence of the shipped kernels on one box: GPU suite, bench line, reference arm, hit-density and rows
# sweeps, the user program, configs[3] and configs[4] (p = 0.1) at full size, then -- after those exited -- the ncu
# launch list of bench.py and one full capture of the streaming kernel with its per-segment stall table.
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 400 python bench.py > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err; echo "bench rc=$?"
timeout 240 python bench.py --impl reference > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; echo "ref rc=$?"
timeout 300 bash tools/sb_sweep.sh > gpurun_out/r2_sb_sweep.log 2>&1; cat gpurun_out/r2_sb_sweep.log
timeout 600 bash tools/rows_sweep.sh > gpurun_out/r2_rows_sweep.log 2>&1; tail -12 gpurun_out/r2_rows_sweep.log
rm -f gpurun_out/user_program.jsonl; timeout 600 python examples/run_benchmark.py > gpurun_out/r2_user_program.log 2>&1; echo "user program rc=$?"
for cfg in c5:0.1 c4; do
  timeout 400 python tests/run_config_full.py --config $cfg > gpurun_out/full_${cfg}.log 2>&1; echo "$cfg rc=$?"; tail -c 400 gpurun_out/full_${cfg}.log
done
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-strong > gpurun_out/r2_ncu_launch.log 2>&1; echo "launch list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:ssv_f16x2 -s 1 -c 1 -f -o gpurun_out/r2_full python tools/quick_bench.py 1e8 35804 2 > gpurun_out/r2_ncu_full.log 2>&1; echo "full rc=$?"
ncu -i gpurun_out/r2_full.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/r2_ncu_full_f16x2.txt
ncu -i gpurun_out/r2_full.ncu-rep --page source --csv --print-source sass 2>/dev/null > gpurun_out/r2_src_f16x2.csv
python tools/src_regions.py gpurun_out/r2_src_f16x2.csv --min 0.3 > gpurun_out/r2_regions_f16x2.txt; cat gpurun_out/r2_regions_f16x2.txt
grep -E "sm__inst_executed.avg.per_cycle_active|issue_inst0.avg|registers_per_thread|gpu__time_duration.sum|dram__bytes_(read|write).sum \[|sm__cycles_active.avg \[" gpurun_out/r2_ncu_full_f16x2.txt
