#!/bin/bash
# Round-2 evidence of the shipped (re-scheduled) kernel on one box: bench line, reference arm, hit-density and rows
# sweeps, then -- after those exited -- the ncu launch list of bench.py and one full capture of the fast kernel.
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
timeout 400 python bench.py > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err; echo "bench rc=$?"
timeout 240 python bench.py --impl reference > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; echo "ref rc=$?"
timeout 200 bash tools/sb_sweep.sh > gpurun_out/r2_sb_sweep.log 2>&1; cat gpurun_out/r2_sb_sweep.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --steps 2 --warmup 1 > gpurun_out/r2_ncu_launch.log 2>&1; echo "launch list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:ssv_f16x2 -s 1 -c 1 -f -o gpurun_out/r2_full python tools/quick_bench.py 1e8 35804 2 > gpurun_out/r2_ncu_full.log 2>&1; echo "full rc=$?"
ncu -i gpurun_out/r2_full.ncu-rep --page raw --csv 2>/dev/null | python tools/ncu_summary.py > gpurun_out/r2_ncu_full_f16x2.txt
ncu -i gpurun_out/r2_full.ncu-rep --page source --csv --print-source sass 2>/dev/null > gpurun_out/r2_src_f16x2.csv
grep -E "issue_stalled.*per_issue_active|sm__inst_executed.avg.per_cycle_active|pipe_alu_cycles_active.avg.pct_of_peak_sustained_active|issue_inst0|registers_per_thread|gpu__time_duration.sum|dram__bytes_(read|write).sum " gpurun_out/r2_ncu_full_f16x2.txt | head -40
ls -la gpurun_out | head -40
