#!/bin/bash
# dev: the ncu evidence of a round in one GPU call (run under gpurun from the repository root; results in gpurun_out/).
#   launch list of the bench command, light metrics of the C2 launches, --set full captures at 1080p x 16 spp
# Summaries are made afterwards, without a GPU: tools/ncu_summary.py / tools/ncu_blocks.py on the .ncu-rep files.
set -x
O=gpurun_out
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,sm__issue_active.avg.pct_of_peak_sustained_elapsed,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__thread_inst_executed_per_inst_executed.ratio
python bench.py --steps 2 --warmup 3 --no-cpu --no-extras > $O/bench_plain.json 2> $O/bench_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-extras > $O/ncu_bench.log 2>&1
ncu --metrics $M --clock-control none -k regex:'primary_scan|render_scanu' --csv --log-file $O/c2_brute_metrics.csv python bench.py --steps 1 --warmup 3 --no-cpu --no-extras > $O/ncu_c2.log 2>&1
F="--set full --clock-control none --import-source on -f"
ncu $F -k regex:'primary_scan|render_scanu' -o $O/brute_s16 python tools/prof_run.py 1920 1080 16 1 0 reference 1 > $O/ncu_brute_s16.log 2>&1
ncu $F -k regex:render_lbvh_flat -o $O/lbvh_485_s16 python tools/prof_run.py 1920 1080 16 2 0 reference 1 > $O/ncu_lbvh_485.log 2>&1
ncu $F -k regex:render_lbvh_flat -o $O/lbvh_100k_s16 python tools/prof_run.py 1920 1080 16 2 0 158 1 > $O/ncu_lbvh_100k.log 2>&1
ls -la $O
