#!/bin/bash
# round-2 final evidence: launch list of the bench command, --set full of the dominant kernel, of the post kernel
# and of one long-transform pass (each after the same command exited 0 without the profiler)
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --no-graph"
$CMD > gpurun_out/prof_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/prof_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:welch|detrend|stft|bigfft|csm|tc_' -c 160 --csv \
    --log-file gpurun_out/r02g_launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:welch_tc_kernel -s 2 -c 1 -f -o gpurun_out/r02g_welch_tc $CMD \
    > gpurun_out/ncu_full1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:tc_post_kernel -s 2 -c 1 -f -o gpurun_out/r02g_tc_post $CMD \
    > gpurun_out/ncu_full2.log 2>&1
PYFFT_BENCH_ONCE=1 python tools/bench_configs.py c4 > gpurun_out/c4_plain.log 2>&1 && \
PYFFT_BENCH_ONCE=1 ncu --set full --clock-control none --import-source on -k regex:bigfft_pass -s 7 -c 1 -f -o gpurun_out/r02g_bigfft \
    python tools/bench_configs.py c4 > gpurun_out/ncu_full3.log 2>&1
tail -n 2 gpurun_out/ncu_list.log gpurun_out/ncu_full1.log gpurun_out/ncu_full2.log gpurun_out/ncu_full3.log
ls -la gpurun_out | tail -8
