#!/bin/bash
# One profiler pass per call, each after the same command exited 0 without the profiler:
#   tools/gpu_profile.sh list                 launch list (gpu__time_duration.sum per launch)
#   tools/gpu_profile.sh full <kernel regex>  one --set full capture of a kernel (3rd matching launch)
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu"
MODE=${1:-list}
KERN=${2:-welch_tc}
$CMD > gpurun_out/prof_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/prof_plain.log; exit 1; }
if [ "$MODE" = list ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:welch|detrend|stft|bigfft|csm|tc_' -c 120 --csv \
      --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
  tail -n 2 gpurun_out/ncu_list.log
else
  ncu --set full --clock-control none --import-source on -k regex:$KERN -s 2 -c 1 -f -o gpurun_out/prof $CMD \
      > gpurun_out/ncu_full.log 2>&1
  tail -n 2 gpurun_out/ncu_full.log
fi
ls -la gpurun_out | tail -8
