#!/bin/bash
# ncu launch list + one full capture of the dominant kernel (each after a plain run exited 0)
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu"
KERN=${1:-welch_pair}
$CMD > gpurun_out/prof_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:welch|detrend|stft|bigfft|csm' -c 60 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
$CMD > gpurun_out/prof_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$KERN -s 2 -c 1 -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
tail -n 2 gpurun_out/ncu_list.log gpurun_out/ncu_full.log; ls -la gpurun_out
