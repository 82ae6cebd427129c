#!/bin/bash
# round-1e GPU pass: tests, smoke, bench, the other configurations, then ncu captures
mkdir -p gpurun_out
bash tools/gpu_check.sh
timeout 600 python tools/bench_configs.py > gpurun_out/configs.jsonl 2> gpurun_out/configs.err
cat gpurun_out/configs.jsonl
bash tools/gpu_profile.sh list > gpurun_out/profile_list.log 2>&1
PYFFT_C3_CH=8 timeout 600 ncu --set full --clock-control none --import-source on -k regex:stft_real_kernel -s 2 -c 1 -f -o gpurun_out/prof_stft_real python tools/bench_configs.py c3 > gpurun_out/ncu_stft_real.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:csm_tile -s 1 -c 1 -f -o gpurun_out/prof_csm_tile python tools/bench_configs.py c5 > gpurun_out/ncu_csm_tile.log 2>&1
ls -la gpurun_out | tail -12
