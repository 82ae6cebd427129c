#!/bin/bash
# round-1f GPU pass: tests, smoke, bench, the other configurations, launch list and a full ncu
# capture of the tensor-core Welch kernel (each profiler pass after the plain command exited 0)
mkdir -p gpurun_out
bash tools/gpu_check.sh
timeout 600 python tools/bench_configs.py > gpurun_out/configs.jsonl 2> gpurun_out/configs.err
cat gpurun_out/configs.jsonl
bash tools/gpu_profile.sh list > gpurun_out/profile_list.log 2>&1
bash tools/gpu_profile.sh full welch_tc > gpurun_out/profile_full.log 2>&1
ls -la gpurun_out | tail -8
