#!/bin/bash
# round-1e GPU pass: tests, smoke, bench, then ncu captures of the STFT / long-FFT / CSM kernels
mkdir -p gpurun_out
bash tools/gpu_check.sh
timeout 600 python tools/bench_configs.py > gpurun_out/configs.jsonl 2> gpurun_out/configs.err
cat gpurun_out/configs.jsonl
export PYFFT_C3_CH=8 PYFFT_C4_CH=4 PYFFT_C5_LOG2N=22
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stft_kernel -s 2 -c 1 -f -o gpurun_out/prof_stft python tools/bench_configs.py c3 > gpurun_out/ncu_stft.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:bigfft_pass -s 6 -c 3 -f -o gpurun_out/prof_bigfft python tools/bench_configs.py c4 > gpurun_out/ncu_bigfft.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:csm_product -s 2 -c 1 -f -o gpurun_out/prof_csm python tools/bench_configs.py c5 > gpurun_out/ncu_csm.log 2>&1
ls -la gpurun_out | tail
