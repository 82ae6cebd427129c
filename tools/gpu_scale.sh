#!/bin/bash
# weak-scaling line of bench.py at NG GPUs of one box (as the driver launches it)
mkdir -p gpurun_out
NG=${NG:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus $NG --steps 20 --warmup 5 2> gpurun_out/bench_scale_n$NG.err | tail -1 > gpurun_out/bench_scale_n$NG.json
python -c "
import json; d=json.loads(open('gpurun_out/bench_scale_n$NG.json').read()); print(d['n_gpus'], 'GPUs: ms_per_step', d['ms_per_step'], 'value', d['value'], d['config'].get('collective'))"
