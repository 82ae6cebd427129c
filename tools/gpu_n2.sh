#!/bin/bash
# 2-GPU pass: peer reduction check, then the bench with the NVLink peer reduction and with NCCL
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NG:-2} --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29512 tools/peer_check.py 2>&1 | grep -E "epoch 5|GPUs,|MISMATCH|Error|error" | tee gpurun_out/peer_check_n${NG:-2}.txt
timeout 600 $TR --master-port 29513 bench.py --gpus ${NG:-2} --steps 20 --warmup 5 --no-cpu --no-e2e 2> gpurun_out/bench_peer.err | tail -1 > gpurun_out/bench_n${NG:-2}_peer.json
PYFFT_PEER_REDUCE=0 timeout 600 $TR --master-port 29514 bench.py --gpus ${NG:-2} --steps 20 --warmup 5 --no-cpu --no-e2e 2> gpurun_out/bench_nccl.err | tail -1 > gpurun_out/bench_n${NG:-2}_nccl.json
python - <<'PY'
import json, os
ng = os.environ.get("NG", "2")
for k in ("peer", "nccl"):
    try:
        d = json.loads(open("gpurun_out/bench_n%s_%s.json" % (ng, k)).read())
        print(k, "ms_per_step", d["ms_per_step"], "value", d["value"], "n_gpus", d["n_gpus"])
    except Exception as e:
        print(k, "failed", e); print(open("gpurun_out/bench_%s.err" % k).read()[-1500:])
PY
