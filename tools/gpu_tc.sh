#!/bin/bash
# tensor-core Welch kernel: its tests (under a time limit: a broken hand-shake must not hang the box), then the bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_welch_tc.py tests/test_gpu_fullsize.py -m gpu -x -q > gpurun_out/pytest_tc.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_tc.log
tail -6 gpurun_out/pytest_tc.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench_tc.json 2> gpurun_out/bench_tc.err
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/bench_tc.json").read().strip().splitlines()[-1])
    print("kernel_ms", d["roofline"]["kernel_ms"], "frac", d["roofline"]["frac"], "step_ms", d["ms_per_step"], "err", d["roofline"].get("vs_two_pass_cuda_core_max_rel"))
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/bench_tc.err").read()[-2000:])
PY
