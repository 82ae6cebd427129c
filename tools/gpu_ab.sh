#!/bin/bash
# A/B of two builds of the library on the same box: ab/lib_base.so against ab/lib_variant.so
for round in 1 2 3; do
  for v in base variant; do
    cp ab/lib_$v.so pyfft_b200/libpyfft_b200.so
    python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', 'kernel_ms %.4f step_ms %.4f' % (d['roofline']['kernel_ms'], d['ms_per_step']))"
  done
done
