#!/bin/bash
# timing experiments: rebuild with -DPINQ_EXP=<n> on the GPU box and print the per-kernel times
for e in "$@"; do
  PINQ_NVCC_EXTRA="-DPINQ_EXP=$e" python -m pin_slam_b200.build --force > /dev/null 2>&1
  PINQ_NVCC_EXTRA="-DPINQ_EXP=$e" python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('EXP $e', 'step', round(d['ms_per_step'],3), {k: round(v,3) for k,v in d['roofline']['kernel_ms'].items()})"
done
