#!/bin/bash
# timing experiments: rebuild with -DPINQ_EXP=<n> on the GPU box and print the per-kernel times
for e in "$@"; do
  PINQ_NVCC_EXTRA="-DPINQ_EXP=$e" python -m pin_slam_b200.build --force > /dev/null 2>&1
  PINQ_NVCC_EXTRA="-DPINQ_EXP=$e" python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('EXP $e', 'step', round(d['ms_per_step'],3), 'fwd', round(d['roofline']['ms_fwd_kernel'],3), 'bwd', round(d['roofline']['ms_bwd_kernel'],3))"
done
