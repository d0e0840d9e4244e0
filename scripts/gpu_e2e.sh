#!/bin/bash
# e2e leg of bench.py for several chunk counts.  bash scripts/gpu_e2e.sh 1 2 4 8
for c in "$@"; do
  PINQ_E2E_CHUNKS=$c python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('chunks $c', 'resident', round(d['value']/1e9,3), 'e2e', round(d['e2e']['value']/1e9,3))"
done
