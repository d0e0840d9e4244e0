#!/bin/bash
# one optimisation iteration on the GPU box: GPU parity tests, then the bench step with per-kernel times.  bash scripts/gpu_iter.sh <tag>
TAG=${1:-it}
OUT=gpurun_out/$TAG
mkdir -p "$OUT"
timeout 900 python -m pytest tests -x -q -m gpu > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -3 "$OUT/pytest_gpu.log"
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-extra-legs --slam-frames 0 --driver-frames 0 > "$OUT/bench.json" 2> "$OUT/bench.err"; echo "bench rc=$?"
python - "$OUT/bench.json" <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(d["ms_per_step"], d["roofline"]["frac"], {k:round(v["ms"],4) for k,v in d["roofline"]["kernels"].items()})
P
