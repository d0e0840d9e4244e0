#!/bin/bash
# N-GPU bench, peer-memory exchange vs NCCL all-reduce.  bash scripts/gpu_multi.sh <tag> <N>
TAG=${1:-m}; N=${2:-2}
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3"
for ex in ${EXCH:-peer nccl}; do
  PINQ_EXCHANGE=$ex timeout 600 $RUN > "$OUT/bench_n${N}_$ex.json" 2> "$OUT/bench_n${N}_$ex.err"; echo "$ex rc=$?"
  python - "$OUT/bench_n${N}_$ex.json" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(d["n_gpus"], "GPUs", round(d["value"] / 1e9, 3), "G q/s", round(d["ms_per_step"], 3), "ms/step e2e", round(d["e2e"]["value"] / 1e9, 3), "exchange_ms", d.get("exchange_ms"), d.get("ms_per_step_by_rank"), "no-exchange", d.get("ms_per_step_without_exchange"), d["config"]["parallelism"][-60:], d.get("peer_exchange_error"))
except Exception as e:
    print("no line:", e)
PY
  tail -3 "$OUT/bench_n${N}_$ex.err"
done
