#!/bin/bash
# exchange experiments on N GPUs.  bash scripts/exp_peer.sh <N> <tag> name=ENV=V,ENV2=V2 ...
N=${1:-2}; TAG=${2:-peer}; shift 2; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
for spec in "$@"; do
  name=${spec%%=*}; envs=${spec#*=}
  env ${envs//,/ } timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-extra-legs --slam-frames 0 --driver-frames 0 \
    > "$OUT/$name.json" 2> "$OUT/$name.err"
  python - "$OUT/$name.json" "$name" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2], "step", round(d["ms_per_step"],4), "no_exchange", round(d.get("ms_per_step_without_exchange") or 0,4), "exchange_alone", round(d.get("exchange_ms") or 0,4), "err", d.get("peer_exchange_error"), d.get("overlap_trace"))
except Exception as e:
    print(sys.argv[2], "FAILED", e)
P
done
