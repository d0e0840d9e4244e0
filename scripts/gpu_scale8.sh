#!/bin/bash
# On an 8-GPU box: bench at N=8 (peer, nccl) and N=4 (peer), configs 4/5 at N=8.  bash scripts/gpu_scale8.sh <tag>
TAG=${1:-s8}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
EXCH="peer nccl" bash scripts/gpu_multi.sh $TAG 8 2>&1 | grep -v "^\*\*\*\|OMP_NUM"
EXCH="peer" bash scripts/gpu_multi.sh $TAG 4 2>&1 | grep -v "^\*\*\*\|OMP_NUM"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 scripts/bench_configs.py > "$OUT/configs_n8.jsonl" 2> "$OUT/configs_n8.err"; echo "configs rc=$?"
cat "$OUT/configs_n8.jsonl"; grep -v "^\*\*\*\|OMP_NUM" "$OUT/configs_n8.err" | tail -5
