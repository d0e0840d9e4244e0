#!/bin/bash
# 8-GPU box: bench at N=8 (peer exchange) and config 4 through the device-side data-parallel loop.  bash scripts/gpu_scale8b.sh <tag>
TAG=${1:-s8b}; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
EXCH="peer" bash scripts/gpu_multi.sh $TAG 8 2>&1 | grep -v "^\*\*\*\|OMP_NUM"
for n in ${CFG4_N:-8 4}; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 scripts/bench_configs.py --skip 5 > "$OUT/config4_n$n.jsonl" 2> "$OUT/config4_n$n.err"; echo "config4 n=$n rc=$?"
cat "$OUT/config4_n$n.jsonl"; grep -v "^\*\*\*\|OMP_NUM" "$OUT/config4_n$n.err" | grep -v "^$" | tail -4
done
