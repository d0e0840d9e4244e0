#!/bin/bash
# ncu launch list + full capture of the hot kernels.  bash scripts/gpu_prof.sh <tag> [kernel-regex] [skip]
TAG=${1:-p}
KREGEX=${2:-'k_search|k_decode|k_bwd|k_brick|k_sort|k_key'}
SKIP=${3:-12}
OUT=gpurun_out/$TAG
mkdir -p "$OUT"
SMALL="bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 python $SMALL > "$OUT/plain.log" 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file "$OUT/launches.csv" python $SMALL > "$OUT/ncu_launches.log" 2>&1
echo "launches rc=$?"
timeout 600 python $SMALL > "$OUT/plain2.log" 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k "regex:$KREGEX" -s $SKIP -c 4 -f -o "$OUT/prof" python $SMALL > "$OUT/ncu_full.log" 2>&1
echo "ncu full rc=$?"
cat "$OUT/plain.log" | tail -2
