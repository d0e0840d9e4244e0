#!/bin/bash
# One gpurun call: smoke, GPU parity tests, the bench (both arms), the ncu launch list and one full capture of the
# hot kernels.  Usage (from the repo root, on the GPU box):  bash scripts/gpu_round.sh <tag> [ncu-kernel-regex]
set -u
TAG=${1:-rX}
KREGEX=${2:-'k_search_brick|k_decode_tc|k_bwd_tc'}
OUT=gpurun_out/$TAG
mkdir -p "$OUT"
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > "$OUT/smoke.log" 2>&1; echo "smoke rc=$?"
timeout 1500 python -m pytest tests -x -q -m gpu > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -3 "$OUT/pytest_gpu.log"
timeout 1200 python bench.py --steps 20 --warmup 5 > "$OUT/bench.json" 2> "$OUT/bench.err"; echo "bench rc=$?"; head -c 1500 "$OUT/bench.json"; echo
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > "$OUT/bench_reference.json" 2> "$OUT/bench_reference.err"; echo "ref rc=$?"; head -c 800 "$OUT/bench_reference.json"; echo
SMALL="bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --slam-frames 0 --no-extra-legs --driver-frames 0"
timeout 600 python $SMALL > "$OUT/plain.log" 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file "$OUT/launches.csv" python $SMALL > "$OUT/ncu_launches.log" 2>&1
echo "launches rc=$?"
timeout 600 python $SMALL > "$OUT/plain2.log" 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k "regex:$KREGEX" -s 9 -c 3 -f -o "$OUT/prof" python $SMALL > "$OUT/ncu_full.log" 2>&1
echo "ncu full rc=$?"
ls -la "$OUT"
