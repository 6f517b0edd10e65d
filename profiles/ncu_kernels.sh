#!/usr/bin/env bash
# ncu --set full capture of selected kernels of one bench step.  Usage: bash profiles/ncu_kernels.sh TAG 'regex' SKIP COUNT
set -uo pipefail
TAG="$1"; RE="$2"; SKIP="${3:-0}"; CNT="${4:-8}"
OUT=gpurun_out/$TAG
mkdir -p "$OUT"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > "$OUT/plain.log" 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$RE" -s "$SKIP" -c "$CNT" -f -o "$OUT/prof" $CMD > "$OUT/ncu.log" 2>&1
echo "ncu rc=$?"; tail -3 "$OUT/ncu.log"
