#!/usr/bin/env bash
# ncu launch list (gpu__time_duration per launch) of one bench step.  Usage: bash profiles/ncu_launches.sh TAG [SKIP] [COUNT]
set -uo pipefail
TAG="$1"; SKIP="${2:-0}"; CNT="${3:-40}"
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > "$OUT/plain.log" 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s "$SKIP" -c "$CNT" --csv --log-file "$OUT/launches.csv" $CMD > "$OUT/ncu_launches.log" 2>&1
echo "ncu rc=$?"
