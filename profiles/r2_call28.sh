#!/usr/bin/env bash
# GPU call 28: ncu --set full of the split-bf16 (default) decode kernel; decode timings of the final tree.
set -uo pipefail
OUT=gpurun_out/r2ab; mkdir -p "$OUT"
CMD="python profiles/tools/decoder_perf.py bf16x2 f16"
$CMD > "$OUT/decoder_perf.log" 2>&1; echo "perf rc=$?"; tail -8 "$OUT/decoder_perf.log"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'decode|fold|pack' -c 60 --csv --log-file "$OUT/decode_launches.csv" $CMD > "$OUT/ncu_launches.log" 2>&1; echo "launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'decode_tc_kernel' -s 1 -c 2 -f -o "$OUT/prof_decode" $CMD > "$OUT/ncu_full.log" 2>&1; echo "ncu full rc=$?"
ncu -i "$OUT/prof_decode.ncu-rep" --page raw --csv > "$OUT/prof_decode_raw.csv" 2>/dev/null; echo "raw rc=$?"
ls -la "$OUT"
