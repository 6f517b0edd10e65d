#!/usr/bin/env bash
# launch list (device time of every launch of one step).  Usage: bash profiles/ncu_launches.sh TAG [skip] [count]
set -uo pipefail
TAG="$1"; SKIP="${2:-42}"; CNT="${3:-28}"
OUT=gpurun_out/$TAG; mkdir -p "$OUT"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > "$OUT/plain.log" 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'preprocess_fwd_kernel|preprocess_bwd_kernel|scan_tiles_kernel|duplicate_kernel|sort_hist_kernel|onesweep_kernel|tile_ranges_kernel|render_fwd_kernel|render_bwd_kernel|unpack' -s "$SKIP" -c "$CNT" --csv --log-file "$OUT/launches.csv" $CMD > "$OUT/ncu.log" 2>&1
echo "rc=$?"
python - <<PY
import csv
rows=[r for r in csv.reader(open("$OUT/launches.csv")) if len(r)>5]
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value')
for r in rows[1:]: print("%-60s %10.1f us" % (r[ki].split('(')[0][-60:], float(r[vi].replace(',',''))/1e3))
PY
