#!/usr/bin/env bash
# side configs of BASELINE.json (3: feature-only training, 5: 4K forward sweep, 4: 8 views/step) on one GPU, both arms
set -uo pipefail
TAG="$1"; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
for c in 3 5 4; do
  for impl in ours reference; do
    python bench.py --config $c --impl $impl --steps 10 --warmup 3 --no-cpu-baseline > "$OUT/bench_c${c}_${impl}.json" 2> "$OUT/bench_c${c}_${impl}.err"
    echo "config $c $impl rc=$?"
    python - <<PY
import json
try:
    j=json.load(open("$OUT/bench_c${c}_${impl}.json"))
    print("  %s: %.1f %s  ms/step %.3f  fwd-only %.3f ms  e2e %.1f" % (j["config"]["workload"][:60], j["value"], j["unit"], j["ms_per_step"], j["fwd_only_ms_per_frame"], j["e2e"]["value"]))
    for k,v in j.get("stages",{}).items(): print("     %-16s %8.1f us" % (k, v["ms"]*1e3))
except Exception as e: print("  parse failed", e); print(open("$OUT/bench_c${c}_${impl}.err").read()[-1500:])
PY
  done
done
