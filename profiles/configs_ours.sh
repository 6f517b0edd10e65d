#!/usr/bin/env bash
# side configs, our arm only (quick A/B of a kernel change across the workloads)
set -uo pipefail
TAG="$1"; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
for c in 2 3 5 4; do
  python bench.py --config $c --steps 10 --warmup 3 --no-cpu-baseline > "$OUT/bench_c${c}_ours.json" 2> "$OUT/bench_c${c}_ours.err"
  python - <<PY
import json
j=json.load(open("$OUT/bench_c${c}_ours.json"))
print("c$c %.1f %s  ms/step %.3f  e2e %.1f  " % (j["value"], j["unit"], j["ms_per_step"], j["e2e"]["value"]) + " ".join("%s %.0f" % (k[:10], v["ms"]*1e3) for k,v in j.get("stages",{}).items()))
PY
done
