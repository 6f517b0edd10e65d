#!/usr/bin/env bash
# quick GPU check: sort + golden + parity tests, then one bench line.  Usage: bash profiles/quick.sh TAG [pytest -k expr]
set -uo pipefail
TAG="$1"; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
python -m pytest tests -m gpu -x -q ${2:+-k "$2"} > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -4 "$OUT/pytest_gpu.log"
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > "$OUT/bench_ours.json" 2> "$OUT/bench_ours.err"; echo "bench rc=$?"
python - <<PY
import json
j=json.load(open("$OUT/bench_ours.json"))
print("fps %.1f ms %.3f fwd %.3f e2e %.1f" % (j["value"], j["ms_per_step"], j["fwd_only_ms_per_frame"], j["e2e"]["value"]))
for k,v in j["stages"].items(): print("  %-16s %8.1f us" % (k, v["ms"]*1e3))
PY
