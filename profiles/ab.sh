#!/usr/bin/env bash
# A/B on the GPU box: GPU tests + bench for the in-tree liblsr.so, then bench only for each build/variants/NAME.so.
# Usage: bash profiles/ab.sh TAG [variant names...]
set -uo pipefail
TAG="$1"; shift; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
show() { python - "$1" <<'PY'
import json, sys
j = json.load(open(sys.argv[1]))
print("fps %.1f ms %.3f fwd %.3f e2e %.1f" % (j["value"], j["ms_per_step"], j["fwd_only_ms_per_frame"], j["e2e"]["value"]))
print("  " + "  ".join("%s %.1f" % (k, v["ms"] * 1e3) for k, v in j["stages"].items()))
PY
}
python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -4 "$OUT/pytest_gpu.log"
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > "$OUT/bench_main.json" 2> "$OUT/bench_main.err"; echo "main rc=$?"; show "$OUT/bench_main.json"
cp langsplat_b200/liblsr.so /tmp/liblsr_main.so
for v in "$@"; do
  cp "build/variants/$v.so" langsplat_b200/liblsr.so
  python bench.py --steps 30 --warmup 5 --no-cpu-baseline > "$OUT/bench_$v.json" 2> "$OUT/bench_$v.err"; echo "$v rc=$?"; show "$OUT/bench_$v.json"
done
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
