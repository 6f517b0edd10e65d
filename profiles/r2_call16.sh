#!/usr/bin/env bash
# GPU call 16: timing-only experiments on the backward's accumulation (no REDs / one vector RED per lane).
set -uo pipefail
OUT=gpurun_out/r2p; mkdir -p "$OUT"
show() { python - "$1" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1]))
    print("fps %.1f ms %.3f  render_bwd %.1f" % (j["value"], j["ms_per_step"], j["stages_us"]["render_bwd"]))
except Exception as e:
    print("no line:", e)
PY
}
B="python bench.py --steps 20 --warmup 5 --no-cpu-baseline"
cp langsplat_b200/liblsr.so /tmp/liblsr_main.so
for v in main b2_nored b2_redv4; do
  if [ $v = main ]; then cp /tmp/liblsr_main.so langsplat_b200/liblsr.so; else cp build/variants/$v.so langsplat_b200/liblsr.so; fi
  $B > "$OUT/bench_$v.json" 2> "$OUT/bench_$v.err"; echo "$v rc=$?"; show "$OUT/bench_$v.json"
done
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
