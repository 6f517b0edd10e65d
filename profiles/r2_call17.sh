#!/usr/bin/env bash
# GPU call 17: packed accumulation record in the blend backward (vector REDs): full GPU tests, A/B against LSR_BLEND_ACC=0.
set -uo pipefail
OUT=gpurun_out/r2q; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so > "$OUT/sha.txt"
show() { python - "$1" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1]))
    print("fps %.1f ms %.3f  render_bwd %.1f pre_bwd %.1f zero %.1f  e2e %s" % (j["value"], j["ms_per_step"], j["stages_us"]["render_bwd"], j["stages_us"]["preprocess_bwd"], j["stages_us"]["bwd_zero_fill"], (j.get("e2e") or {}).get("value")))
except Exception as e:
    print("no line:", e)
PY
}
rm -f gpurun_out/grad_errors.json
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -3 "$OUT/pytest_gpu.log"
cp gpurun_out/grad_errors.json "$OUT/grad_errors.json" 2>/dev/null
B="python bench.py --steps 20 --warmup 5 --no-cpu-baseline"
for v in acc noacc; do
  if [ $v = noacc ]; then export LSR_BLEND_ACC=0; else unset LSR_BLEND_ACC; fi
  $B > "$OUT/bench_$v.json" 2> "$OUT/bench_$v.err"; echo "$v rc=$?"; show "$OUT/bench_$v.json"
  $B --config 3 > "$OUT/bench_c3_$v.json" 2> "$OUT/bench_c3_$v.err"; echo "c3 $v rc=$?"; show "$OUT/bench_c3_$v.json"
  $B --config 4 > "$OUT/bench_c4_$v.json" 2> "$OUT/bench_c4_$v.err"; echo "c4 $v rc=$?"; show "$OUT/bench_c4_$v.json"
done
unset LSR_BLEND_ACC
LSR_BLEND_ACC=0 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_reference.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_noacc.log" 2>&1; echo "pytest noacc rc=$?"; tail -2 "$OUT/pytest_noacc.log"
