#!/usr/bin/env bash
# GPU call 31: forward blend with one weight per pair (LSR_FWD_WEIGHT=1): speed, image difference to the reference, tests.
set -uo pipefail
OUT=gpurun_out/r2ae; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so build/variants/fwd_weight.so > "$OUT/sha.txt"
show() { python - "$1" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1])); s = j["stages_us"]
    print("fps %.1f ms %.3f | render_fwd %.1f fwd_only %.3f ms e2e %.1f" % (j["value"], j["ms_per_step"], s["render_fwd"], j["fwd_only_ms_per_frame"], j["e2e"]["value"]))
except Exception as e:
    print("no line:", e)
PY
}
B="python bench.py --steps 20 --warmup 5 --no-cpu-baseline"
cp langsplat_b200/liblsr.so /tmp/liblsr_main.so
for v in main fwd_weight; do
  if [ $v = main ]; then cp /tmp/liblsr_main.so langsplat_b200/liblsr.so; else cp build/variants/$v.so langsplat_b200/liblsr.so; fi
  $B > "$OUT/bench_$v.json" 2> "$OUT/bench_$v.err"; echo "$v rc=$?"; show "$OUT/bench_$v.json"
  $B --config 5 --steps 10 > "$OUT/bench_c5_$v.json" 2> "$OUT/bench_c5_$v.err"; echo "c5 $v rc=$?"; show "$OUT/bench_c5_$v.json"
  python profiles/tools/image_diff.py 2 3 > "$OUT/image_diff_$v.log" 2>&1; cat "$OUT/image_diff_$v.log" | tail -2
done
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_vs_reference.py tests/test_gpu_full_size.py tests/test_gpu_reference_caller.py tests/test_gpu_async.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_fwd_weight.log" 2>&1; echo "pytest fwd_weight rc=$?"; tail -2 "$OUT/pytest_fwd_weight.log"
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
