#!/usr/bin/env bash
# GPU call 15: guard-banded fast exp in the backward (A/B + the full-size parity tests); ncu --set full of one step.
set -uo pipefail
OUT=gpurun_out/r2o; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so build/variants/b2_fexp2.so > "$OUT/sha.txt"
show() { python - "$1" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1]))
    print("fps %.1f ms %.3f  render_bwd %.1f  e2e %s" % (j["value"], j["ms_per_step"], j["stages_us"]["render_bwd"], (j.get("e2e") or {}).get("value")))
except Exception as e:
    print("no line:", e)
PY
}
B="python bench.py --steps 20 --warmup 5 --no-cpu-baseline"
cp langsplat_b200/liblsr.so /tmp/liblsr_main.so
for v in main b2_fexp2; do
  if [ $v = main ]; then cp /tmp/liblsr_main.so langsplat_b200/liblsr.so; else cp build/variants/$v.so langsplat_b200/liblsr.so; fi
  $B > "$OUT/bench_$v.json" 2> "$OUT/bench_$v.err"; echo "$v rc=$?"; show "$OUT/bench_$v.json"
  $B --config 3 > "$OUT/bench_c3_$v.json" 2> "$OUT/bench_c3_$v.err"; echo "c3 $v rc=$?"; show "$OUT/bench_c3_$v.json"
done
cp build/variants/b2_fexp2.so langsplat_b200/liblsr.so
rm -f gpurun_out/grad_errors.json
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_reference.py tests/test_gpu_full_size.py tests/test_gpu_reference_caller.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_fexp2.log" 2>&1; echo "pytest fexp2 rc=$?"; tail -3 "$OUT/pytest_fexp2.log"
cp gpurun_out/grad_errors.json "$OUT/grad_errors_fexp2.json" 2>/dev/null
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
rm -f gpurun_out/grad_errors.json
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_reference.py tests/test_gpu_full_size.py tests/test_gpu_async.py tests/test_gpu_canaries.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_main.log" 2>&1; echo "pytest main rc=$?"; tail -3 "$OUT/pytest_main.log"
cp gpurun_out/grad_errors.json "$OUT/grad_errors_main.json" 2>/dev/null
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'preprocess_fwd_kernel|preprocess_bwd_kernel|depth_words_kernel|dup_scan_kernel|onesweep_kernel|finalize_tiles_kernel|render_fwd_kernel|render_bwd2_kernel|poison_outputs_kernel' -s 524 -c 14 -f -o "$OUT/prof_step" $CMD > "$OUT/ncu_full.log" 2>&1
echo "ncu full rc=$?"; tail -3 "$OUT/ncu_full.log"
ls -la "$OUT"
