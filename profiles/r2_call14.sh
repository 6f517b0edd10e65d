#!/usr/bin/env bash
# GPU call 14: backward A/B (no vote, fast exp), uint8 ground truth in the loss; full GPU tests; ncu launch list.
set -uo pipefail
OUT=gpurun_out/r2n; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so build/variants/b2_novote.so build/variants/b2_fexp.so build/variants/b2_fexp_nv.so > "$OUT/sha.txt"
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
for v in main b2_novote b2_fexp b2_fexp_nv; do
  if [ $v = main ]; then cp /tmp/liblsr_main.so langsplat_b200/liblsr.so; else cp build/variants/$v.so langsplat_b200/liblsr.so; fi
  $B > "$OUT/bench_$v.json" 2> "$OUT/bench_$v.err"; echo "$v rc=$?"; show "$OUT/bench_$v.json"
  $B --config 3 > "$OUT/bench_c3_$v.json" 2> "$OUT/bench_c3_$v.err"; echo "c3 $v rc=$?"; show "$OUT/bench_c3_$v.json"
done
cp build/variants/b2_fexp_nv.so langsplat_b200/liblsr.so
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_reference.py tests/test_gpu_full_size.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_fexp_nv.log" 2>&1; echo "pytest fexp_nv rc=$?"; tail -3 "$OUT/pytest_fexp_nv.log"
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
timeout 1200 python -m pytest tests -m gpu -x -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest main rc=$?"; tail -3 "$OUT/pytest_gpu.log"
cp gpurun_out/grad_errors.json "$OUT/" 2>/dev/null
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file "$OUT/launches.csv" $CMD > "$OUT/ncu_launches.log" 2>&1
echo "ncu launches rc=$?"; wc -l "$OUT/launches.csv"
ls -la "$OUT"
