#!/usr/bin/env bash
# GPU call 13: chain shorteners in the two-phase backward (early colour record, one vote, record prefetch): A/B.
set -uo pipefail
OUT=gpurun_out/r2m; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so build/variants/b2_*.so > "$OUT/sha.txt"
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
for v in main b2_early b2_vote b2_both b2_all; do
  if [ $v = main ]; then cp /tmp/liblsr_main.so langsplat_b200/liblsr.so; else cp build/variants/$v.so langsplat_b200/liblsr.so; fi
  $B > "$OUT/bench_$v.json" 2> "$OUT/bench_$v.err"; echo "$v rc=$?"; show "$OUT/bench_$v.json"
  $B --config 3 > "$OUT/bench_c3_$v.json" 2> "$OUT/bench_c3_$v.err"; echo "c3 $v rc=$?"; show "$OUT/bench_c3_$v.json"
done
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_reference.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_main.log" 2>&1; echo "pytest main rc=$?"; tail -2 "$OUT/pytest_main.log"
cp build/variants/b2_all.so langsplat_b200/liblsr.so
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_reference.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_b2_all.log" 2>&1; echo "pytest b2_all rc=$?"; tail -2 "$OUT/pytest_b2_all.log"
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
