#!/usr/bin/env bash
# GPU call 25: binning A/B — keys per thread of the onesweep sort (12 / 20 / 24 against 16), 128-Gaussian blocks in the
# duplication kernel (twice the resident blocks to hide the look-back wait ncu shows, profiles/r2t).
set -uo pipefail
OUT=gpurun_out/r2y; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so build/variants/sort_items*.so build/variants/dup128.so > "$OUT/sha.txt"
show() { python - "$1" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1])); s = j["stages_us"]
    print("fps %.1f ms %.3f | depth_order %.1f duplicate %.1f sort %.1f" % (j["value"], j["ms_per_step"], s["depth_order"], s["duplicate"], s["sort"]))
except Exception as e:
    print("no line:", e)
PY
}
B="python bench.py --steps 20 --warmup 5 --no-cpu-baseline"
cp langsplat_b200/liblsr.so /tmp/liblsr_main.so
for v in main sort_items12 sort_items20 sort_items24 dup128; do
  if [ $v = main ]; then cp /tmp/liblsr_main.so langsplat_b200/liblsr.so; else cp build/variants/$v.so langsplat_b200/liblsr.so; fi
  $B > "$OUT/bench_$v.json" 2> "$OUT/bench_$v.err"; echo "$v rc=$?"; show "$OUT/bench_$v.json"
  if [ $v != main ]; then
    timeout 300 python -m pytest tests/test_gpu_sort.py tests/test_gpu_parity.py tests/test_gpu_async.py tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_$v.log" 2>&1; echo "pytest $v rc=$?"; tail -1 "$OUT/pytest_$v.log"
  fi
done
cp langsplat_b200/liblsr.so /dev/null; cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
$B --config 5 --steps 10 > "$OUT/bench_c5_main.json" 2> "$OUT/bench_c5_main.err"; echo "c5 main rc=$?"; show "$OUT/bench_c5_main.json"
cp build/variants/dup128.so langsplat_b200/liblsr.so
$B --config 5 --steps 10 > "$OUT/bench_c5_dup128.json" 2> "$OUT/bench_c5_dup128.err"; echo "c5 dup128 rc=$?"; show "$OUT/bench_c5_dup128.json"
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
