#!/usr/bin/env bash
# GPU call 27: the sort's ranking as two loops (ballots first, then a leader-only shared-memory chain): A/B + tests.
set -uo pipefail
OUT=gpurun_out/r2aa; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so build/variants/rank2*.so > "$OUT/sha.txt"
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
for v in main rank2 rank2_minb3; do
  if [ $v = main ]; then cp /tmp/liblsr_main.so langsplat_b200/liblsr.so; else cp build/variants/$v.so langsplat_b200/liblsr.so; fi
  $B > "$OUT/bench_$v.json" 2> "$OUT/bench_$v.err"; echo "$v rc=$?"; show "$OUT/bench_$v.json"
  $B --config 5 --steps 10 > "$OUT/bench_c5_$v.json" 2> "$OUT/bench_c5_$v.err"; echo "c5 $v rc=$?"; show "$OUT/bench_c5_$v.json"
  if [ $v != main ]; then
    timeout 300 python -m pytest tests/test_gpu_sort.py tests/test_gpu_parity.py tests/test_gpu_vs_reference.py tests/test_gpu_async.py tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_$v.log" 2>&1; echo "pytest $v rc=$?"; tail -1 "$OUT/pytest_$v.log"
  fi
done
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
