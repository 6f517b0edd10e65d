#!/usr/bin/env bash
# GPU call 19: the allocation-free sort entry point alone in a fresh process, then the whole GPU suite (it runs last),
# then the bench A/B of the packed accumulation record (acc / LSR_BLEND_ACC=0).
set -uo pipefail
OUT=gpurun_out/r2s; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so > "$OUT/sha.txt"
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_no_alloc_alone.log" 2>&1; echo "no_alloc alone rc=$?"; tail -3 "$OUT/pytest_no_alloc_alone.log"
rm -f gpurun_out/grad_errors.json
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -4 "$OUT/pytest_gpu.log"
cp gpurun_out/grad_errors.json "$OUT/grad_errors.json" 2>/dev/null
show() { python - "$1" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1]))
    print("fps %.1f ms %.3f  render_bwd %.1f pre_bwd %.1f zero %.1f  e2e %s" % (j["value"], j["ms_per_step"], j["stages_us"]["render_bwd"], j["stages_us"]["preprocess_bwd"], j["stages_us"]["bwd_zero_fill"], (j.get("e2e") or {}).get("value")))
except Exception as e:
    print("no line:", e)
PY
}
B="python bench.py --steps 20 --warmup 5 --no-cpu-baseline"
for v in acc noacc; do
  if [ $v = noacc ]; then export LSR_BLEND_ACC=0; else unset LSR_BLEND_ACC; fi
  $B > "$OUT/bench_$v.json" 2> "$OUT/bench_$v.err"; echo "$v rc=$?"; show "$OUT/bench_$v.json"
  $B --config 3 > "$OUT/bench_c3_$v.json" 2> "$OUT/bench_c3_$v.err"; echo "c3 $v rc=$?"; show "$OUT/bench_c3_$v.json"
  $B --config 4 > "$OUT/bench_c4_$v.json" 2> "$OUT/bench_c4_$v.err"; echo "c4 $v rc=$?"; show "$OUT/bench_c4_$v.json"
done
unset LSR_BLEND_ACC
LSR_BLEND_ACC=0 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_reference.py tests/test_gpu_canaries.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_noacc.log" 2>&1; echo "pytest noacc rc=$?"; tail -2 "$OUT/pytest_noacc.log"
