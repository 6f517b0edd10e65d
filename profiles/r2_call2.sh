#!/usr/bin/env bash
# GPU call 2 of round 2: full GPU test suite on the device-planned forward, gradient-error diagnostic (main, IEEE
# division, direct sums instead of moments), bench A/B (async / sync / legacy binning / look-back window 4).
set -uo pipefail
OUT=gpurun_out/r2b; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so build/variants/*.so > "$OUT/sha.txt"
timeout 900 python -m pytest tests -m gpu -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -15 "$OUT/pytest_gpu.log"
show() { python - "$1" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1]))
    print("fps %.1f ms %.3f fwd %.3f e2e %.1f" % (j["value"], j["ms_per_step"], j["fwd_only_ms_per_frame"], j["e2e"]["value"]))
    print("  " + "  ".join("%s %.1f" % (k, v) for k, v in j.get("stages_us", {}).items()))
except Exception as e:
    print("no line:", e)
PY
}
B="python bench.py --steps 30 --warmup 5 --no-cpu-baseline"
$B > "$OUT/bench_async.json" 2> "$OUT/bench_async.err"; echo "async rc=$?"; show "$OUT/bench_async.json"
$B --sync-forward > "$OUT/bench_sync.json" 2> "$OUT/bench_sync.err"; echo "sync rc=$?"; show "$OUT/bench_sync.json"
LSR_LEGACY_BINNING=1 $B --sync-forward > "$OUT/bench_legacy.json" 2> "$OUT/bench_legacy.err"; echo "legacy rc=$?"; show "$OUT/bench_legacy.json"
cp langsplat_b200/liblsr.so /tmp/liblsr_main.so
cp build/variants/lbwin4.so langsplat_b200/liblsr.so
$B > "$OUT/bench_lbwin4.json" 2> "$OUT/bench_lbwin4.err"; echo "lbwin4 rc=$?"; show "$OUT/bench_lbwin4.json"
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
for v in main ieeediv nomoments; do
  if [ $v = main ]; then L=/tmp/liblsr_main.so; else L=build/variants/$v.so; fi
  timeout 600 python profiles/tools/grad_diag.py --views 2 --lib $L > "$OUT/diag_$v.log" 2>&1; echo "diag $v rc=$?"; tail -18 "$OUT/diag_$v.log"
done
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
