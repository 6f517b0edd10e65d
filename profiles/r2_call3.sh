#!/usr/bin/env bash
# GPU call 3 of round 2: two-phase render_bwd (tests, A/B against the one-phase kernel, register-cap variants, ncu),
# 8-view gradient diagnostic.
set -uo pipefail
OUT=gpurun_out/r2c; mkdir -p "$OUT"
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
$B > "$OUT/bench_bwd2.json" 2> "$OUT/bench_bwd2.err"; echo "bwd2 rc=$?"; show "$OUT/bench_bwd2.json"
LSR_BWD_V1=1 $B > "$OUT/bench_bwd1.json" 2> "$OUT/bench_bwd1.err"; echo "bwd1 rc=$?"; show "$OUT/bench_bwd1.json"
$B --config 3 > "$OUT/bench_c3_bwd2.json" 2> "$OUT/bench_c3_bwd2.err"; echo "c3 bwd2 rc=$?"; show "$OUT/bench_c3_bwd2.json"
LSR_BWD_V1=1 $B --config 3 > "$OUT/bench_c3_bwd1.json" 2> "$OUT/bench_c3_bwd1.err"; echo "c3 bwd1 rc=$?"; show "$OUT/bench_c3_bwd1.json"
cp langsplat_b200/liblsr.so /tmp/liblsr_main.so
for v in bwd2_minb20 bwd2_minb28; do
  cp build/variants/$v.so langsplat_b200/liblsr.so
  $B > "$OUT/bench_$v.json" 2> "$OUT/bench_$v.err"; echo "$v rc=$?"; show "$OUT/bench_$v.json"
done
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
timeout 900 python profiles/tools/grad_diag.py --views 8 > "$OUT/diag8_main.log" 2>&1; echo "diag8 rc=$?"; tail -18 "$OUT/diag8_main.log"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > "$OUT/plain.log" 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'render_bwd2_kernel' -s 3 -c 2 -f -o "$OUT/prof_bwd2" $CMD > "$OUT/ncu.log" 2>&1
echo "ncu rc=$?"; tail -3 "$OUT/ncu.log"
