#!/usr/bin/env bash
# GPU call 4 of round 2: software-pipelined two-phase render_bwd (tests, A/B), sort variants, ncu of the new kernel.
set -uo pipefail
OUT=gpurun_out/r2d; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so build/variants/*.so > "$OUT/sha.txt"
timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -8 "$OUT/pytest_gpu.log"
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
$B > "$OUT/bench_main.json" 2> "$OUT/bench_main.err"; echo "main rc=$?"; show "$OUT/bench_main.json"
$B --config 3 > "$OUT/bench_c3_main.json" 2> "$OUT/bench_c3_main.err"; echo "c3 main rc=$?"; show "$OUT/bench_c3_main.json"
cp langsplat_b200/liblsr.so /tmp/liblsr_main.so
for v in bwd2_nopipe bwd2_pipe20 sort_minb3 depth_nb8; do
  cp build/variants/$v.so langsplat_b200/liblsr.so
  $B > "$OUT/bench_$v.json" 2> "$OUT/bench_$v.err"; echo "$v rc=$?"; show "$OUT/bench_$v.json"
done
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > "$OUT/plain.log" 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'render_bwd2_kernel' -s 3 -c 1 -f -o "$OUT/prof_bwd2" $CMD > "$OUT/ncu.log" 2>&1
echo "ncu rc=$?"; tail -2 "$OUT/ncu.log"
