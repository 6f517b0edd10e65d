#!/usr/bin/env bash
# GPU call 5 of round 2: defaults after r2d (two-phase backward without pipelining, ballots in the depth passes, status
# pool), branch-free back half as a variant, C4 arbitration details, compute-sanitizer memcheck on smoke() and the
# small parity tests.
set -uo pipefail
OUT=gpurun_out/r2e; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so build/variants/*.so > "$OUT/sha.txt"
timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -8 "$OUT/pytest_gpu.log"
cp gpurun_out/grad_errors.json "$OUT/grad_errors.json" 2>/dev/null
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
$B > "$OUT/bench_main2.json" 2> "$OUT/bench_main2.err"; echo "main again rc=$?"; show "$OUT/bench_main2.json"
$B --sync-forward > "$OUT/bench_sync.json" 2> "$OUT/bench_sync.err"; echo "sync rc=$?"; show "$OUT/bench_sync.json"
cp langsplat_b200/liblsr.so /tmp/liblsr_main.so
for v in bwd2_flat; do
  cp build/variants/$v.so langsplat_b200/liblsr.so
  $B > "$OUT/bench_$v.json" 2> "$OUT/bench_$v.err"; echo "$v rc=$?"; show "$OUT/bench_$v.json"
done
cp /tmp/liblsr_main.so langsplat_b200/liblsr.so
timeout 900 compute-sanitizer --tool memcheck --log-file "$OUT/memcheck_smoke.log" python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/memcheck_smoke.out" 2>&1; echo "memcheck smoke rc=$?"; tail -3 "$OUT/memcheck_smoke.log"
timeout 1500 compute-sanitizer --tool memcheck --log-file "$OUT/memcheck_parity.log" python -m pytest tests/test_gpu_parity.py tests/test_gpu_async.py -q -p no:cacheprovider -k "config1 or no_feature or feature_only or all_culled or bit_identical or overflow_poisons" > "$OUT/memcheck_parity.out" 2>&1; echo "memcheck parity rc=$?"; tail -3 "$OUT/memcheck_parity.log"; tail -3 "$OUT/memcheck_parity.out"
