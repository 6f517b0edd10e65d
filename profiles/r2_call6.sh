#!/usr/bin/env bash
# GPU call 6 of round 2: bench with the graph-replayed end-to-end step and the u8 / cached ground truth, side configs,
# reference arm, C4 arbitration test.
set -uo pipefail
OUT=gpurun_out/r2f; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so > "$OUT/sha.txt"
timeout 900 python -m pytest tests/test_gpu_async.py tests/test_gpu_full_size.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -8 "$OUT/pytest_gpu.log"
cp gpurun_out/grad_errors.json "$OUT/grad_errors.json" 2>/dev/null
show() { python - "$1" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1]))
    e = j["e2e"]
    print("fps %.1f ms %.3f fwd %.3f e2e %.1f (%s; eager %.1f) h2d %d" % (j["value"], j["ms_per_step"], j["fwd_only_ms_per_frame"], e["value"], e.get("mode"), e.get("eager_value", 0), e["h2d_bytes_per_step"]))
except Exception as ex:
    print("no line:", ex)
PY
}
python bench.py > "$OUT/bench_ours.json" 2> "$OUT/bench_ours.err"; echo "ours rc=$?"; show "$OUT/bench_ours.json"; tail -3 "$OUT/bench_ours.err"
python bench.py --impl reference > "$OUT/bench_reference.json" 2> "$OUT/bench_reference.err"; echo "ref rc=$?"; show "$OUT/bench_reference.json"
B="python bench.py --steps 20 --warmup 5 --no-cpu-baseline"
$B --gt f32 > "$OUT/bench_ours_f32gt.json" 2> "$OUT/bench_ours_f32gt.err"; echo "f32 gt rc=$?"; show "$OUT/bench_ours_f32gt.json"
for c in 3 4 5; do
  $B --config $c > "$OUT/bench_c${c}_ours.json" 2> "$OUT/bench_c${c}_ours.err"; echo "c$c ours rc=$?"; show "$OUT/bench_c${c}_ours.json"; tail -2 "$OUT/bench_c${c}_ours.err"
  $B --config $c --impl reference > "$OUT/bench_c${c}_reference.json" 2> "$OUT/bench_c${c}_reference.err"; echo "c$c ref rc=$?"; show "$OUT/bench_c${c}_reference.json"
done
$B --config 5 --separate-levels > "$OUT/bench_c5_ours_separate.json" 2> "$OUT/bench_c5_ours_separate.err"; echo "c5 separate rc=$?"; show "$OUT/bench_c5_ours_separate.json"
