#!/usr/bin/env bash
# GPU call 7 of round 2: GraphedStep test fix, C4 arbitration, bench with graph-replayed value / forward-only / e2e.
set -uo pipefail
OUT=gpurun_out/r2g; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so > "$OUT/sha.txt"
timeout 900 python -m pytest tests/test_gpu_async.py tests/test_gpu_full_size.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -6 "$OUT/pytest_gpu.log"
cp gpurun_out/grad_errors.json "$OUT/grad_errors.json" 2>/dev/null
show() { python - "$1" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1]))
    e = j["e2e"]
    print("fps %.1f (eager %.1f) ms %.3f fwd %.3f (eager %.3f) e2e %.1f (eager %.1f) h2d %d launches %s" % (
        j["value"], j.get("value_eager", 0), j["ms_per_step"], j["fwd_only_ms_per_frame"],
        j.get("fwd_only_ms_per_frame_eager", 0), e["value"], e.get("eager_value", 0), e["h2d_bytes_per_step"], j["gpu_launches"]))
except Exception as ex:
    print("no line:", ex)
PY
}
python bench.py > "$OUT/bench_ours.json" 2> "$OUT/bench_ours.err"; echo "ours rc=$?"; show "$OUT/bench_ours.json"; tail -3 "$OUT/bench_ours.err"
python bench.py --config 3 --no-cpu-baseline > "$OUT/bench_c3_ours.json" 2> "$OUT/bench_c3_ours.err"; echo "c3 rc=$?"; show "$OUT/bench_c3_ours.json"
python bench.py --no-graph --no-cpu-baseline > "$OUT/bench_ours_nograph.json" 2> "$OUT/bench_ours_nograph.err"; echo "nograph rc=$?"; show "$OUT/bench_ours_nograph.json"
