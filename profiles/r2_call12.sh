#!/usr/bin/env bash
# GPU call 12: preprocess_bwd rewrite, guard-band tests, full suite, bench (C2, C5 with the three-call sequence).
set -uo pipefail
OUT=gpurun_out/r2l; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so > "$OUT/sha.txt"
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -6 "$OUT/pytest_gpu.log"
cp gpurun_out/grad_errors.json "$OUT/grad_errors.json" 2>/dev/null
python bench.py --no-cpu-baseline > "$OUT/bench_ours.json" 2> "$OUT/bench_ours.err"; echo "ours rc=$?"
python - "$OUT/bench_ours.json" <<'PY'
import json, sys
j = json.load(open(sys.argv[1])); e = j["e2e"]
print("fps %.1f (eager %.1f) fwd %.3f e2e %.1f (eager %.1f)" % (j["value"], j["value_eager"], j["fwd_only_ms_per_frame"], e["value"], e["eager_value"]))
print(j["stages_us"])
PY
python bench.py --config 5 --steps 20 --no-cpu-baseline > "$OUT/bench_c5_ours.json" 2> "$OUT/bench_c5_ours.err"; echo "c5 rc=$?"
python - "$OUT/bench_c5_ours.json" <<'PY'
import json, sys
j = json.load(open(sys.argv[1])); print("c5 fps %.1f ms %.3f" % (j["value"], j["ms_per_step"]), j.get("three_call_sequence"))
PY
