#!/usr/bin/env bash
# GPU call 9 of round 2: split-bf16 decode (tests + bench row), full bench line, ncu of the binning kernels.
set -uo pipefail
OUT=gpurun_out/r2i; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so > "$OUT/sha.txt"
timeout 900 python -m pytest tests/test_decoder.py tests/test_abi.py tests/test_gpu_async.py tests/test_gpu_golden.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -6 "$OUT/pytest_gpu.log"
python bench.py > "$OUT/bench_ours.json" 2> "$OUT/bench_ours.err"; echo "ours rc=$?"
python - "$OUT/bench_ours.json" <<'PY'
import json, sys
j = json.load(open(sys.argv[1])); e = j["e2e"]
print("fps %.1f (eager %.1f) fwd %.3f e2e %.1f (eager %.1f)" % (j["value"], j["value_eager"], j["fwd_only_ms_per_frame"], e["value"], e["eager_value"]))
print("roofline", j.get("roofline"))
d = j["adjacent"]["decode_relevancy"]
print({k: d[k] for k in d if k.endswith("_ms")}, d.get("decode_only"), d.get("max_abs_diff_vs_torch"), d.get("roofline_bf16x2"))
PY
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph"
$CMD > "$OUT/plain.log" 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'dup_scan_kernel|onesweep_kernel|finalize_tiles_kernel|depth_words_kernel' -s 16 -c 9 -f -o "$OUT/prof_binning" $CMD > "$OUT/ncu.log" 2>&1
echo "ncu rc=$?"; tail -2 "$OUT/ncu.log"
