#!/usr/bin/env bash
# GPU call 11: duplication kernel with the fill beside the look-back (ticket at block start): parity subset + bench.
set -uo pipefail
OUT=gpurun_out/r2k; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so > "$OUT/sha.txt"
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_reference.py tests/test_gpu_golden.py tests/test_gpu_async.py tests/test_gpu_full_size.py -m gpu -q -p no:cacheprovider -k "not config4 and not config3_full" > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -4 "$OUT/pytest_gpu.log"
for c in 2 5; do
python bench.py --config $c --steps 20 --no-cpu-baseline > "$OUT/bench_c${c}_ours.json" 2> "$OUT/bench_c${c}_ours.err"; echo "c$c rc=$?"
python - "$OUT/bench_c${c}_ours.json" <<'PY'
import json, sys
j = json.load(open(sys.argv[1])); print("fps %.1f ms %.3f" % (j["value"], j["ms_per_step"])); print(j["stages_us"])
PY
done
