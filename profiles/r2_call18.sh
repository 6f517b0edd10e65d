#!/usr/bin/env bash
# GPU call 18: full GPU tests without -x (which test fails with the packed accumulation record?)
set -uo pipefail
OUT=gpurun_out/r2r; mkdir -p "$OUT"
rm -f gpurun_out/grad_errors.json
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -5 "$OUT/pytest_gpu.log"
cp gpurun_out/grad_errors.json "$OUT/grad_errors.json" 2>/dev/null
grep -E "^(FAILED|ERROR)|Error|assert " "$OUT/pytest_gpu.log" | head -40
