#!/usr/bin/env bash
# GPU call 30: the tests behind the one that stopped r2ac's `pytest -x` (loss tolerance), three times for flakiness.
set -uo pipefail
OUT=gpurun_out/r2ad; mkdir -p "$OUT"
for i in 1 2 3; do
  timeout 600 python -m pytest tests/test_loss.py tests/test_zz_no_alloc.py tests/test_gpu_canaries.py -m gpu -q -p no:cacheprovider > "$OUT/pytest_$i.log" 2>&1; echo "pytest $i rc=$?"; tail -1 "$OUT/pytest_$i.log"
done
