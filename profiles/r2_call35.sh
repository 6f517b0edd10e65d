#!/usr/bin/env bash
# GPU call 35: the round-end command once more on the final tree (flakiness check of `pytest -m gpu -x`).
set -uo pipefail
OUT=gpurun_out/r2ai; mkdir -p "$OUT"
timeout 1500 python -m pytest tests/ -x -q -m gpu > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -3 "$OUT/pytest_gpu.log"
