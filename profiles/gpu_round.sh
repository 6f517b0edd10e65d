#!/usr/bin/env bash
# One gpurun call: GPU tests, golden vectors from the reference extension, both bench arms, ncu launch list and one
# `--set full` capture of one whole step's kernels.  Usage: gpurun --timeout 1500 -- 'bash profiles/gpu_round.sh TAG'
set -uo pipefail
TAG="${1:-r01}"
OUT=gpurun_out/$TAG
mkdir -p "$OUT"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "$OUT/smi.csv" 2>&1
python -m pytest tests -m gpu -x -q > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?" | tee -a "$OUT/pytest_gpu.log"
tail -5 "$OUT/pytest_gpu.log"
if [ "${GOLDEN:-0}" = 1 ]; then python tests/golden/make_golden.py "$OUT/golden" > "$OUT/golden.log" 2>&1; tail -4 "$OUT/golden.log"; fi
python bench.py --impl reference --steps 20 --warmup 5 > "$OUT/bench_reference.json" 2> "$OUT/bench_reference.err"; echo "ref rc=$?"
python bench.py --steps 30 --warmup 5 > "$OUT/bench_ours.json" 2> "$OUT/bench_ours.err"; echo "ours rc=$?"
cat "$OUT/bench_ours.json" | head -c 3000; echo
if [ "${NCU:-1}" = 1 ]; then
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
  $CMD > "$OUT/plain.log" 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'preprocess_fwd_kernel|preprocess_bwd_kernel|depth_words_kernel|scan_tiles_kernel|duplicate_kernel|dup_scan_kernel|onesweep_kernel|tile_ranges_kernel|tile_order_kernel|finalize_tiles_kernel|render_fwd_kernel|render_bwd_kernel' -s 42 -c 28 --csv \
      --log-file "$OUT/launches.csv" $CMD > "$OUT/ncu_launches.log" 2>&1
  echo "ncu launches rc=$?"
  ncu --set full --clock-control none --import-source on -k regex:'preprocess_fwd_kernel|preprocess_bwd_kernel|depth_words_kernel|scan_tiles_kernel|duplicate_kernel|dup_scan_kernel|onesweep_kernel|tile_ranges_kernel|tile_order_kernel|finalize_tiles_kernel|render_fwd_kernel|render_bwd_kernel' -s 42 -c 14 -f -o "$OUT/prof_step" $CMD \
      > "$OUT/ncu_full.log" 2>&1
  echo "ncu full rc=$?"
fi
ls -la "$OUT"
