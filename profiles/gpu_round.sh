#!/usr/bin/env bash
# One gpurun call: GPU tests, both bench arms, the ncu launch list of one bench run and one `--set full` capture of the
# kernels of its last forward+backward step.  Usage: gpurun --timeout 2400 -- 'bash profiles/gpu_round.sh TAG'
set -uo pipefail
TAG="${1:-r2z}"
OUT=gpurun_out/$TAG
mkdir -p "$OUT"
RE='preprocess_fwd_kernel|preprocess_bwd_kernel|depth_words_kernel|dup_scan_kernel|onesweep_kernel|finalize_tiles_kernel|render_fwd_kernel|render_bwd2_kernel|poison_outputs_kernel|unpack_acc_kernel'
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "$OUT/smi.csv" 2>&1
sha256sum langsplat_b200/liblsr.so > "$OUT/sha.txt"
rm -f gpurun_out/grad_errors.json
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?" | tee -a "$OUT/pytest_gpu.log"
tail -4 "$OUT/pytest_gpu.log"
cp gpurun_out/grad_errors.json "$OUT/grad_errors.json" 2>/dev/null
if [ "${GOLDEN:-0}" = 1 ]; then python tests/golden/make_golden.py "$OUT/golden" > "$OUT/golden.log" 2>&1; tail -4 "$OUT/golden.log"; fi
python bench.py --impl reference --steps 20 --warmup 5 > "$OUT/bench_reference.json" 2> "$OUT/bench_reference.err"; echo "ref rc=$?"
python bench.py --steps 30 --warmup 5 > "$OUT/bench_ours.json" 2> "$OUT/bench_ours.err"; echo "ours rc=$?"
cat "$OUT/bench_ours.json" | head -c 2500; echo
if [ "${NCU:-1}" = 1 ]; then
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
  $CMD > "$OUT/plain.log" 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file "$OUT/launches.csv" $CMD > "$OUT/ncu_launches.log" 2>&1
  echo "ncu launches rc=$?"
  read SKIP CNT < <(python profiles/skip_for_last_step.py "$OUT/launches.csv" "$RE")
  echo "full capture: skip $SKIP count $CNT"
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"$RE" -s "$SKIP" -c "$CNT" -f -o "$OUT/prof_step" $CMD > "$OUT/ncu_full.log" 2>&1
  echo "ncu full rc=$?"
  ncu -i "$OUT/prof_step.ncu-rep" --page raw --csv > "$OUT/prof_step_raw.csv" 2> /dev/null; echo "raw export rc=$?"
fi
ls -la "$OUT"
