#!/usr/bin/env bash
# GPU call 8 of round 2 (8 GPUs): peer-exchange check, C2 weak scaling at N=8, exchange breakdown at N=8, C4 (3 M
# Gaussians, 8 views per step) strong scaling at N = 8, 4, 2, second-device test.
set -uo pipefail
OUT=gpurun_out/r2h; mkdir -p "$OUT"
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > "$OUT/smi.csv" 2>&1
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider -k "second_device" > "$OUT/pytest_second_device.log" 2>&1; echo "second device rc=$?"; tail -2 "$OUT/pytest_second_device.log"
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
show() { python - "$1" <<'PY'
import json, sys
try:
    j = json.load(open(sys.argv[1]))
    print("  n=%d %s %.1f  ms/step %.3f  e2e %.1f  h2d %d" % (j["n_gpus"], j["unit"], j["value"], j["ms_per_step"], j["e2e"]["value"], j["e2e"]["h2d_bytes_per_step"]))
except Exception as ex:
    print("  no line:", ex)
PY
}
timeout 300 $TR --nproc-per-node 8 --master-port 29534 tests/peer_exchange_check.py > "$OUT/peer_check_n8.log" 2>&1; echo "peer check rc=$?"; grep -h "PEER_EXCHANGE_OK\|Error\|unavailable" "$OUT/peer_check_n8.log" | head -3
timeout 300 $TR --nproc-per-node 8 --master-port 29535 profiles/tools/exchange_times.py > "$OUT/exchange_times_n8.log" 2>&1; echo "exchange times rc=$?"; grep -v "^\[" "$OUT/exchange_times_n8.log" | tail -12
for n in 8 4 2; do
  timeout 400 $TR --nproc-per-node $n --master-port 2954$n bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline > "$OUT/bench_c2_n$n.json" 2> "$OUT/bench_c2_n$n.err"; echo "c2 n=$n rc=$?"; show "$OUT/bench_c2_n$n.json"
  timeout 400 $TR --nproc-per-node $n --master-port 2955$n bench.py --gpus $n --config 4 --steps 10 --warmup 3 --no-cpu-baseline > "$OUT/bench_c4_n$n.json" 2> "$OUT/bench_c4_n$n.err"; echo "c4 n=$n rc=$?"; show "$OUT/bench_c4_n$n.json"
done
timeout 400 $TR --nproc-per-node 8 --master-port 29561 bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu-baseline --gt f32 > "$OUT/bench_c2_n8_f32gt.json" 2> "$OUT/bench_c2_n8_f32gt.err"; echo "c2 n=8 f32 gt rc=$?"; show "$OUT/bench_c2_n8_f32gt.json"
