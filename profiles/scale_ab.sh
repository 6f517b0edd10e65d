#!/usr/bin/env bash
# Multi-GPU A/B on the box: the 2+-rank peer-exchange check, then bench at N ranks with the three SH-gradient exchanges
# (default: factors read from peer memory + NCCL for the rest; FULL=1 adds: everything over peer memory / NCCL
# all-gather of factors / dense all-reduce).
# Usage: bash profiles/scale_ab.sh TAG N [config]
set -uo pipefail
TAG="$1"; N="$2"; CFG="${3:-2}"; OUT=gpurun_out/$TAG; mkdir -p "$OUT"
run() { # name, extra flags
  python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus "$N" --config "$CFG" --steps 30 --warmup 5 --no-cpu-baseline $2 \
    > "$OUT/bench_c${CFG}_n${N}_$1.json" 2> "$OUT/bench_c${CFG}_n${N}_$1.err"
  echo "$1 rc=$?"
  python - "$OUT/bench_c${CFG}_n${N}_$1.json" <<'PY'
import json, sys
j = json.load(open(sys.argv[1]))
print("  %s %.1f  ms/step %.3f  e2e %.1f  launches %s" % (j["unit"], j["value"], j["ms_per_step"], j["e2e"]["value"], j["gpu_launches"]))
PY
}
if [ "$CFG" = 2 ]; then
  python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29534 \
    tests/peer_exchange_check.py > "$OUT/peer_check_n${N}.log" 2>&1
  echo "peer check rc=$?"; grep -h "PEER_EXCHANGE_OK\|Error\|unavailable" "$OUT/peer_check_n${N}.log" | head -5
fi
run peer ""
if [ "${FULL:-0}" = 1 ]; then
  run peer_allreduce "--peer-allreduce"
  run allgather "--allgather-sh-exchange"
  run dense "--dense-sh-exchange"
fi
