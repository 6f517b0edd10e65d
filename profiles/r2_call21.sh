#!/usr/bin/env bash
# GPU call 21 (N GPUs): multi-GPU bench with the final tree (accumulation record): C2 weak scaling, default exchange and
# the all-peer-memory exchange; C4 strong scaling.  Usage: gpurun --gpus N -- 'bash profiles/r2_call21.sh N TAG'
set -uo pipefail
N="${1:-2}"; OUT=gpurun_out/${2:-r2u}; mkdir -p "$OUT"
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
show() { python - "$1" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("value %.1f ms %.3f e2e %s" % (j["value"], j["ms_per_step"], (j.get("e2e") or {}).get("value")))
except Exception as e:
    print("no line:", e)
PY
}
$T bench.py --gpus $N --steps 20 --warmup 5 > "$OUT/bench_c2_n$N.json" 2> "$OUT/bench_c2_n$N.err"; echo "c2 rc=$?"; show "$OUT/bench_c2_n$N.json"
$T bench.py --gpus $N --steps 20 --warmup 5 --peer-allreduce > "$OUT/bench_c2_n${N}_peer.json" 2> "$OUT/bench_c2_n${N}_peer.err"; echo "c2 peer rc=$?"; show "$OUT/bench_c2_n${N}_peer.json"
$T bench.py --gpus $N --steps 10 --warmup 3 --config 4 > "$OUT/bench_c4_n$N.json" 2> "$OUT/bench_c4_n$N.err"; echo "c4 rc=$?"; show "$OUT/bench_c4_n$N.json"
$T bench.py --gpus $N --steps 10 --warmup 3 --impl reference > "$OUT/bench_ref_n$N.json" 2> "$OUT/bench_ref_n$N.err"; echo "ref rc=$?"; show "$OUT/bench_ref_n$N.json"
