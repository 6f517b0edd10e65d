#!/usr/bin/env bash
# GPU call 24 (4 GPUs): C2 and C4 at N = 4 on the final tree (graph-replayed multi-rank step).
set -uo pipefail
N=4; OUT=gpurun_out/r2x; mkdir -p "$OUT"
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
show() { python - "$1" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("value %.1f (eager %.1f) ms %.3f e2e %s (eager %s)" % (j["value"], j.get("value_eager", 0), j["ms_per_step"], (j.get("e2e") or {}).get("value"), (j.get("e2e") or {}).get("eager_value")))
except Exception as e:
    print("no line:", e)
PY
}
timeout 600 $T bench.py --gpus $N --steps 20 --warmup 5 > "$OUT/bench_c2_n$N.json" 2> "$OUT/bench_c2_n$N.err"; echo "c2 rc=$?"; show "$OUT/bench_c2_n$N.json"
timeout 600 $T bench.py --gpus $N --steps 10 --warmup 3 --config 4 > "$OUT/bench_c4_n$N.json" 2> "$OUT/bench_c4_n$N.err"; echo "c4 rc=$?"; show "$OUT/bench_c4_n$N.json"
timeout 600 $T bench.py --gpus $N --steps 10 --warmup 3 --impl reference > "$OUT/bench_ref_n$N.json" 2> "$OUT/bench_ref_n$N.err"; echo "ref rc=$?"; show "$OUT/bench_ref_n$N.json"
