#!/usr/bin/env bash
# GPU call 34 (2 GPUs): the driver's multi-GPU command on the final tree (default flags), both arms.
set -uo pipefail
N=2; OUT=gpurun_out/r2ah; mkdir -p "$OUT"
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519"
timeout 600 $T bench.py --gpus $N --steps 20 --warmup 5 > "$OUT/bench_c2_n$N.json" 2> "$OUT/bench_c2_n$N.err"; echo "ours rc=$?"
timeout 600 $T bench.py --impl reference --gpus $N --steps 10 --warmup 3 > "$OUT/bench_ref_n$N.json" 2> "$OUT/bench_ref_n$N.err"; echo "ref rc=$?"
python - "$OUT" <<'PY'
import json, sys
for n in ("c2_n2", "ref_n2"):
    j = json.loads(open(sys.argv[1] + "/bench_%s.json" % n).read().strip().splitlines()[-1])
    print(n, "value %.1f e2e %.1f" % (j["value"], j["e2e"]["value"]), j.get("value_mode", "")[:50], j.get("clocks"))
PY
