#!/usr/bin/env bash
# GPU call 22 (N GPUs): multi-rank step replayed from CUDA graphs (device-epoch barriers); the new single-process tests.
set -uo pipefail
N="${1:-2}"; OUT=gpurun_out/${2:-r2v}; mkdir -p "$OUT"
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
show() { python - "$1" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("value %.1f (eager %.1f) ms %.3f e2e %s (eager %s) | %s" % (j["value"], j.get("value_eager", 0), j["ms_per_step"], (j.get("e2e") or {}).get("value"), (j.get("e2e") or {}).get("eager_value"), j.get("value_mode", "")[:40]))
except Exception as e:
    print("no line:", e)
PY
}
timeout 600 python -m pytest tests/test_gpu_sh_factors.py tests/test_gpu_canaries.py -m gpu -x -q -p no:cacheprovider > "$OUT/pytest.log" 2>&1; echo "pytest rc=$?"; tail -3 "$OUT/pytest.log"
timeout 600 $T bench.py --gpus $N --steps 20 --warmup 5 > "$OUT/bench_c2_n$N.json" 2> "$OUT/bench_c2_n$N.err"; echo "c2 rc=$?"; show "$OUT/bench_c2_n$N.json"; tail -5 "$OUT/bench_c2_n$N.err"
timeout 600 $T bench.py --gpus $N --steps 10 --warmup 3 --config 4 > "$OUT/bench_c4_n$N.json" 2> "$OUT/bench_c4_n$N.err"; echo "c4 rc=$?"; show "$OUT/bench_c4_n$N.json"; tail -5 "$OUT/bench_c4_n$N.err"
