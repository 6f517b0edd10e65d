#!/usr/bin/env bash
# GPU call 26: final tree — smoke(), full GPU suite, both bench arms with default flags.
set -uo pipefail
OUT=gpurun_out/${1:-r2z}; mkdir -p "$OUT"
sha256sum langsplat_b200/liblsr.so > "$OUT/sha.txt"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > "$OUT/smoke.log" 2>&1; echo "smoke rc=$?"; tail -2 "$OUT/smoke.log"
rm -f gpurun_out/grad_errors.json
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > "$OUT/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -3 "$OUT/pytest_gpu.log"
cp gpurun_out/grad_errors.json "$OUT/grad_errors.json" 2>/dev/null
python bench.py --impl reference > "$OUT/bench_reference.json" 2> "$OUT/bench_reference.err"; echo "ref rc=$?"
python bench.py > "$OUT/bench_ours.json" 2> "$OUT/bench_ours.err"; echo "ours rc=$?"
python - "$OUT" <<'PY'
import json, sys
for n in ("reference", "ours"):
    j = json.loads(open(sys.argv[1] + "/bench_%s.json" % n).read().strip().splitlines()[-1])
    print(n, "value %.1f e2e %.1f" % (j["value"], j["e2e"]["value"]), j.get("roofline", {}) and j["roofline"].get("frac"))
PY
