#!/usr/bin/env bash
# Builds a variant of liblsr.so with extra nvcc flags into build/variants/NAME.so (for A/B runs on the GPU box:
# `cp build/variants/NAME.so langsplat_b200/liblsr.so` inside the gpurun command).  Usage: build_variant.sh NAME "-DX=1 ..."
set -euo pipefail
NAME="$1"; FLAGS="${2:-}"
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
D="$ROOT/build/variants/$NAME.d"; mkdir -p "$D"
cd "$ROOT/langsplat_b200/csrc"
for f in api preprocess_fwd binning radix_sort render_fwd render_bwd render_bwd2 preprocess_bwd loss peer activations gt_features adam decoder; do
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC \
    -Xcompiler -fvisibility=hidden -Xptxas -v $FLAGS -c $f.cu -o "$D/$f.o" 2> "$D/$f.ptxas.log" &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$ROOT/build/variants/$NAME.so" "$D"/*.o
grep -h "Used" "$D/render_fwd.ptxas.log" "$D/render_bwd.ptxas.log" | sort | uniq -c
