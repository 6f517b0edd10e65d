#!/usr/bin/env bash
# Stages the reference's own CALLER of the rasterizer — gaussian_renderer/, scene/, utils/, arguments/ of
# /root/reference, byte for byte — next to the built reference extension, under baseline/_ref/langsplat_caller/.
# Like baseline/_ref itself this is a reference install: git-ignored (never part of the history), but it travels to
# the GPU box with the snapshot, where /root/reference does not exist.  tests/test_gpu_reference_caller.py imports
# gaussian_renderer.render() from there UNMODIFIED and runs it against this repo's diff_gaussian_rasterization
# package and against the reference extension.
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
SRC=/root/reference
DST="$ROOT/baseline/_ref/langsplat_caller"
rm -rf "$DST"
mkdir -p "$DST"
for d in gaussian_renderer scene utils arguments; do
  mkdir -p "$DST/$d"
  cp "$SRC/$d"/*.py "$DST/$d/"
done
( cd "$SRC" && sha256sum gaussian_renderer/*.py scene/*.py utils/*.py arguments/*.py ) > "$DST/SHA256SUMS"
( cd "$DST" && sha256sum -c SHA256SUMS --quiet )
echo "staged $(wc -l < "$DST/SHA256SUMS") files into $DST"
