#!/usr/bin/env bash
# Builds the UNMODIFIED reference extension (submodules/langsplat-rasterization) for sm_100a into baseline/_ref.
# /root/reference is read-only and setup.py writes build/, so the tree is copied to a scratch dir first.
# The only flag added is `-include cstdint` (rasterizer_impl.h uses std::uintptr_t without the header; g++ 13 no
# longer pulls it in transitively).  Numerics flags stay at nvcc defaults, as in the reference's setup.py:30.
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
SRC=/root/reference/submodules/langsplat-rasterization
TMP="$(mktemp -d /tmp/refbuild.XXXX)"
cp -r "$SRC" "$TMP/rast"
cd "$TMP/rast"
mkdir -p "$ROOT/baseline/_ref"
unset CXX CC
TORCH_CUDA_ARCH_LIST=10.0a NVCC_APPEND_FLAGS="-include cstdint" MAX_JOBS=6 \
  python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
  --target "$ROOT/baseline/_ref" --upgrade . 
rm -rf "$TMP"
