#!/usr/bin/env bash
# Builds the UNMODIFIED reference rasterizer (RaDe-GS diff-gaussian-rasterization, used by
# IGS infer_batch.py:23) for sm_100a into baseline/_ref/ (git-ignored; travels with gpurun).
# The reference source tree is read-only, so the build runs on a copy under /tmp.
# The only deviation from `python setup.py build_ext` is `-include cstdint`, needed because
# cuda_rasterizer/rasterizer_impl.h uses uintptr_t/uint32_t without including <cstdint>
# (gcc 13 rejects it).  No reference source is copied into this repository.
set -euo pipefail
REF=${REF:-/root/reference/submodules/RaDe-GS/submodules/diff-gaussian-rasterization}
OUT="$(cd "$(dirname "$0")" && pwd)/_ref"
[ -d "$REF" ] || { echo "reference tree not found: $REF" >&2; exit 3; }
TMP=$(mktemp -d /tmp/ref_build.XXXXXX)
cp -rL "$REF" "$TMP/src"
cd "$TMP/src"
export NVCC_APPEND_FLAGS="-include cstdint"
export TORCH_CUDA_ARCH_LIST="${TORCH_CUDA_ARCH_LIST:-10.0a+PTX}"
export MAX_JOBS=${MAX_JOBS:-8}
python setup.py build_ext --inplace > "$TMP/build.log" 2>&1 || { tail -50 "$TMP/build.log"; exit 1; }
mkdir -p "$OUT"
rm -rf "$OUT/diff_gaussian_rasterization_rade"
cp -r diff_gaussian_rasterization_rade "$OUT/"
ls -la "$OUT/diff_gaussian_rasterization_rade"
echo "reference built into $OUT (log: $TMP/build.log)"
