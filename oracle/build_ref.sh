#!/usr/bin/env bash
# TEST INFRASTRUCTURE -- not part of the product.
#
# Builds the reference's own CPU branch of the word-count path, unmodified except
# for the one compile-time switch the reference itself provides
# (MapReduce/src/main.cu:21, `#define GPU_IMPLEMENTATION 1` -> 0; it is a plain
# #define, so -D cannot override it).  The sources are compiled from where they
# lie under $REF (default /root/reference); the only thing this script keeps is
# the binary oracle/_ref/MapReduce_cpu (git-ignored).  The patched main.cu lives
# in a scratch directory under oracle/_ref/ for the duration of the compile and
# is removed again, so no reference source stays in this repository.
#
# nvcc is needed as the compiler driver (the CPU branch still contains an unused
# __global__ kernel, main.cu:319-324); no GPU is needed to run the result.
set -euo pipefail
REF="${REF:-/root/reference}"
SRC="$REF/MapReduce/src"
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/_ref"
if [ ! -f "$SRC/main.cu" ]; then
  echo "build_ref: $SRC/main.cu not found (reference absent) -- keeping any prebuilt $OUT" >&2
  exit 0
fi
mkdir -p "$OUT"
TMP="$OUT/.scratch.$$"
mkdir -p "$TMP"
trap 'rm -rf "$TMP"' EXIT
sed 's/^#define GPU_IMPLEMENTATION 1/#define GPU_IMPLEMENTATION 0/' "$SRC/main.cu" > "$TMP/main.cu"
grep -q '^#define GPU_IMPLEMENTATION 0' "$TMP/main.cu"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"$NVCC" -std=c++14 -O2 -w -Wno-deprecated-gpu-targets -rdc=true -I"$SRC" -o "$OUT/MapReduce_cpu" \
    "$TMP/main.cu" "$SRC/KeyValue.cu" "$SRC/util.cu"
echo "build_ref: built $OUT/MapReduce_cpu"
