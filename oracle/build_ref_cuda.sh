#!/usr/bin/env bash
# oracle/build_ref_cuda.sh -- MEASUREMENT INFRASTRUCTURE.  Builds oracle/_ref/ref_cuda_legacy: the reference's own
# CUDA path (src/CUDA/cuda_impl.cu) compiled for sm_100a from the sources where they lie under /root/reference, plus
# oracle/ref_cuda_harness.cxx.  Mechanical, non-arithmetic patches applied to a temporary copy that is deleted
# afterwards (no reference source enters this repository; only the binary lands in oracle/_ref/, git-ignored):
#   * `* = 0>`  -> `* = nullptr>`   (MSVC-only non-type template default, as in build_ref.sh)
#   * namespace `cuda` -> `rcuda`   (the name collides with libcu++'s ::cuda pulled in by thrust of CUDA 12.9)
# Informational comparator only: different spp/depth/RNG/scene than the oracle.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${RTIOW_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/src/CUDA" ]; then
    echo "build_ref_cuda: $REF/src/CUDA not present; keeping prebuilt $OUT/ref_cuda_legacy (if any)" >&2
    exit 0
fi
if [ -x "$OUT/ref_cuda_legacy" ] && [ "$OUT/ref_cuda_legacy" -nt "$REF/src/CUDA/cuda_impl.cu" ] \
   && [ "$OUT/ref_cuda_legacy" -nt "$HERE/ref_cuda_harness.cxx" ] && [ "$OUT/ref_cuda_legacy" -nt "$HERE/build_ref_cuda.sh" ]; then
    echo "build_ref_cuda: $OUT/ref_cuda_legacy is up to date"
    exit 0
fi
TMP="$(mktemp -d)"
trap 'rm -rf "$TMP"' EXIT
for f in "$REF"/src/*.hxx; do
    sed 's/\* = 0>/* = nullptr>/g' "$f" > "$TMP/$(basename "$f")"
done
sed -e 's/\* = 0>/* = nullptr>/g' -e 's/namespace cuda\b/namespace rcuda/g' -e 's/\bcuda::/rcuda::/g' \
    "$REF/src/CUDA/cuda_impl.cu" > "$TMP/cuda_impl.cu"
mkdir -p "$OUT"
nvcc -std=c++17 -O3 --expt-extended-lambda --expt-relaxed-constexpr -gencode arch=compute_100a,code=sm_100a \
    -include cfloat -I"$TMP" -I"$REF/include" -w \
    "$TMP/cuda_impl.cu" -x cu "$HERE/ref_cuda_harness.cxx" -o "$OUT/ref_cuda_legacy" -cudart shared
echo "build_ref_cuda: wrote $OUT/ref_cuda_legacy"
