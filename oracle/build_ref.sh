#!/usr/bin/env bash
# oracle/build_ref.sh -- TEST INFRASTRUCTURE.  Builds oracle/_ref/libref_oracle.so from the
# reference sources where they lie under /root/reference (read-only).  The reference headers
# need one mechanical, non-arithmetic patch to compile with g++ (MSVC-only `* = 0` non-type
# template default -> `* = nullptr`, 21 sites); the patched copy lives in a temporary
# directory that is deleted afterwards -- no reference source enters this repository, only the
# compiled library lands in oracle/_ref/ (git-ignored, travels to the GPU box).
#
#   -O2 -ffp-contract=off : no FMA contraction, so the bit-exact primary-hit check is meaningful.
#   -include cfloat       : the reference uses FLT_MIN/FLT_MAX without including <cfloat>.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${RTIOW_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/src" ]; then
    echo "build_ref: $REF/src not present; keeping prebuilt $OUT (if any)" >&2
    exit 0
fi
TMP="$(mktemp -d)"
trap 'rm -rf "$TMP"' EXIT
for f in "$REF"/src/*.hxx "$REF"/src/main.cxx; do
    sed 's/\* = 0>/* = nullptr>/g' "$f" > "$TMP/$(basename "$f")"
done
mkdir -p "$OUT"
g++ -std=c++20 -O2 -ffp-contract=off -fPIC -shared -pthread -include cfloat \
    -I"$TMP" "$HERE/ref_harness.cxx" -o "$OUT/libref_oracle.so"
echo "build_ref: wrote $OUT/libref_oracle.so"
