#!/usr/bin/env bash
# tools/build_dropin.sh -- proof that the backend drops into the reference's main.cxx:
# compiles the reference's OWN src/main.cxx (with its own headers, from a sed-patched temporary copy: the
# MSVC-only `* = 0` template default, see oracle/build_ref.sh) and links it against librtiow_shim.so, whose
# cuda_impl has the reference's exact signature.  Output: raytracinginoneweekend_b200/_build/reference_main_dropin
# (git-ignored; travels to the GPU box).  Running it writes image_cuda.ppm like the reference's main().
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
REF="${RTIOW_REFERENCE:-/root/reference}"
PKG="$HERE/raytracinginoneweekend_b200"
if [ ! -d "$REF/src" ]; then
    echo "build_dropin: $REF/src not present; keeping prebuilt binary (if any)" >&2
    exit 0
fi
TMP="$(mktemp -d)"
trap 'rm -rf "$TMP"' EXIT
for f in "$REF"/src/*.hxx "$REF"/src/main.cxx; do
    sed 's/\* = 0>/* = nullptr>/g' "$f" > "$TMP/$(basename "$f")"
done
mkdir -p "$PKG/_build"
g++ -std=c++20 -O2 -ffp-contract=off -include cfloat -I"$TMP" "$TMP/main.cxx" -o "$PKG/_build/reference_main_dropin" \
    -L"$PKG" -lrtiow_shim -lrtiow_b200 -Wl,-rpath,"$PKG" -Wl,-rpath,'$ORIGIN/..' -pthread
echo "build_dropin: wrote $PKG/_build/reference_main_dropin"
