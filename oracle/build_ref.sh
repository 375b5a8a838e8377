#!/usr/bin/env bash
# TEST INFRASTRUCTURE — builds the compiled reference (oracle/_ref/libinkref.so).
#
# Compiles the reference's own hot-path sources from where they lie under /root/reference,
# through a scratch copy in a temp dir (NOT in this repo), with a Qt-free C-ABI shim
# (oracle/ref_shim.cpp).  Only edit applied to the scratch copy: the three grid-size constants
# in src/system.h:15-17 become `inline int` so one .so serves all grid sizes.  `-DNDEBUG` is
# mandatory: with asserts on the reference aborts in step 0 (src/system.h:65 via
# src/systemHelpers.cpp:78).  No FMA contraction (x86-64 baseline has no FMA, and we say so
# explicitly with -ffp-contract=off).
#
# Usage: oracle/build_ref.sh [REFERENCE_ROOT]   (default /root/reference)
# Exits 0 with a message if the reference is absent (e.g. on the GPU box: the prebuilt .so travels).
set -euo pipefail
REF="${1:-/root/reference}"
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/_ref"
if [ ! -f "$REF/src/system.h" ]; then
    echo "build_ref: $REF not present; keeping prebuilt $OUT (if any)"
    exit 0
fi
mkdir -p "$OUT"
if [ -f "$OUT/libinkref.so" ] && [ "$OUT/libinkref.so" -nt "$HERE/ref_shim.cpp" ] && [ "$OUT/libinkref.so" -nt "$HERE/build_ref.sh" ]; then
    echo "build_ref: $OUT/libinkref.so up to date"
    exit 0
fi
TMP="$(mktemp -d /tmp/inkref.XXXXXX)"
trap 'rm -rf "$TMP"' EXIT
cp "$REF"/src/system.h "$REF"/src/system.cpp "$REF"/src/systemHelpers.cpp "$REF"/src/waterSolver.cpp \
   "$REF"/src/inkSolver.cpp "$REF"/src/inkSim.h "$REF"/src/inkSim.cpp "$TMP"/
sed -i -E 's/^const int (WATERGRID_[XYZ]) *=/inline int \1 =/' "$TMP/system.h"
grep -c '^inline int WATERGRID_' "$TMP/system.h" | grep -qx 3 || { echo "build_ref: sed did not patch 3 lines"; exit 1; }
g++ -std=c++20 -O2 -DNDEBUG -ffp-contract=off -fPIC -shared -w -Wl,-Bsymbolic \
    -I"$REF" -I"$TMP" \
    "$HERE/ref_shim.cpp" "$TMP/system.cpp" "$TMP/systemHelpers.cpp" "$TMP/waterSolver.cpp" \
    "$TMP/inkSolver.cpp" "$TMP/inkSim.cpp" \
    -o "$OUT/libinkref.so"
echo "build_ref: built $OUT/libinkref.so"
