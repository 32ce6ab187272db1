#!/usr/bin/env bash
# Builds the REFERENCE runtime from /root/reference (CPU box, offline, ~1-4 min on 8 cores) in
# two flavours and leaves binaries only under oracle/_ref/iree-rt/ (git-ignored, shipped to
# the GPU box by gpurun):
#   stock/  the unmodified checkout: iree-benchmark-executable, iree-run-module,
#           iree-benchmark-module, iree-as-module (the VM assembler: iree_b200/vmasm.py turns our
#           kernels into a real .vmfb with it), iree-dump-module, plus our two
#           checkers linked against it (integration/ref_runtime/): cda1_ref_verify (the
#           reference's flatcc verifier + header parser on a CDA1 file, CPU only) and
#           ref_hal_harness (one dispatch through iree_hal_* with file-backed bindings).
#   tma/    the same tools from a symlink copy of the checkout with
#           integration/cuda_driver_tma.patch applied to hal/drivers/cuda (the TMA hook of
#           INTEGRATION.md §3) — nothing under /root/reference is written.
# SURVEY.md §0 records the cmake flag set; integration/ref_runtime/CMakeLists.txt carries it.
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
REF="${IREE_B200_REFERENCE_DIR:-/root/reference}"
WORK="${IREE_B200_REF_BUILD_DIR:-/tmp/iree_b200_ref_rt}"
OUT="$ROOT/oracle/_ref/iree-rt"
[ -d "$REF/runtime/src/iree/hal/drivers/cuda" ] || { echo "no reference checkout at $REF"; exit 3; }
TARGETS="ref_hal_harness cda1_ref_verify iree-benchmark-executable iree-run-module iree-as-module iree-benchmark-module iree-dump-module"

build() {  # <flavour> <source dir>
  local flavour="$1" src="$2" b="$WORK/$1"
  mkdir -p "$b"
  cmake -G Ninja -S "$ROOT/integration/ref_runtime" -B "$b" -DCMAKE_BUILD_TYPE=Release \
        -DIREE_B200_REFERENCE_DIR="$src" > "$b/cmake.log" 2>&1 || { tail -30 "$b/cmake.log"; exit 1; }
  ninja -C "$b" $TARGETS > "$b/ninja.log" 2>&1 || { tail -40 "$b/ninja.log"; exit 1; }
  mkdir -p "$OUT/$flavour"
  cp "$b/ref_hal_harness" "$b/cda1_ref_verify" "$b/iree/tools/iree-benchmark-executable" \
     "$b/iree/tools/iree-run-module" "$b/iree/tools/iree-as-module" \
     "$b/iree/tools/iree-benchmark-module" "$b/iree/tools/iree-dump-module" "$OUT/$flavour/"
  strip "$OUT/$flavour"/* || true
}

build stock "$REF"

# Patched flavour: directories real, files symlinked, the patched files real copies.
PSRC="$WORK/patched_src"
rm -rf "$PSRC"
cp -as "$REF" "$PSRC"
DRV="$PSRC/runtime/src/iree/hal/drivers/cuda"
for f in $(grep '^+++ ' "$ROOT/integration/cuda_driver_tma.patch" | sed 's|^+++ b/||; s|\t.*||'); do
  cp --remove-destination "$REF/$f" "$PSRC/$f"; chmod u+w "$PSRC/$f"
done
patch -s -p1 -d "$PSRC" < "$ROOT/integration/cuda_driver_tma.patch"
cp "$ROOT/integration/iree_b200_tma_hook.h" "$ROOT/include/iree_b200_tma_recipe.h" "$DRV/"
sed -i 's|#include "iree_b200_tma_recipe.h"|#include "iree/hal/drivers/cuda/iree_b200_tma_recipe.h"|' "$DRV/iree_b200_tma_hook.h"
build tma "$PSRC"
echo "built $(ls "$OUT"/*/* | tr '\n' ' ')"
