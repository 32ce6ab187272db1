#!/bin/sh
# Builds oracle/_ref/libiree_uk_ref.so: the reference's own CPU micro-kernels
# (runtime/src/iree/builtins/ukernel: query_tile_sizes, pack, mmt4d, unpack — generic tiles
# and the x86-64 AVX2 / AVX-512 / VNNI / BF16 variants), compiled with gcc from the sources
# where they lie under /root/reference, plus oracle/ref_glue.c (ours) which drives them as a
# row-major matmul the way builtins/ukernel/tools/e2e_matmul_benchmark.c:200-330 does.
# Nothing is copied: outputs go only to oracle/_ref/ (git-ignored, shipped to the GPU box).
# The one generated file is config_x86_64.h, which the reference's cmake would configure from
# arch/x86_64/config_x86_64.h.in (four #cmakedefine lines); we enable all four variants and
# let the run-time cpu_data bits (from cpuid, ref_glue.c) pick one.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${IREE_REFERENCE_DIR:-/root/reference}"
UK="$REF/runtime/src/iree/builtins/ukernel"
OUT="$HERE/_ref"
CC="$(command -v /usr/bin/gcc || echo gcc)"
[ -d "$UK" ] || { echo "reference ukernel sources not found under $REF" >&2; exit 1; }
mkdir -p "$OUT/gen/iree/builtins/ukernel/arch/x86_64" "$OUT/obj"
cat > "$OUT/gen/iree/builtins/ukernel/arch/x86_64/config_x86_64.h" <<'CFG'
// Written by oracle/build_ref.sh (stands in for the cmake-configured config_x86_64.h.in).
#ifndef ORACLE_REF_CONFIG_X86_64_H_
#define ORACLE_REF_CONFIG_X86_64_H_
#define IREE_UK_BUILD_X86_64_AVX2_FMA
#define IREE_UK_BUILD_X86_64_AVX512_BASE
#define IREE_UK_BUILD_X86_64_AVX512_VNNI
#define IREE_UK_BUILD_X86_64_AVX512_BF16
#endif
CFG
INC="-I$REF/runtime/src -I$OUT/gen"
BASE="-O3 -fPIC -fvisibility=hidden -std=c11 -DNDEBUG -w"
AVX2="-mavx2 -mfma -mf16c"
AVX512="-mavx512f -mavx512vl -mavx512cd -mavx512bw -mavx512dq"
OBJS=""
cc() { # cc <flags> <source>
  o="$OUT/obj/$(basename "$2" .c).o"
  "$CC" $BASE $1 $INC -c "$2" -o "$o"
  OBJS="$OBJS $o"
}
for f in mmt4d mmt4d_tile_generic pack pack_tile unpack unpack_tile query_tile_sizes; do
  cc "" "$UK/$f.c"
done
A="$UK/arch/x86_64"
for f in mmt4d pack unpack query_tile_sizes; do cc "" "$A/${f}_x86_64_entry_point.c"; done
for f in mmt4d pack unpack; do
  cc "$AVX2" "$A/${f}_x86_64_avx2_fma.c"
  cc "$AVX2 $AVX512" "$A/${f}_x86_64_avx512_base.c"
done
cc "$AVX2 $AVX512 -mavx512vnni" "$A/mmt4d_x86_64_avx512_vnni.c"
cc "$AVX2 $AVX512 -mavx512bf16" "$A/mmt4d_x86_64_avx512_bf16.c"
"$CC" $BASE -fopenmp $INC -c "$HERE/ref_glue.c" -o "$OUT/obj/ref_glue.o"
"$CC" $BASE $INC -c "$HERE/ref_math_glue.c" -o "$OUT/obj/ref_math_glue.o"
"$CC" -shared -fopenmp -o "$OUT/libiree_uk_ref.so" $OBJS "$OUT/obj/ref_glue.o" "$OUT/obj/ref_math_glue.o" -lm
echo "built $OUT/libiree_uk_ref.so"
