#!/bin/bash
# Builds the launch layer's host sources + tools/fuzz_cda1.cc with ASAN + UBSAN and runs the
# mutation fuzzer over the CDA1 reader (no GPU needed).  Usage: tools/fuzz_cda1.sh [iterations]
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/iree_b200/_build
python -m iree_b200.build > /dev/null
CXX=/usr/bin/g++; [ -x $CXX ] || CXX=g++
SRC="$ROOT/iree_b200/csrc/hal/device.cc $ROOT/iree_b200/csrc/hal/executable.cc \
     $ROOT/iree_b200/csrc/hal/command_buffer.cc $ROOT/iree_b200/csrc/hal/channel.cc"
$CXX -O1 -g -std=c++17 -fsanitize=address,undefined -fno-sanitize-recover=undefined \
  -I $ROOT/include -I /usr/local/cuda/include -o $OUT/fuzz_cda1 $ROOT/tools/fuzz_cda1.cc $SRC \
  -ldl -lpthread
# Seed: the builtin executable's export table with the module images cut down to stubs, so
# the mutations land on tables, vtables, strings and vectors instead of cubin bytes.
python - "$OUT" <<'PY'
import sys
from iree_b200 import cda1
out = sys.argv[1]
exe = cda1.deserialize(open(out + "/iree_b200_kernels.cda1", "rb").read())
exe.modules = [b"\x7fELF" + bytes([i]) * 11 for i in range(len(exe.modules))]
open(out + "/fuzz_seed.cda1", "wb").write(cda1.serialize(exe))
PY
ASAN_OPTIONS=detect_leaks=1 $OUT/fuzz_cda1 $OUT/fuzz_seed.cda1 ${1:-200000}
