#!/bin/bash
# Builds the host side of libiree_b200.so with ASAN + UBSAN, then (1) runs tools/sanitize_harness.cc
# against it with leak detection and (2) runs the CPU-side C-ABI tests (tests/test_cabi.py,
# tests/test_selection_cpu.py) against it under LD_PRELOADed sanitizer runtimes. No GPU needed.
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd); cd "$ROOT"
python -m iree_b200.build > /dev/null
CXX=/usr/bin/g++; [ -x $CXX ] || CXX=g++
T=$(mktemp -d)
$CXX -O1 -g -std=c++17 -fPIC -fsanitize=address,undefined -fno-omit-frame-pointer -I include \
  -I /usr/local/cuda/include -shared -o $T/libiree_b200_asan.so iree_b200/csrc/hal/device.cc \
  iree_b200/csrc/hal/executable.cc iree_b200/csrc/hal/command_buffer.cc iree_b200/csrc/hal/channel.cc \
  iree_b200/csrc/ops/ops.cc iree_b200/_build/builtin_blob.o -ldl -lpthread
$CXX -O1 -g -std=c++17 -fsanitize=address,undefined -I include -I /usr/local/cuda/include \
  -o $T/harness tools/sanitize_harness.cc $T/libiree_b200_asan.so -Wl,-rpath,$T
ASAN_OPTIONS=detect_leaks=1 $T/harness
cat > $T/run.py <<PY
import sys
sys.path.insert(0, "$ROOT")
from iree_b200 import build as B
B.lib_path = lambda: "$T/libiree_b200_asan.so"
import iree_b200.hal as H
H._build.lib_path = B.lib_path
assert H.lib()._name == "$T/libiree_b200_asan.so"
import pytest
sys.exit(pytest.main(["-x", "-q", "tests/test_selection_cpu.py", "tests/test_cabi.py",
                      "-p", "no:cacheprovider"]))
PY
LD_PRELOAD="$(gcc -print-file-name=libasan.so) $(gcc -print-file-name=libubsan.so)" \
  ASAN_OPTIONS=detect_leaks=0 python $T/run.py
rm -rf $T
