#!/bin/bash
# Builds the host side of libiree_b200.so with ASAN + UBSAN, then (1) runs tools/sanitize_harness.cc
# against it with leak detection and (2) runs the CPU-side C-ABI tests (tests/test_cabi.py,
# tests/test_selection_cpu.py) against it under LD_PRELOADed sanitizer runtimes and (3) runs the
# launch layer's host logic (command buffers in both kinds, semaphores, executable loading,
# the ops layer's dispatch recording) against it on top of the mock CUDA driver
# (tests/mock_cuda/). No GPU needed.
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
$CXX -O1 -g -std=c++17 -fPIC -shared -Wl,-soname,libcuda.so.1 -I /usr/local/cuda/include \
  -o $T/libcuda.so.1 tests/mock_cuda/mock_cuda.cc -ldl
# The kernels' host builds are instrumented too: an out-of-bounds access by a SIMT kernel's own
# code (ragged tiles, sub-range bindings, packed nibbles) aborts like under compute-sanitizer.
$CXX -O1 -g -std=c++17 -ffp-contract=off -fPIC -shared -fsanitize=address,undefined -x c++ \
  -I /usr/local/cuda/include -o $T/libmock_kernels.so tests/mock_cuda/mock_kernels_generic.cc tests/mock_cuda/mock_kernels_mx.cc tests/mock_cuda/mock_kernels_mmt4d.cc -lpthread
cat > $T/run_mock.py <<PY
import sys
sys.path.insert(0, "$ROOT")
from iree_b200 import build as B
B.lib_path = lambda: "$T/libiree_b200_asan.so"
import iree_b200.hal as H
H._build.lib_path = B.lib_path
assert H.lib()._name == "$T/libiree_b200_asan.so"
sys.path.insert(0, "$ROOT/tests")
from test_hal_mock_cpu import HOST_LOGIC_TESTS
import pytest
rc = pytest.main(["-x", "-q", "-m", "gpu", "-p", "no:cacheprovider", "tests/test_hal_gpu.py",
                  "-k", HOST_LOGIC_TESTS])
rc = rc or pytest.main(["-x", "-q", "-m", "gpu", "-p", "no:cacheprovider",
                        "tests/mock_cuda/mock_suite.py"])
sys.exit(rc)
PY
LD_LIBRARY_PATH=$T:$LD_LIBRARY_PATH MOCK_CUDA_KERNELS=$T/libmock_kernels.so \
  LD_PRELOAD="$(gcc -print-file-name=libasan.so) $(gcc -print-file-name=libubsan.so)" \
  ASAN_OPTIONS=detect_leaks=0:allocator_may_return_null=1 python $T/run_mock.py
rm -rf $T
