"""mx_kernels.cu's device code executed on the host (tests/host_shims/mx_kernels_host.cc:
CUDA keywords shimmed, 256 OS threads per block, __syncthreads() = a pthread barrier) and
compared bit for bit with the oracle. The kernel has not run on a GPU yet; this is what
stands behind it until it does."""

import ctypes
import os
import shutil
import subprocess

import numpy as np
import pytest

import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def shim(tmp_path_factory):
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else shutil.which("g++")
    out = str(tmp_path_factory.mktemp("mx") / "libmx_host.so")
    subprocess.run([cxx, "-O1", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-x", "c++",
                    "-o", out, os.path.join(ROOT, "tests", "host_shims", "mx_kernels_host.cc"),
                    "-lpthread"], check=True)
    return ctypes.CDLL(out)


@pytest.mark.parametrize("shape", [(1, 32, 1, False), (70, 96, 50, True), (64, 256, 130, False),
                                   (129, 64, 65, True)], ids=str)
def test_mxfp4_device_code_on_host_threads(shim, shape):
    M, K, N, accumulate = shape
    lhs = oracle.generate_random_matrix("u8", (M, K // 2), 1)
    ls = oracle.generate_random_matrix("f8e8m0fnu", (M, K // 32), 2)
    rhs = oracle.generate_random_matrix("u8", (N, K // 2), 3)
    rs = oracle.generate_random_matrix("f8e8m0fnu", (N, K // 32), 4)
    ls[0, 0] = 0        # 2^-127
    if M > 1:
        ls[M - 1, K // 32 - 1] = 0xFF   # NaN
    acc = oracle.generate_random_matrix("f32", (M, N), 5) if accumulate else None
    out = acc.copy() if accumulate else np.full((M, N), 77, dtype=np.float32)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    shim.mx_host_launch(p(lhs), p(ls), p(rhs), p(rs), p(out), M, N, K, 1 if accumulate else 0)
    assert np.array_equal(out, oracle.matmul_mxfp4(lhs, ls, rhs, rs, acc=acc), equal_nan=True)
