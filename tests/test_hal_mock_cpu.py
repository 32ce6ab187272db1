"""The launch layer's host logic on a machine without a GPU: a mock libcuda.so.1
(tests/mock_cuda/mock_cuda.cc — memory is host memory, copies and fills run immediately,
graphs replay their nodes, launches are logged, and the SIMT kernels — generic matmuls,
MXFP4 — run as their own source compiled for the host) is put in front of a pytest
subprocess, which then runs (1) the tests of tests/test_hal_gpu.py that do not depend on a
tcgen05 kernel's result and (2) tests/mock_cuda/mock_suite.py, which checks what reaches
cuLaunchKernel / cuTensorMapEncodeTiled and runs generic / fused-bias / MXFP4 dispatches
end to end against the oracle. Skipped where a real GPU exists: there the GPU
suite runs against the real driver."""

import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

HOST_LOGIC_TESTS = ("allocator_and_transfers or host_local or imported_buffers or fill_buffer or "
                    "update_buffer or copy_buffer or execution_barriers or argument_checks or "
                    "recording_state_machine or keeps_its_resources or device_queries or "
                    "semaphores_and_queue_ordering or semaphore_host_signal or queue_alloca or "
                    "dispatch_constants_and_binding or "
                    "dispatch_reuse or executable_cache")


def _have_gpu():
    return os.path.exists("/dev/nvidia0") or os.path.exists("/dev/nvidiactl")


@pytest.fixture(scope="module")
def mock_env(tmp_path_factory):
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else shutil.which("g++")
    out = tmp_path_factory.mktemp("mock_cuda")
    subprocess.run([cxx, "-O1", "-g", "-std=c++17", "-fPIC", "-shared", "-Wl,-soname,libcuda.so.1",
                    "-I", "/usr/local/cuda/include", "-o", str(out / "libcuda.so.1"),
                    os.path.join(ROOT, "tests", "mock_cuda", "mock_cuda.cc"), "-ldl"], check=True)
    # The SIMT kernels' own source, compiled for the host: launches of these kernels run.
    mock = os.path.join(ROOT, "tests", "mock_cuda")
    subprocess.run([cxx, "-O1", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-x", "c++",
                    "-I", "/usr/local/cuda/include", "-o", str(out / "libmock_kernels.so"),
                    os.path.join(mock, "mock_kernels_generic.cc"),
                    os.path.join(mock, "mock_kernels_mx.cc"),
                    os.path.join(mock, "mock_kernels_mmt4d.cc"), "-lpthread"], check=True)
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = str(out) + os.pathsep + env.get("LD_LIBRARY_PATH", "")
    env["MOCK_CUDA_KERNELS"] = str(out / "libmock_kernels.so")
    return env


def _run(env, *args):
    p = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-p", "no:cacheprovider",
                        "-m", "gpu", *args], cwd=ROOT, env=env, capture_output=True, text=True,
                       timeout=900)
    assert p.returncode == 0, p.stdout[-4000:] + p.stderr[-2000:]
    return p.stdout


@pytest.mark.skipif(_have_gpu(), reason="a real GPU is present: the GPU suite covers this")
def test_hal_host_logic_under_the_mock_driver(mock_env):
    out = _run(mock_env, "tests/test_hal_gpu.py", "-k", HOST_LOGIC_TESTS)
    passed = int(out.strip().splitlines()[-1].split(" passed")[0].split()[-1])
    assert passed >= 40, out      # 20 tests x {stream, graph}


@pytest.mark.skipif(_have_gpu(), reason="a real GPU is present: the GPU suite covers this")
def test_dispatches_as_seen_by_the_driver(mock_env):
    out = _run(mock_env, "tests/mock_cuda/mock_suite.py")
    assert " passed" in out and "failed" not in out


@pytest.mark.skipif(_have_gpu(), reason="a real GPU is present: the GPU suite covers this")
def test_ops_guards_and_selection_under_the_mock_driver(mock_env):
    """The GPU suites' guard / selection / empty-shape tests need a device but no kernel."""
    out = _run(mock_env, "tests/test_matmul_gpu.py", "tests/test_attention_gpu.py",
               "tests/test_matmul_bias_gpu.py", "-k", "guards or selection or empty_and_zero_k")
    assert "6 passed" in out, out
