"""iree_b200_ops_matmul_bias_host through the three-stream pipeline, on the GPU.

PENDING VALIDATION: the pipelined form of the fused-bias host call was written after this
round's GPU budget was spent. Its whole data path (bias staging, per-block bias slices, panel
uploads, packed result blocks) is checked bit for bit on the CPU with the generic kernel under
the mock driver (tests/mock_cuda/mock_suite.py::test_pipelined_host_call_end_to_end_on_the_host
and ::test_pipelined_bias_call_hands_each_block_its_bias_columns, also under ASAN); the tcgen05
fused exports it dispatches were validated on the GPU as whole-problem dispatches
(tests/test_matmul_bias_gpu.py). What has not run is the two together. Skipped unless
IREE_B200_PENDING_VALIDATION=1; first thing next round: run, then fold into
test_matmul_bias_gpu.py.
"""

import os

import numpy as np
import pytest

import oracle
from iree_b200 import hal as H

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("t", ["i8", "f16", "bf16"])
@pytest.mark.parametrize("case", [
    # batch, M, K, N, accumulate, nt, truncate, relu, block
    (1, 1000, 272, 528, True, False, False, True, 256),
    (1, 528, 144, 1008, False, True, False, False, 256),
    (3, 300, 528, 272, False, False, False, True, 256),
    (2, 513, 64, 768, False, False, True, True, 256),
    (1, 1536, 144, 1280, False, True, True, False, 512),
], ids=str)
def test_pipelined_bias_host_call(ops, t, case):
    batch, M, K, N, accumulate, nt, truncate, relu, block = case
    if truncate and t == "i8":
        pytest.skip("TRUNCATE_RESULT is f16 / bf16 only")
    acc_t = "i32" if t == "i8" else "f32"
    lhs = oracle.generate_random_matrix(t, (batch, M, K), 1)
    rhs = oracle.generate_random_matrix(t, (batch, N, K) if nt else (batch, K, N), 2)
    acc = oracle.generate_random_matrix(acc_t, (batch, M, N), 3) if accumulate else None
    bias = oracle.generate_random_matrix(acc_t, (N,), 4)
    if truncate:
        out = np.full((batch, M, N), 0x7777, dtype=np.uint16)
    else:
        out = acc.copy() if accumulate else np.full((batch, M, N), 77, dtype=bias.dtype)
    flags = ((H.MATMUL_ACCUMULATE if accumulate else 0) | (H.MATMUL_TRANSPOSE_RHS if nt else 0)
             | (H.MATMUL_TRUNCATE_RESULT if truncate else 0) | (H.MATMUL_RELU if relu else 0))
    before = ops.dispatch_count
    ops.set_host_pipeline(0, block, block)
    try:
        ops.matmul_bias_host(t, batch, M, N, K, lhs.ctypes.data, rhs.ctypes.data, bias.ctypes.data,
                             out.ctypes.data, flags)
    finally:
        ops.set_host_pipeline(16 << 20, 2048, 2048)
    assert ops.dispatch_count - before == batch * -(-M // block) * -(-N // block)
    assert np.array_equal(out, oracle.matmul_bias(t, lhs, rhs, bias, acc=acc, transpose_rhs=nt,
                                                  relu=relu, truncate=truncate))


def test_linear_layer_at_bench_size_matches_the_serial_form(ops):
    """8192 x 4096 x 4096 bf16 linear (above the default threshold: pipelined) against the same
    call with the pipeline off."""
    M, K, N = 8192, 4096, 4096
    rng = np.random.Generator(np.random.PCG64(7))
    lhs = rng.integers(-2, 3, (M, K)).astype(np.float32)
    rhs = rng.integers(-2, 3, (K, N)).astype(np.float32)
    lhs, rhs = oracle.from_f32("bf16", lhs), oracle.from_f32("bf16", rhs)
    bias = rng.standard_normal(N, dtype=np.float32)
    outs = []
    for min_bytes in (16 << 20, 1 << 62):
        out = np.zeros((M, N), dtype=np.uint16)
        ops.set_host_pipeline(min_bytes, 2048, 2048)
        try:
            ops.matmul_bias_host("bf16", 1, M, N, K, lhs.ctypes.data, rhs.ctypes.data,
                                 bias.ctypes.data, out.ctypes.data,
                                 H.MATMUL_RELU | H.MATMUL_TRUNCATE_RESULT)
        finally:
            ops.set_host_pipeline(16 << 20, 2048, 2048)
        outs.append(out)
    assert np.array_equal(outs[0], outs[1])
