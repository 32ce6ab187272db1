"""linalg.mmt4d on data-tiled operands, on the GPU, against the oracle (bit-exact) and the
reference's own iree_uk_mmt4d when oracle/_ref travelled.

PENDING VALIDATION: kernels/mmt4d_kernels.cu was written after this round's GPU budget was
spent; its device code has run only on host threads under the mock driver
(tests/mock_cuda/mock_suite.py::test_mmt4d_end_to_end_on_the_host, also under ASAN). Skipped
unless IREE_B200_PENDING_VALIDATION=1; first thing next round: run, then drop the gate.
"""

import os

import numpy as np
import pytest

import oracle
from iree_b200 import hal as H

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("t", ["i8", "f16", "bf16"])
@pytest.mark.parametrize("case", [
    (37, 50, 29, 16, 16, 2, False), (70, 33, 130, 8, 8, 1, True), (64, 64, 64, 16, 16, 16, False),
    (100, 40, 33, 8, 32, 4, True), (33, 17, 40, 1, 1, 1, False), (20, 24, 90, 7, 5, 3, True),
    (1000, 528, 272, 16, 16, 2, False), (512, 1024, 768, 16, 16, 16, True),
    (2048, 2048, 2048, 16, 16, 2, False), (130, 136, 200, 8, 8, 8, True),
], ids=str)
def test_mmt4d_vs_oracle(ops, t, case):
    M, K, N, M0, N0, K0, accumulate = case
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    l4 = oracle.pack(lhs, M0, K0)
    r4 = oracle.pack(np.ascontiguousarray(rhs.T), N0, K0)
    M1, K1, N1 = l4.shape[0], l4.shape[1], r4.shape[0]
    acc_t = "i32" if t == "i8" else "f32"
    acc = oracle.generate_random_matrix(acc_t, (M1, N1, M0, N0), 3) if accumulate else None
    out = acc.copy() if accumulate else np.full((M1, N1, M0, N0), 77, dtype=np.dtype(
        np.int32 if t == "i8" else np.float32))
    before = ops.dispatch_count
    ops.mmt4d_host(t, M1, N1, K1, M0, N0, K0, l4.ctypes.data, r4.ctypes.data, out.ctypes.data,
                   flags=H.MATMUL_ACCUMULATE if accumulate else 0)
    # Tile-filling shapes take the tensor-core route: unpack lhs, rhs (, acc), GEMM, pack.
    routed = ops.mmt4d_workspace_size(t, M1, N1, K1, M0, N0, K0) != 0
    assert ops.dispatch_count - before == ((5 if accumulate else 4) if routed else 1)
    assert np.array_equal(out, oracle.mmt4d(t, l4, r4, acc=acc))
    if oracle.ref_available():
        assert np.array_equal(out, oracle.ref_mmt4d(t, l4, r4, acc=acc))
    if not accumulate:
        assert np.array_equal(oracle.unpack(out, M, N), oracle.matmul(t, lhs[None], rhs[None])[0])


def test_mmt4d_edge_cases(ops):
    t, M0, N0, K0 = "bf16", 16, 16, 2
    out = np.full((2, 1, M0, N0), 5, dtype=np.float32)
    ops.mmt4d_host(t, 2, 1, 0, M0, N0, K0, 0, 0, out.ctypes.data)
    assert (out == 0).all()
    out[:] = 5
    before = ops.dispatch_count
    ops.mmt4d_host(t, 2, 1, 0, M0, N0, K0, 0, 0, out.ctypes.data, flags=H.MATMUL_ACCUMULATE)
    assert ops.dispatch_count == before and (out == 5).all()
    with pytest.raises(H.Status):
        ops.mmt4d_host(t, 2, 1, 4, 0, N0, K0, 1, 1, out.ctypes.data)
    with pytest.raises(H.Status):
        ops.mmt4d_host("f32", 2, 1, 4, M0, N0, K0, 1, 1, out.ctypes.data)
