"""linalg.mmt4d on data-tiled operands, on the GPU, against the oracle (bit-exact) and the
reference's own iree_uk_mmt4d when oracle/_ref travelled.

First GPU run: round 2 (profiles/mmt4d_mxfp4_first_gpu_run_r2.json), bit-exact; ungated since.
"""

import os

import numpy as np
import pytest

import oracle
from iree_b200 import hal as H

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("t", ["i8", "f16", "bf16", "f32"])
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
    # Tile-filling shapes take a tensor-core route: tiles read directly (one dispatch) when
    # the tile shape allows, else unpack lhs, rhs (, acc), GEMM, pack.
    routed = ops.mmt4d_workspace_size(t, M1, N1, K1, M0, N0, K0) != 0
    direct = _direct(t, M1, N1, K1, M0, N0, K0)
    assert ops.dispatch_count - before == (1 if direct else (5 if accumulate else 4) if routed else 1)
    assert np.array_equal(out, oracle.mmt4d(t, l4, r4, acc=acc))
    if oracle.ref_available():
        assert np.array_equal(out, oracle.ref_mmt4d(t, l4, r4, acc=acc))
    if not accumulate:
        want_mm = (lhs.astype(np.float64) @ rhs.astype(np.float64)).astype(np.float32) if t == "f32" \
            else oracle.matmul(t, lhs[None], rhs[None])[0]
        assert np.array_equal(oracle.unpack(out, M, N), want_mm)


def _direct(t, M1, N1, K1, M0, N0, K0):
    """iree_b200_ops_record_mmt4d's guard for the tile-direct tcgen05 export."""
    es = {"i8": 1, "f16": 2, "bf16": 2}.get(t)
    return (es is not None and K0 * es == 16 and 128 % M0 == 0 and 128 % N0 == 0 and M0 % 8 == 0
            and N0 % 8 == 0 and min(M1 * M0, N1 * N0, K1 * K0) >= 128)


@pytest.mark.parametrize("t", ["bf16", "f16", "i8"])
@pytest.mark.parametrize("case", [
    # M, K, N (logical), M0, N0, accumulate   (K0 = 16 bytes / element size)
    (128, 128, 128, 16, 16, False), (1000, 528, 272, 16, 16, False), (300, 256, 520, 8, 32, True),
    (257, 1032, 260, 32, 64, False), (2048, 256, 2560, 16, 16, True), (4096, 512, 1024, 128, 8, False),
    (520, 384, 1000, 64, 128, True),
], ids=str)
def test_mmt4d_tiles_read_directly_by_the_tcgen05_gemm(ops, t, case):
    """iree_b200_mmt4d_*: rank-4 tensor maps walk the tiled operands / result, one dispatch, no
    workspace; bit-exact against the oracle, the reference's iree_uk_mmt4d and the SIMT kernel."""
    M, K, N, M0, N0, accumulate = case
    K0 = 16 if t == "i8" else 8
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    l4 = oracle.pack(lhs, M0, K0)
    r4 = oracle.pack(np.ascontiguousarray(rhs.T), N0, K0)
    M1, K1, N1 = l4.shape[0], l4.shape[1], r4.shape[0]
    assert _direct(t, M1, N1, K1, M0, N0, K0)
    acc_t = "i32" if t == "i8" else "f32"
    acc = oracle.generate_random_matrix(acc_t, (M1, N1, M0, N0), 3) if accumulate else None
    dt = np.int32 if t == "i8" else np.float32
    flags = H.MATMUL_ACCUMULATE if accumulate else 0
    out = acc.copy() if accumulate else np.full((M1, N1, M0, N0), 77, dtype=dt)
    before = ops.dispatch_count
    ops.mmt4d_host(t, M1, N1, K1, M0, N0, K0, l4.ctypes.data, r4.ctypes.data, out.ctypes.data,
                   flags=flags)
    assert ops.dispatch_count - before == 1
    want = oracle.mmt4d(t, l4, r4, acc=acc)
    assert np.array_equal(out, want)
    if oracle.ref_available():
        assert np.array_equal(out, oracle.ref_mmt4d(t, l4, r4, acc=acc))
    simt = acc.copy() if accumulate else np.zeros((M1, N1, M0, N0), dtype=dt)
    ops.mmt4d_host(t, M1, N1, K1, M0, N0, K0, l4.ctypes.data, r4.ctypes.data, simt.ctypes.data,
                   flags=flags | H.MATMUL_FORCE_GENERIC)
    assert np.array_equal(simt, want)


@pytest.mark.parametrize("t", ["f16", "bf16", "i8"])
def test_tensor_route_vs_simt_on_random_data(ops, t):
    """Non-integer operands (N(0,1), where f32 partial sums round): the routed result (relayout
    -> tcgen05 GEMM -> relayout) against the SIMT kernel's k-ascending result. i8 stays
    bit-exact; f16 / bf16 differ by f32 summation order only: <= 1e-5 of the result's scale."""
    M, K, N, M0, N0, K0 = 512, 1024, 384, 16, 16, 2
    g = np.random.Generator(np.random.PCG64(0xB200))
    if t == "i8":
        lhs = g.integers(-128, 128, (M, K), dtype=np.int8)
        rhs = g.integers(-128, 128, (K, N), dtype=np.int8)
    else:
        lhs = oracle.from_f32(t, g.standard_normal((M, K)).astype(np.float32))
        rhs = oracle.from_f32(t, g.standard_normal((K, N)).astype(np.float32))
    l4 = oracle.pack(lhs, M0, K0)
    r4 = oracle.pack(np.ascontiguousarray(rhs.T), N0, K0)
    M1, K1, N1 = l4.shape[0], l4.shape[1], r4.shape[0]
    assert ops.mmt4d_workspace_size(t, M1, N1, K1, M0, N0, K0) != 0
    dt = np.int32 if t == "i8" else np.float32
    routed, simt = np.empty((M1, N1, M0, N0), dt), np.empty((M1, N1, M0, N0), dt)
    before = ops.dispatch_count
    ops.mmt4d_host(t, M1, N1, K1, M0, N0, K0, l4.ctypes.data, r4.ctypes.data, routed.ctypes.data)
    assert ops.dispatch_count - before == 4
    ops.mmt4d_host(t, M1, N1, K1, M0, N0, K0, l4.ctypes.data, r4.ctypes.data, simt.ctypes.data,
                   flags=H.MATMUL_FORCE_GENERIC)
    assert ops.dispatch_count - before == 5
    want = oracle.mmt4d(t, l4, r4)
    assert np.array_equal(simt, want)              # the SIMT kernel IS the reference order
    if t == "i8":
        assert np.array_equal(routed, want)
    else:
        scale = np.abs(want).max()
        assert np.abs(routed - want).max() <= 1e-5 * scale


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
        ops.mmt4d_host("f8e4m3fn", 2, 1, 4, M0, N0, K0, 1, 1, out.ctypes.data)
