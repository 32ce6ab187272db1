"""MXFP4 matmul on the GPU against the oracle (bit-exact on the reference generator's inputs).

Three routes (iree_b200_ops_matmul_mxfp4_host picks; flags force): the block-scaled tcgen05
kernel (kind::mxf4.block_scale; K % 256 == 0, N % 4 == 0), the exact bf16 expansion + bf16
tcgen05 GEMM (FLAG_MX_EXPAND; tile-filling shapes), the SIMT kernel (FLAG_FORCE_GENERIC; any
shape). First GPU run of the SIMT / expansion routes: round 2
(profiles/mmt4d_mxfp4_first_gpu_run_r2.json).
"""

import os

import numpy as np
import pytest

import oracle
from iree_b200 import hal as H

pytestmark = pytest.mark.gpu


def _inputs(M, K, N, accumulate):
    lhs = oracle.generate_random_matrix("u8", (M, K // 2), 1)
    ls = oracle.generate_random_matrix("f8e8m0fnu", (M, K // 32), 2)
    rhs = oracle.generate_random_matrix("u8", (N, K // 2), 3)
    rs = oracle.generate_random_matrix("f8e8m0fnu", (N, K // 32), 4)
    acc = oracle.generate_random_matrix("f32", (M, N), 5) if accumulate else None
    return lhs, ls, rhs, rs, acc


@pytest.mark.parametrize("shape", [(1, 32, 1), (3, 64, 5), (70, 96, 50), (64, 256, 130),
                                   (512, 1024, 384), (1000, 32, 8),
                                   # tensor-core route (bf16 expansion + tcgen05 GEMM)
                                   (128, 128, 128), (1000, 512, 264), (2048, 2048, 2048)], ids=str)
@pytest.mark.parametrize("accumulate", [False, True], ids=["", "acc"])
def test_mxfp4_matmul_vs_oracle(ops, shape, accumulate):
    M, K, N = shape
    lhs, ls, rhs, rs, acc = _inputs(M, K, N, accumulate)
    out = acc.copy() if accumulate else np.full((M, N), 77, dtype=np.float32)
    before = ops.dispatch_count
    ops.matmul_mxfp4_host(M, N, K, lhs.ctypes.data, ls.ctypes.data, rhs.ctypes.data,
                          rs.ctypes.data, out.ctypes.data,
                          flags=H.MATMUL_ACCUMULATE if accumulate else 0)
    # Block-scaled kernel: two scale repacks + the GEMM; bf16-expansion route: 3; SIMT: 1.
    block_scaled = ops.matmul_mxfp4_scale_workspace_size(M, N, K) != 0
    expected = 3 if block_scaled else (3 if ops.matmul_mxfp4_workspace_size(M, N, K) else 1)
    assert ops.dispatch_count - before == expected
    assert np.array_equal(out, oracle.matmul_mxfp4(lhs, ls, rhs, rs, acc=acc))


@pytest.mark.parametrize("shape", [(128, 256, 256), (129, 256, 4), (1000, 512, 264), (300, 768, 520),
                                   (128, 2048, 256), (257, 1024, 1028), (2048, 2048, 2048),
                                   (4096, 4096, 1536)], ids=str)
@pytest.mark.parametrize("accumulate", [False, True], ids=["", "acc"])
def test_block_scaled_kernel_and_the_other_routes_agree_with_the_oracle(ops, shape, accumulate):
    """tcgen05.mma.kind::mxf4.block_scale (iree_b200_matmul_mxfp4_f32) against the oracle, bit
    for bit on the reference generator's inputs, ragged M / N tiles, many K stages; and the
    same call forced onto the expansion route and the SIMT kernel."""
    M, K, N = shape
    lhs, ls, rhs, rs, acc = _inputs(M, K, N, accumulate)
    want = oracle.matmul_mxfp4(lhs, ls, rhs, rs, acc=acc)
    base = H.MATMUL_ACCUMULATE if accumulate else 0
    for route, flag, dispatches in (("block_scaled", 0, 3),
                                    ("block_scaled_1cta", H.MATMUL_FORCE_1CTA, 3),
                                    ("block_scaled_2cta", H.MATMUL_FORCE_2CTA, 3),
                                    ("expand", H.MATMUL_MX_EXPAND, None),
                                    ("simt", H.MATMUL_FORCE_GENERIC, 1)):
        if route == "simt" and M * N * K > 1 << 28:
            continue
        if route == "expand" and M * N * K > 1 << 33:
            continue
        out = acc.copy() if accumulate else np.full((M, N), 77, dtype=np.float32)
        before = ops.dispatch_count
        ops.matmul_mxfp4_host(M, N, K, lhs.ctypes.data, ls.ctypes.data, rhs.ctypes.data,
                              rs.ctypes.data, out.ctypes.data, flags=base | flag)
        if dispatches is not None:
            assert ops.dispatch_count - before == dispatches, route
        assert np.array_equal(out, want), route
    # The workspace-free recorded form (scales gathered inside the kernel): one dispatch.
    for family in (0, H.MATMUL_FORCE_1CTA, H.MATMUL_FORCE_2CTA):
        out = _recorded_without_workspace(ops, M, N, K, lhs, ls, rhs, rs, acc, base | family)
        assert np.array_equal(out, want), ("gather", family)


def _recorded_without_workspace(ops, M, N, K, lhs, ls, rhs, rs, acc, flags):
    from iree_b200 import hal
    dev = _device(ops)
    host = [lhs, ls, rhs, rs]
    out = acc.copy() if acc is not None else np.full((M, N), 77, dtype=np.float32)
    bufs = [dev.allocate(max(x.nbytes, 16)) for x in host] + [dev.allocate(out.nbytes)]
    try:
        for b, x in zip(bufs, host):
            dev.h2d(x.ctypes.data, b, x.nbytes)
        dev.h2d(out.ctypes.data, bufs[4], out.nbytes)
        before = ops.dispatch_count
        cb = hal.CommandBuffer(dev).begin()
        ops.record_matmul_mxfp4(cb, M, N, K, *(b.ref() for b in bufs), flags=flags)
        cb.end()
        assert ops.dispatch_count - before == 1
        dev.execute(cb)
        dev.d2h(bufs[4], out.ctypes.data, out.nbytes)
        dev.wait_idle()
        cb.release()
    finally:
        for b in bufs:
            b.release()
    return out


def _device(ops):
    return ops.device   # the session fixture's device (tests/conftest.py)


def test_block_scaled_kernel_random_scales_and_values(ops):
    """Full-range f4 codes and scales spread over 2^-10 .. 2^10 (PCG): sums are no longer exact,
    so the bar is the f32 summation-order tolerance against the oracle (k-ascending)."""
    M, K, N = 384, 1024, 512
    g = np.random.Generator(np.random.PCG64(0xB200))
    lhs = g.integers(0, 256, (M, K // 2), dtype=np.uint8)
    rhs = g.integers(0, 256, (N, K // 2), dtype=np.uint8)
    ls = g.integers(117, 138, (M, K // 32), dtype=np.uint8)
    rs = g.integers(117, 138, (N, K // 32), dtype=np.uint8)
    out = np.zeros((M, N), dtype=np.float32)
    ops.matmul_mxfp4_host(M, N, K, lhs.ctypes.data, ls.ctypes.data, rhs.ctypes.data,
                          rs.ctypes.data, out.ctypes.data)
    want = oracle.matmul_mxfp4(lhs, ls, rhs, rs)
    assert np.abs(out - want).max() <= 1e-5 * np.abs(want).max()


def test_block_scaled_kernel_special_scales(ops):
    """NaN scale (0xFF) poisons exactly the rows / columns it scales; the smallest scale
    (2^-127) gives products deep in the f32 subnormals (the tensor core may flush those)."""
    M, K, N = 128, 256, 256
    lhs, ls, rhs, rs, _ = _inputs(M, K, N, False)
    ls[2, 1] = 0xFF
    rs[5, 0] = 0
    ls[7, :] = 1
    out = np.zeros((M, N), dtype=np.float32)
    before = ops.dispatch_count
    ops.matmul_mxfp4_host(M, N, K, lhs.ctypes.data, ls.ctypes.data, rhs.ctypes.data,
                          rs.ctypes.data, out.ctypes.data)
    assert ops.dispatch_count - before == 3
    want = oracle.matmul_mxfp4(lhs, ls, rhs, rs)
    assert np.isnan(out[2]).all() and np.array_equal(np.isnan(out), np.isnan(want))
    ok = ~np.isnan(want)
    assert np.allclose(out[ok], want[ok], rtol=1e-6, atol=1e-30)


def test_mxfp4_special_scales_and_guards(ops):
    M, K, N = 8, 64, 8
    lhs, ls, rhs, rs, _ = _inputs(M, K, N, False)
    ls[2, 1] = 0xFF          # NaN scale poisons row 2
    rs[5, 0] = 0             # 2^-127: products underflow to f32 subnormals, still exact
    out = np.zeros((M, N), dtype=np.float32)
    ops.matmul_mxfp4_host(M, N, K, lhs.ctypes.data, ls.ctypes.data, rhs.ctypes.data,
                          rs.ctypes.data, out.ctypes.data)
    want = oracle.matmul_mxfp4(lhs, ls, rhs, rs)
    assert np.isnan(out[2]).all() and np.array_equal(out, want, equal_nan=True)
    with pytest.raises(H.Status):    # K not a multiple of the block
        ops.matmul_mxfp4_host(M, N, 48, lhs.ctypes.data, ls.ctypes.data, rhs.ctypes.data,
                              rs.ctypes.data, out.ctypes.data)
    with pytest.raises(H.Status):    # transpose / truncate flags do not apply
        ops.matmul_mxfp4_host(M, N, K, lhs.ctypes.data, ls.ctypes.data, rhs.ctypes.data,
                              rs.ctypes.data, out.ctypes.data, flags=H.MATMUL_TRANSPOSE_RHS)
    # K == 0 and empty results.
    out = np.full((M, N), 5, dtype=np.float32)
    ops.matmul_mxfp4_host(M, N, 0, 0, 0, 0, 0, out.ctypes.data)
    assert (out == 0).all()


def test_mxfp4_tensor_route_handles_special_scales(ops):
    """NaN scales, the subnormal end (2^-127) and overflow through the bf16 expansion and the
    tcgen05 GEMM: same values as the oracle (NaN where the oracle has NaN)."""
    M, K, N = 128, 128, 128
    lhs, ls, rhs, rs, _ = _inputs(M, K, N, False)
    ls[2, 1] = 0xFF
    rs[5, 0] = 0
    ls[7, :] = 1          # 2^-126: products with rs[5] land deep in the f32 subnormals
    out = np.zeros((M, N), dtype=np.float32)
    ops.matmul_mxfp4_host(M, N, K, lhs.ctypes.data, ls.ctypes.data, rhs.ctypes.data,
                          rs.ctypes.data, out.ctypes.data)
    want = oracle.matmul_mxfp4(lhs, ls, rhs, rs)
    assert np.isnan(out[2]).all() and np.array_equal(np.isnan(out), np.isnan(want))
    ok = ~np.isnan(want)
    assert np.allclose(out[ok], want[ok], rtol=1e-6, atol=1e-30)
