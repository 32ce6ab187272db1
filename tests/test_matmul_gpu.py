"""Parity of the CUDA dispatches with the oracle, through the C-ABI (B200 only).

Shape sets are the reference's (tests/e2e/matmul/generate_e2e_matmul_tests.py:33-101,
generate_e2e_batch_matmul_tests.py:29-31); inputs come from the reference generator
(seeds lhs=1, rhs=2, acc=3), for which f32 results are exact, so — as in the reference
runner (--require_exact_results, test_utils.c:287-290) — the bar is bit equality for
every element type. Random-input cases (PCG, N(0,1) / full-range i8): i8 stays bit-exact;
bf16 within 1e-2 and f16 within 1e-3 relative (BASELINE.json north_star).
"""

import numpy as np
import pytest

import oracle
from iree_b200 import hal as H

pytestmark = pytest.mark.gpu

SMALL = [(1, 1, 1, True), (1, 1, 1, False), (2, 2, 2, True), (4, 4, 4, True), (8, 8, 8, True),
         (9, 9, 9, True), (6, 13, 3, True), (15, 37, 7, False), (81, 19, 41, True),
         (1, 10, 10, True), (1, 10, 10, False), (10, 1, 10, True), (10, 10, 1, True),
         (10, 10, 1, False)]
LARGE = [(457, 330, 512, False), (438, 331, 513, False), (540, 332, 516, False),
         (1000, 4, 512, False), (4, 1000, 512, False), (512, 1000, 4, False),
         (513, 128, 55, False), (7, 160, 31, False), (512, 330, 33, False),
         (1, 1000, 1000, True), (1000, 1000, 1, True), (1000, 1000, 1, False)]
EASY_LARGE_STATIC = [(512, 128, 512, True), (512, 128, 512, False)]
# TMA-eligible shapes exercising ragged M/N/K tiles, multi-tile rasters, both families.
TILED = [(128, 64, 256, False), (256, 64, 256, True), (130, 72, 264, False),
         (1000, 520, 264, True), (384, 1024, 768, False), (2048, 256, 1280, False),
         (129, 8, 4, False), (8, 16, 2052, True)]


def _run(ops, t, batch, M, N, K, flags=0, accumulate=False, nt=False, seeds=(1, 2, 3),
         lhs=None, rhs=None):
    if lhs is None:
        lhs = oracle.generate_random_matrix(t, (batch, M, K), seeds[0])
        rhs = oracle.generate_random_matrix(t, (batch, N, K) if nt else (batch, K, N), seeds[1])
    out_t = "i32" if t == "i8" else "f32"
    acc = oracle.generate_random_matrix(out_t, (batch, M, N), seeds[2]) if accumulate else None
    out = acc.copy() if accumulate else np.full((batch, M, N), 77, dtype=acc_dtype(t))
    f = flags | (H.MATMUL_ACCUMULATE if accumulate else 0) | (H.MATMUL_TRANSPOSE_RHS if nt else 0)
    ops.matmul_host(t, batch, M, N, K, lhs.ctypes.data, rhs.ctypes.data, out.ctypes.data, f)
    want = oracle.matmul(t, lhs, rhs, acc=acc, transpose_rhs=nt)
    return out, want


def acc_dtype(t):
    return np.int32 if t == "i8" else np.float32


ALL_TYPES = ["i8", "f16", "bf16", "f8e4m3fn", "f8e5m2"]


def elem_size(t):
    return 2 if t in ("f16", "bf16") else 1


@pytest.mark.parametrize("t", ALL_TYPES)
@pytest.mark.parametrize("shape", SMALL + LARGE + EASY_LARGE_STATIC,
                         ids=lambda s: "%dx%dx%d%s" % (s[0], s[1], s[2], "+acc" if s[3] else ""))
def test_reference_shape_sets(ops, t, shape):
    M, K, N, accumulate = shape
    out, want = _run(ops, t, 1, M, N, K, accumulate=accumulate)
    assert np.array_equal(out[0], want[0])


@pytest.mark.parametrize("t", ALL_TYPES)
@pytest.mark.parametrize("family", ["1cta", "2cta", "generic"])
@pytest.mark.parametrize("nt", [False, True], ids=["nn", "nt"])
@pytest.mark.parametrize("shape", TILED, ids=lambda s: "%dx%dx%d%s" % (s[0], s[1], s[2], "+acc" if s[3] else ""))
def test_every_kernel_family(ops, t, family, nt, shape):
    M, K, N, accumulate = shape
    es = elem_size(t)
    if family != "generic" and ((K * es) % 16 or (N * 4) % 16 or (not nt and (N * es) % 16)):
        pytest.skip("shape does not satisfy the TMA guards (goes to the generic kernel)")
    flag = {"1cta": H.MATMUL_FORCE_1CTA, "2cta": H.MATMUL_FORCE_2CTA,
            "generic": H.MATMUL_FORCE_GENERIC}[family]
    out, want = _run(ops, t, 1, M, N, K, flags=flag, accumulate=accumulate, nt=nt)
    assert np.array_equal(out, want)


@pytest.mark.parametrize("t", ["i8", "f16", "bf16", "f8e4m3fn", "f8e5m2"])
@pytest.mark.parametrize("family", ["1cta", "2cta"])
@pytest.mark.parametrize("nt", [False, True], ids=["nn", "nt"])
@pytest.mark.parametrize("shape", TILED + [(4, 512, 384, 640, False)],
                         ids=lambda s: "x".join(str(v) for v in s))
def test_device_built_tensor_maps(ops, t, family, nt, shape):
    """`*_dtm` exports (FLAG_DEVICE_MAPS): the tcgen05 GEMM with no tensor-map parameters — it
    patches host-encoded templates with tensormap.replace — i.e. what an unmodified IREE CUDA
    driver can launch. Same bits as the oracle, every family, ragged tiles, batch, accumulate;
    back-to-back launches with different shapes reuse the per-CTA slots."""
    batch, (M, K, N, accumulate) = (shape[0], shape[1:]) if len(shape) == 5 else (1, shape)
    es = elem_size(t)
    if (K * es) % 16 or (N * 4) % 16 or (not nt and (N * es) % 16):
        pytest.skip("shape does not satisfy the TMA guards (goes to the generic kernel)")
    flag = {"1cta": H.MATMUL_FORCE_1CTA, "2cta": H.MATMUL_FORCE_2CTA}[family] | H.MATMUL_DEVICE_MAPS
    sel = ops.select_matmul(t, batch, M, N, K, flags=flag | (H.MATMUL_TRANSPOSE_RHS if nt else 0))
    info = ops.executable.export_info(sel["export_ordinal"])
    assert "_dtm" in info["kernel_name"] and info["tma_map_count"] == 0
    out, want = _run(ops, t, batch, M, N, K, flags=flag, accumulate=accumulate, nt=nt)
    assert np.array_equal(out, want)


@pytest.mark.parametrize("t", ["f16", "bf16"])
@pytest.mark.parametrize("family", ["1cta", "2cta"])
@pytest.mark.parametrize("nt", [False, True], ids=["nn", "nt"])
@pytest.mark.parametrize("shape", [(256, 64, 256), (1000, 520, 264), (2048, 256, 1280)], ids=str)
def test_device_built_tensor_maps_truncating_epilogue(ops, t, family, nt, shape):
    """The truncating `*_dtm` twins: the 16-bit result map is patched from its own template."""
    M, K, N = shape
    flag = {"1cta": H.MATMUL_FORCE_1CTA, "2cta": H.MATMUL_FORCE_2CTA}[family]
    f = flag | H.MATMUL_TRUNCATE_RESULT | H.MATMUL_DEVICE_MAPS | (H.MATMUL_TRANSPOSE_RHS if nt else 0)
    name = ops.executable.export_info(ops.select_matmul(t, 1, M, N, K, flags=f)["export_ordinal"])
    assert name["kernel_name"].endswith("_dtm") and f"_{t}_{t}" in name["kernel_name"]
    lhs = oracle.generate_random_matrix(t, (1, M, K), 1)
    rhs = oracle.generate_random_matrix(t, (1, N, K) if nt else (1, K, N), 2)
    out = np.full((1, M, N), 0x7777, dtype=np.uint16)
    ops.matmul_host(t, 1, M, N, K, lhs.ctypes.data, rhs.ctypes.data, out.ctypes.data, f)
    assert np.array_equal(out, oracle.from_f32(t, oracle.matmul(t, lhs, rhs, transpose_rhs=nt)))


@pytest.mark.parametrize("t", ["i8", "bf16", "f8e4m3fn"])
@pytest.mark.parametrize("bshape", [(32, 128, 128, 128), (32, 256, 64, 256), (32, 96, 96, 96),
                                    (3, 300, 136, 520)], ids=str)
def test_batch_matmul_reference_shapes(ops, t, bshape):
    batch, M, K, N = bshape
    out, want = _run(ops, t, batch, M, N, K)
    assert np.array_equal(out, want)


def test_selection_guards(ops):
    """The matcher guards: unaligned pitches / pointers go to the generic export; big
    aligned shapes to the 2-CTA family with a persistent, cluster-divisible grid."""
    exe = ops.executable
    name = lambda sel: exe.export_info(sel["export_ordinal"])["kernel_name"]
    s = ops.select_matmul("bf16", 1, 8192, 8192, 8192)
    assert name(s) == "iree_b200_matmul_bf16_f32"
    assert s["workgroup_count"][0] % 2 == 0 and s["workgroup_count"][0] <= 148
    assert name(ops.select_matmul("bf16", 1, 128, 512, 512)).endswith("_1cta")
    assert "generic" in name(ops.select_matmul("bf16", 1, 457, 330, 512))   # N pitch 660 B
    assert "generic" in name(ops.select_matmul("bf16", 1, 512, 512, 512, alignment=2))
    assert name(ops.select_matmul("i8", 1, 512, 512, 512, flags=H.MATMUL_TRANSPOSE_RHS)) \
        .startswith("iree_b200_matmul_nt_i8")
    assert ops.select_matmul("bf16", 1, 0, 16, 16)["export_ordinal"] == -1
    with pytest.raises(H.Status):
        ops.select_matmul("bf16", 1, 457, 330, 512, flags=H.MATMUL_FORCE_2CTA)
    with pytest.raises(H.Status):
        ops.select_matmul(7, 1, 8, 8, 8)


def test_empty_and_zero_k(ops):
    out = np.full((1, 4, 4), 5, np.float32)
    a = np.zeros((1,), np.uint16)
    ops.matmul_host("bf16", 1, 4, 4, 0, a.ctypes.data, a.ctypes.data, out.ctypes.data)
    assert (out == 0).all()                      # linalg.fill 0 seeds the accumulator
    out[:] = 5
    ops.matmul_host("bf16", 1, 4, 4, 0, a.ctypes.data, a.ctypes.data, out.ctypes.data,
                    H.MATMUL_ACCUMULATE)
    assert (out == 5).all()
    ops.matmul_host("bf16", 1, 0, 4, 8, a.ctypes.data, a.ctypes.data, out.ctypes.data)


def _pcg(seed):
    return np.random.Generator(np.random.PCG64(seed))


def test_i8_full_range_bit_exact(ops):
    g = _pcg(0xB200)
    M, N, K = 512, 768, 2048
    lhs = g.integers(-128, 128, (1, M, K), dtype=np.int8)
    rhs = g.integers(-128, 128, (1, K, N), dtype=np.int8)
    for flag in (H.MATMUL_FORCE_1CTA, H.MATMUL_FORCE_2CTA):
        out, want = _run(ops, "i8", 1, M, N, K, flags=flag, lhs=lhs, rhs=rhs)
        assert np.array_equal(out, want)


@pytest.mark.parametrize("t", ["f8e4m3fn", "f8e5m2"])
def test_f8_random_inputs(ops, t):
    """N(0,1) inputs rounded to the 8-bit type (so also subnormal and large codes), both
    tcgen05 families and the generic kernel. Products of two 8-bit floats are exact in f32;
    what differs from the oracle is the f32 summation order only."""
    g = _pcg(0xB200 + 8)
    M, N, K = 384, 512, 4096
    scale_in = 1.0 if t == "f8e4m3fn" else 16.0
    lhs = oracle.from_f32(t, (g.standard_normal((1, M, K)) * scale_in).astype(np.float32))
    rhs = oracle.from_f32(t, (g.standard_normal((1, K, N)) * scale_in).astype(np.float32))
    for flag in (H.MATMUL_FORCE_1CTA, H.MATMUL_FORCE_2CTA, H.MATMUL_FORCE_GENERIC):
        out, want = _run(ops, t, 1, M, N, K, flags=flag, lhs=lhs, rhs=rhs)
        scale = np.abs(want).max()
        assert np.isfinite(out).all()
        assert np.abs(out - want).max() <= 1e-3 * scale, (flag, np.abs(out - want).max(), scale)


def test_f8_every_code_decodes_like_the_oracle(ops):
    """A [256, 32] x [32, 256] product with a one-hot RHS reads back every 8-bit code through
    the tensor core's own decoder (and the generic kernel's): it must equal the oracle's
    table bit for bit (NaN codes compare as NaN)."""
    for t in ("f8e4m3fn", "f8e5m2"):
        codes = np.arange(256, dtype=np.uint8)
        one = oracle.from_f32(t, np.array([1.0], np.float32))[0]
        lhs = np.zeros((1, 256, 32), np.uint8)
        lhs[0, :, 0] = codes
        rhs = np.zeros((1, 32, 256), np.uint8)
        rhs[0, 0, :] = one
        want = oracle.to_f32(t, codes)
        for flag in (H.MATMUL_FORCE_1CTA, H.MATMUL_FORCE_GENERIC):
            out, _ = _run(ops, t, 1, 256, 256, 32, flags=flag, lhs=lhs, rhs=rhs)
            got = out[0, :, 7]
            fin = np.isfinite(want)
            assert np.array_equal(got[fin], want[fin]), (t, flag)
            assert np.isnan(got[np.isnan(want)]).all() and np.array_equal(got[np.isinf(want)], want[np.isinf(want)])


@pytest.mark.parametrize("t,rtol", [("bf16", 1e-2), ("f16", 1e-3)])
def test_random_normal_inputs_within_tolerance(ops, t, rtol):
    g = _pcg(0xB200 + 1)
    M, N, K = 384, 512, 4096
    lhs = oracle.from_f32(t, np.clip(g.standard_normal((1, M, K)), -4, 4).astype(np.float32))
    rhs = oracle.from_f32(t, np.clip(g.standard_normal((1, K, N)), -4, 4).astype(np.float32))
    out, want = _run(ops, t, 1, M, N, K, lhs=lhs, rhs=rhs)
    # Tolerance from BASELINE.json north_star, relative to the result's scale (sqrt(K)).
    scale = np.abs(want).max()
    assert np.abs(out - want).max() <= rtol * scale
    # In practice both sides accumulate exact products in f32: the difference is
    # summation order only.
    assert np.abs(out - want).max() <= 1e-5 * scale


@pytest.mark.parametrize("t", ["i8", "bf16", "f8e4m3fn"])
def test_full_size_sampled_like_the_reference(ops, t):
    """BASELINE config sizes: compare the reference checker's sampled index set
    (iree_test_utils_calculate_check_every, test_utils.c:41-68) at 4096^3."""
    M = N = K = 4096
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    out = np.empty((M, N), acc_dtype(t))
    ops.matmul_host(t, 1, M, N, K, lhs.ctypes.data, rhs.ctypes.data, out.ctypes.data)
    idx, every = oracle.sampled_indices(M, N)
    for i in idx[:: max(1, len(idx) // 2000)]:
        m, n = divmod(int(i), N)
        assert out[m, n] == oracle.matmul_element(t, lhs, rhs, m, n)
    # Size-independent property: row/column checksums (exact in f64 for these inputs).
    a = lhs.astype(np.float64) if t == "i8" else oracle.to_f32(t, lhs).astype(np.float64)
    b = rhs.astype(np.float64) if t == "i8" else oracle.to_f32(t, rhs).astype(np.float64)
    assert np.array_equal(out.astype(np.float64).sum(axis=0), a.sum(axis=0) @ b)
    assert np.array_equal(out.astype(np.float64).sum(axis=1), a @ b.sum(axis=1))


@pytest.mark.parametrize("t", ["f16", "bf16"])
@pytest.mark.parametrize("family", ["1cta", "2cta", "generic"])
@pytest.mark.parametrize("nt", [False, True], ids=["nn", "nt"])
@pytest.mark.parametrize("shape", [(256, 64, 256), (130, 72, 264), (1000, 520, 264),
                                   (384, 1024, 768)], ids=str)
def test_truncating_epilogue(ops, t, family, nt, shape):
    """matmul + arith.truncf in one dispatch: out = round_T(f32 accumulate), bit-exact for the
    reference generator's inputs (the f32 result is exact, so one rounding on both sides)."""
    M, K, N = shape
    flag = {"1cta": H.MATMUL_FORCE_1CTA, "2cta": H.MATMUL_FORCE_2CTA,
            "generic": H.MATMUL_FORCE_GENERIC}[family]
    lhs = oracle.generate_random_matrix(t, (1, M, K), 1)
    rhs = oracle.generate_random_matrix(t, (1, N, K) if nt else (1, K, N), 2)
    out = np.full((1, M, N), 0x7777, dtype=np.uint16)
    f = flag | H.MATMUL_TRUNCATE_RESULT | (H.MATMUL_TRANSPOSE_RHS if nt else 0)
    ops.matmul_host(t, 1, M, N, K, lhs.ctypes.data, rhs.ctypes.data, out.ctypes.data, f)
    want = oracle.from_f32(t, oracle.matmul(t, lhs, rhs, transpose_rhs=nt))
    assert np.array_equal(out, want)


def test_truncating_epilogue_guards(ops):
    with pytest.raises(H.Status):
        ops.select_matmul("i8", 1, 256, 256, 256, flags=H.MATMUL_TRUNCATE_RESULT)
    with pytest.raises(H.Status):
        ops.select_matmul("bf16", 1, 256, 256, 256,
                          flags=H.MATMUL_TRUNCATE_RESULT | H.MATMUL_ACCUMULATE)
    exe = ops.executable
    s = ops.select_matmul("bf16", 1, 8192, 4096, 4096, flags=H.MATMUL_TRUNCATE_RESULT)
    assert exe.export_info(s["export_ordinal"])["kernel_name"] == "iree_b200_matmul_bf16_bf16"


@pytest.mark.parametrize("t", ALL_TYPES)
@pytest.mark.parametrize("case", [
    # batch, M, K, N, accumulate, nt, truncate
    (1, 1000, 272, 528, False, False, False),   # 4x3 blocks, ragged last row / column panel
    (1, 1000, 272, 528, True, False, False),    # accumulator blocks go up too
    (1, 528, 144, 1008, False, True, False),    # transpose_rhs: panels are rows of [N, K]
    (3, 300, 528, 272, False, False, False),    # batch entries are separate blocks
    (2, 513, 64, 768, False, False, True),      # truncating epilogue (2-byte result blocks)
    (1, 256, 128, 2048, False, False, False),   # one row panel, 8 column panels
    (1, 2048, 128, 256, True, False, False),    # 8 row panels, one column panel
    # 512-blocks
    (1, 1536, 144, 1280, False, False, False, 512),
    (1, 1104, 80, 1040, False, True, False, 512),
    (2, 1024, 64, 1296, True, False, False, 512),
    (1, 1280, 64, 784, False, False, True, 512),
], ids=str)
def test_pipelined_host_call(ops, t, case):
    """iree_b200_ops_matmul_host as a pipeline (panel uploads, per-block dispatch, block
    downloads on three streams): same bits as the single-dispatch form and the oracle."""
    batch, M, K, N, accumulate, nt, truncate = case[:7]
    block = case[7] if len(case) > 7 else 256
    if truncate and elem_size(t) == 1:
        pytest.skip("TRUNCATE_RESULT is f16 / bf16 only")
    before = ops.dispatch_count
    ops.set_host_pipeline(0, block, block)
    try:
        if truncate:
            lhs = oracle.generate_random_matrix(t, (batch, M, K), 1)
            rhs = oracle.generate_random_matrix(t, (batch, K, N), 2)
            out = np.full((batch, M, N), 0x7777, dtype=np.uint16)
            ops.matmul_host(t, batch, M, N, K, lhs.ctypes.data, rhs.ctypes.data, out.ctypes.data,
                            H.MATMUL_TRUNCATE_RESULT)
            want = oracle.from_f32(t, oracle.matmul(t, lhs, rhs))
        else:
            out, want = _run(ops, t, batch, M, N, K, accumulate=accumulate, nt=nt)
    finally:
        ops.set_host_pipeline(16 << 20, 2048, 2048)
    assert np.array_equal(out, want)
    blocks = batch * -(-M // block) * -(-N // block)
    assert ops.dispatch_count - before == blocks


# ---------------------------------------------------------------------------------------
# BASELINE.json configuration sizes, as such (configs[1] and configs[2]).
# ---------------------------------------------------------------------------------------

def _checksums_exact(out, a64, b64):
    """Row / column checksum identity, exact: sum_m C[m,:] = (sum_m A[m,:]) . B and
    sum_n C[:,n] = A . (sum_n B[:,n]). Every intermediate is an integer (or a multiple of a
    small power of two) far below 2^53, so f64 is exact."""
    assert np.array_equal(out.sum(axis=0, dtype=np.float64), a64.sum(axis=0) @ b64)
    assert np.array_equal(out.sum(axis=1, dtype=np.float64), a64 @ b64.sum(axis=1))


@pytest.mark.parametrize("inputs", ["reference_generator", "full_range"])
def test_config_i8_8192_cubed_bit_exact(device, ops, inputs):
    """BASELINE configs[1]: linalg.matmul i8 x i8 -> i32 8192^3, ONE device-resident dispatch
    (the form bench.py times), bit-exact: the reference checker's sampled index set
    (test_utils.c:41-68) against the oracle's element loop + exact row / column checksums of
    the whole result. `full_range` draws i8 over [-128, 127] (|C| < 2^27, no i32 wrap)."""
    M = N = K = 8192
    if inputs == "reference_generator":
        lhs = oracle.generate_random_matrix("i8", (M, K), 1)
        rhs = oracle.generate_random_matrix("i8", (K, N), 2)
    else:
        g = np.random.Generator(np.random.PCG64(0xB200))
        lhs = g.integers(-128, 128, (M, K), dtype=np.int8)
        rhs = g.integers(-128, 128, (K, N), dtype=np.int8)
    out = np.empty((M, N), np.int32)
    bl, br, bo = device.allocate(lhs.nbytes), device.allocate(rhs.nbytes), device.allocate(out.nbytes)
    try:
        device.h2d(lhs.ctypes.data, bl, lhs.nbytes)
        device.h2d(rhs.ctypes.data, br, rhs.nbytes)
        sel = ops.select_matmul("i8", 1, M, N, K)
        assert ops.executable.export_info(sel["export_ordinal"])["kernel_name"] == "iree_b200_matmul_i8_i32"
        cb = H.CommandBuffer(device).begin()
        ops.record_matmul(cb, "i8", 1, M, N, K, bl.ref(), br.ref(), bo.ref())
        cb.end()
        device.execute(cb)
        device.d2h(bo, out.ctypes.data, out.nbytes)
        device.wait_idle()
        cb.release()
    finally:
        for b in (bl, br, bo):
            b.release()
    idx, _ = oracle.sampled_indices(M, N)
    assert len(idx) >= 8000
    for i in idx:
        m, n = divmod(int(i), N)
        assert out[m, n] == oracle.matmul_element("i8", lhs, rhs, m, n), (m, n)
    _checksums_exact(out, lhs.astype(np.float64), rhs.astype(np.float64))


def test_config_bf16_8192_cubed_exact(device, ops):
    """BASELINE configs[1], bf16 half: 8192^3 in one dispatch on the reference generator's
    inputs (f32 result exact => bit equality), sampled like the reference + checksums."""
    M = N = K = 8192
    lhs = oracle.generate_random_matrix("bf16", (M, K), 1)
    rhs = oracle.generate_random_matrix("bf16", (K, N), 2)
    out = np.empty((M, N), np.float32)
    bl, br, bo = device.allocate(lhs.nbytes), device.allocate(rhs.nbytes), device.allocate(out.nbytes)
    try:
        device.h2d(lhs.ctypes.data, bl, lhs.nbytes)
        device.h2d(rhs.ctypes.data, br, rhs.nbytes)
        cb = H.CommandBuffer(device).begin()
        ops.record_matmul(cb, "bf16", 1, M, N, K, bl.ref(), br.ref(), bo.ref())
        cb.end()
        device.execute(cb)
        device.d2h(bo, out.ctypes.data, out.nbytes)
        device.wait_idle()
        cb.release()
    finally:
        for b in (bl, br, bo):
            b.release()
    idx, _ = oracle.sampled_indices(M, N)
    for i in idx:
        m, n = divmod(int(i), N)
        assert out[m, n] == oracle.matmul_element("bf16", lhs, rhs, m, n), (m, n)
    _checksums_exact(out, oracle.to_f32("bf16", lhs).astype(np.float64),
                     oracle.to_f32("bf16", rhs).astype(np.float64))


def test_config_batch_matmul_64x2048_cubed_exact(device, ops):
    """BASELINE configs[2]: linalg.batch_matmul bf16 batch=64 M=N=K=2048 as ONE dispatch
    (batch is a tile coordinate of the persistent grid), every batch entry sampled like the
    reference's per-batch check (iree-e2e-matmul-test.cc:882-911) + exact checksums of each
    entry. The same buffers are then run as the 8-way shard (8 dispatches of batch=8 on
    sub-ranges, what each rank of the batch-sharded config executes): identical bits."""
    B, M, N, K = 64, 2048, 2048, 2048
    lhs = oracle.generate_random_matrix("bf16", (B, M, K), 1)
    rhs = oracle.generate_random_matrix("bf16", (B, K, N), 2)
    out = np.empty((B, M, N), np.float32)
    bl, br, bo = device.allocate(lhs.nbytes), device.allocate(rhs.nbytes), device.allocate(out.nbytes)
    try:
        device.h2d(lhs.ctypes.data, bl, lhs.nbytes)
        device.h2d(rhs.ctypes.data, br, rhs.nbytes)
        cb = H.CommandBuffer(device).begin()
        ops.record_matmul(cb, "bf16", B, M, N, K, bl.ref(), br.ref(), bo.ref())
        cb.end()
        device.execute(cb)
        device.d2h(bo, out.ctypes.data, out.nbytes)
        device.wait_idle()
        cb.release()
        idx, _ = oracle.sampled_indices(M, N, max_elements=400)
        for b in range(B):
            for i in idx:
                m, n = divmod(int(i), N)
                assert out[b, m, n] == oracle.matmul_element("bf16", lhs[b], rhs[b], m, n), (b, m, n)
            _checksums_exact(out[b], oracle.to_f32("bf16", lhs[b]).astype(np.float64),
                             oracle.to_f32("bf16", rhs[b]).astype(np.float64))
        # The sharded form: rank r of 8 owns entries [8r, 8r+8).
        shard = np.empty_like(out)
        fill = H.CommandBuffer(device).begin()
        fill.fill(bo.ref(), b"\xff\xff\xff\xff")
        fill.end()
        device.execute(fill)
        cb = H.CommandBuffer(device).begin()
        per = B // 8
        for r in range(8):
            ops.record_matmul(cb, "bf16", per, M, N, K,
                              bl.ref(r * per * M * K * 2, per * M * K * 2),
                              br.ref(r * per * K * N * 2, per * K * N * 2),
                              bo.ref(r * per * M * N * 4, per * M * N * 4))
        cb.end()
        device.execute(cb)
        device.d2h(bo, shard.ctypes.data, shard.nbytes)
        device.wait_idle()
        cb.release()
        fill.release()
        assert np.array_equal(shard, out)
    finally:
        for b in (bl, br, bo):
            b.release()
