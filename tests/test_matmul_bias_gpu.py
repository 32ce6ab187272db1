"""The fused matmul + elementwise-consumer dispatches: out =
act((out +) lhs·rhs + bias[n]) [truncated], against the oracle's op-by-op restatement
(oracle.matmul_bias), through the C-ABI (B200 only).

Inputs come from the reference generator (test_utils.c:359-420: small integers, so every
f32 value on the way is exact) and the bar is bit equality, as for the plain matmuls; a
random-normal case carries the stated tolerance.
"""

import numpy as np
import pytest

import oracle
from iree_b200 import hal as H

pytestmark = pytest.mark.gpu

FAMILY = {"1cta": H.MATMUL_FORCE_1CTA, "2cta": H.MATMUL_FORCE_2CTA,
          "generic": H.MATMUL_FORCE_GENERIC, "auto": 0}


def _acc_t(t):
    return "i32" if t == "i8" else "f32"


def _run(ops, t, batch, M, N, K, *, family="auto", nt=False, relu=False, accumulate=False,
         truncate=False, bias=None, extra_flags=0):
    lhs = oracle.generate_random_matrix(t, (batch, M, K), 1)
    rhs = oracle.generate_random_matrix(t, (batch, N, K) if nt else (batch, K, N), 2)
    acc = oracle.generate_random_matrix(_acc_t(t), (batch, M, N), 3) if accumulate else None
    if bias is None:
        bias = oracle.generate_random_matrix(_acc_t(t), (N,), 4)
    if truncate:
        out = np.full((batch, M, N), 0x7777, dtype=np.uint16)
    else:
        out = acc.copy() if accumulate else np.full((batch, M, N), 77, dtype=bias.dtype)
    flags = (FAMILY[family] | (H.MATMUL_TRANSPOSE_RHS if nt else 0) | (H.MATMUL_RELU if relu else 0)
             | (H.MATMUL_ACCUMULATE if accumulate else 0)
             | (H.MATMUL_TRUNCATE_RESULT if truncate else 0) | extra_flags)
    ops.matmul_bias_host(t, batch, M, N, K, lhs.ctypes.data, rhs.ctypes.data, bias.ctypes.data,
                         out.ctypes.data, flags)
    want = oracle.matmul_bias(t, lhs, rhs, bias, acc=acc, transpose_rhs=nt, relu=relu,
                              truncate=truncate)
    return out, want


def _tma_ok(t, N, K, nt, out_es=4):
    es = 1 if t == "i8" else 2
    return not ((K * es) % 16 or (N * out_es) % 16 or (not nt and (N * es) % 16))


@pytest.mark.parametrize("t", ["i8", "f16", "bf16"])
@pytest.mark.parametrize("family", ["1cta", "2cta", "generic"])
@pytest.mark.parametrize("nt", [False, True], ids=["nn", "nt"])
@pytest.mark.parametrize("relu", [False, True], ids=["bias", "bias+relu"])
@pytest.mark.parametrize("shape", [(256, 64, 256, False), (130, 72, 264, True),
                                   (1000, 528, 272, False), (384, 1024, 768, True)], ids=str)
def test_bias_epilogue(ops, t, family, nt, relu, shape):
    M, K, N, accumulate = shape
    if family != "generic" and not _tma_ok(t, N, K, nt):
        pytest.skip("shape does not satisfy the TMA guards (goes to the generic kernel)")
    before = ops.dispatch_count
    out, want = _run(ops, t, 1, M, N, K, family=family, nt=nt, relu=relu, accumulate=accumulate)
    assert np.array_equal(out, want)
    assert ops.dispatch_count - before == 1   # one dispatch: nothing runs beside the GEMM
    if relu:
        assert (out >= 0).all() and (out == 0).any() and (out > 0).any()


@pytest.mark.parametrize("t", ["f16", "bf16"])
@pytest.mark.parametrize("family", ["1cta", "2cta", "generic"])
@pytest.mark.parametrize("nt", [False, True], ids=["nn", "nt"])
@pytest.mark.parametrize("relu", [False, True], ids=["bias", "bias+relu"])
@pytest.mark.parametrize("shape", [(256, 64, 256), (130, 72, 264), (1000, 520, 264)], ids=str)
def test_bias_then_truncate(ops, t, family, nt, relu, shape):
    """The bf16 linear layer: bf16 operands, f32 accumulate + f32 bias, [relu,] bf16 result."""
    M, K, N = shape
    if family != "generic" and not _tma_ok(t, N, K, nt, out_es=2):
        pytest.skip("shape does not satisfy the TMA guards")
    out, want = _run(ops, t, 1, M, N, K, family=family, nt=nt, relu=relu, truncate=True)
    assert np.array_equal(out, want)


@pytest.mark.parametrize("t", ["i8", "bf16"])
def test_bias_is_broadcast_over_batch_and_rows(ops, t):
    out, want = _run(ops, t, 3, 300, 272, 528, relu=True)
    assert np.array_equal(out, want)
    # Ragged everything (generic kernel), batch 2.
    out, want = _run(ops, t, 2, 37, 19, 15)
    assert np.array_equal(out, want)


@pytest.mark.parametrize("t", ["i8", "f16", "bf16"])
def test_bias_with_an_empty_reduction(ops, t):
    """K == 0: linalg.fill 0 is the matmul result, so the dispatch stores act(bias)."""
    out, want = _run(ops, t, 1, 20, 24, 0, relu=True)
    assert np.array_equal(out, want)
    bias = oracle.generate_random_matrix(_acc_t(t), (24,), 4)
    assert np.array_equal(out[0], np.broadcast_to(np.maximum(bias, 0), (20, 24)))
    # Empty results: nothing to do, nothing touched.
    for M, N in ((0, 16), (16, 0)):
        before = ops.dispatch_count
        _run(ops, t, 1, M, N, 8)
        assert ops.dispatch_count == before


@pytest.mark.parametrize("family", ["2cta", "1cta", "generic"])
def test_maximumf_semantics(ops, family):
    """arith.maximumf: NaN propagates; -inf clamps to 0; +inf stays."""
    M, K, N = 256, 64, 256
    bias = oracle.generate_random_matrix("f32", (N,), 4)
    bias[3], bias[40], bias[41], bias[200] = np.nan, np.inf, -np.inf, np.nan
    out, want = _run(ops, "bf16", 1, M, N, K, family=family, relu=True, bias=bias)
    assert np.isnan(out[0, :, 3]).all() and np.isnan(out[0, :, 200]).all()
    assert (out[0, :, 40] == np.inf).all() and (out[0, :, 41] == 0).all()
    assert np.array_equal(out, want, equal_nan=True)


def test_i32_bias_wraps_like_addi(ops):
    M, K, N = 128, 64, 256
    bias = np.full((N,), np.iinfo(np.int32).max, dtype=np.int32)
    bias[::2] = np.iinfo(np.int32).min
    out, want = _run(ops, "i8", 1, M, N, K, bias=bias)
    assert np.array_equal(out, want)
    out, want = _run(ops, "i8", 1, M, N, K, bias=bias, relu=True)
    assert np.array_equal(out, want)


@pytest.mark.parametrize("t,rtol", [("f16", 1e-3), ("bf16", 1e-2)])
def test_random_normal_inputs_within_tolerance(ops, t, rtol):
    rng = np.random.Generator(np.random.PCG64(0xB200))
    M, K, N = 512, 1024, 768
    lhs = oracle.from_f32(t, rng.standard_normal((1, M, K), dtype=np.float32))
    rhs = oracle.from_f32(t, rng.standard_normal((1, K, N), dtype=np.float32))
    bias = rng.standard_normal((N,), dtype=np.float32) * 8
    out = np.empty((1, M, N), dtype=np.float32)
    ops.matmul_bias_host(t, 1, M, N, K, lhs.ctypes.data, rhs.ctypes.data, bias.ctypes.data,
                         out.ctypes.data, H.MATMUL_RELU)
    want = oracle.matmul_bias(t, lhs, rhs, bias, relu=True)
    scale = np.abs(want).max()
    assert np.abs(out - want).max() <= rtol * scale


def test_bias_binding_needs_only_element_alignment(ops, device):
    """Recorded form with the bias at a 4-byte offset into its buffer (the ABI promises
    element alignment only, kernels.cu:28-34); the GEMM still takes the tcgen05 export."""
    t, M, K, N = "bf16", 256, 64, 256
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    bias = oracle.generate_random_matrix("f32", (N + 1,), 4)
    bufs = []
    for a in (lhs, rhs, bias):
        b = device.allocate(a.nbytes)
        device.h2d(a.ctypes.data, b, a.nbytes)
        bufs.append(b)
    out_b = device.allocate(M * N * 4)
    cb = H.CommandBuffer(device).begin()
    ops.record_matmul_bias(cb, t, 1, M, N, K, bufs[0].ref(), bufs[1].ref(),
                           bufs[2].ref(4, N * 4), out_b.ref(), flags=H.MATMUL_RELU)
    cb.end()
    device.execute(cb)
    device.wait_idle()
    out = np.empty((M, N), dtype=np.float32)
    device.d2h(out_b, out.ctypes.data, out.nbytes)
    assert np.array_equal(out, oracle.matmul_bias(t, lhs, rhs, bias[1:], relu=True))
    with pytest.raises(H.Status):   # 2-byte offset: not an f32 element boundary
        cb2 = H.CommandBuffer(device).begin()
        ops.record_matmul_bias(cb2, t, 1, M, N, K, bufs[0].ref(), bufs[1].ref(),
                               bufs[2].ref(2, N * 4), out_b.ref())
    with pytest.raises(H.Status):   # bias binding too short
        cb3 = H.CommandBuffer(device).begin()
        ops.record_matmul_bias(cb3, t, 1, M, N, K, bufs[0].ref(), bufs[1].ref(),
                               bufs[2].ref(0, N * 4 - 4), out_b.ref())


def test_bias_selection_and_guards(ops):
    exe = ops.executable

    def name(*a, **k):
        return exe.export_info(ops.select_matmul(*a, **k)["export_ordinal"])["kernel_name"]

    B, R, T = H.MATMUL_BIAS, H.MATMUL_RELU, H.MATMUL_TRUNCATE_RESULT
    assert name("bf16", 1, 8192, 4096, 4096, flags=B) == "iree_b200_matmul_bias_bf16_f32"
    assert name("bf16", 1, 8192, 4096, 4096, flags=B | R | T) == "iree_b200_matmul_bias_bf16_bf16"
    assert name("i8", 1, 128, 4096, 4096, flags=B | H.MATMUL_TRANSPOSE_RHS) == \
        "iree_b200_matmul_bias_nt_i8_i32_1cta"
    assert name("f16", 1, 100, 33, 7, flags=B) == "iree_b200_matmul_generic_bias_f16_f32"
    s = ops.select_matmul("f16", 1, 8192, 4096, 4096, flags=B | R | H.MATMUL_ACCUMULATE)
    assert s["constants"][4] == 3          # tcgen05 kernels: bit 0 accumulate, bit 1 relu
    s = ops.select_matmul("f16", 1, 100, 33, 7, flags=B | R | H.MATMUL_TRANSPOSE_RHS)
    assert s["constants"][4] == 2 | 8      # generic kernels: 2 transpose_rhs, 8 relu
    info = exe.export_info(ops.select_matmul("bf16", 1, 512, 512, 512, flags=B)["export_ordinal"])
    assert info["binding_count"] == 4 and info["constant_count"] == 5
    with pytest.raises(H.Status):
        ops.select_matmul("bf16", 1, 256, 256, 256, flags=R)            # relu needs the bias form
    with pytest.raises(H.Status):
        ops.select_matmul("f8e4m3fn", 1, 256, 256, 256, flags=B)        # no f8 bias exports
    with pytest.raises(H.Status):
        ops.matmul_host("bf16", 1, 16, 16, 16, 0, 0, 0, flags=B)        # wrong entry point


@pytest.mark.parametrize("t", ["i8", "f16", "bf16"])
@pytest.mark.parametrize("family", ["1cta", "2cta"])
@pytest.mark.parametrize("case", [dict(nt=False, relu=True, truncate=True), dict(nt=True, relu=False),
                                  dict(nt=False, relu=True, accumulate=True)], ids=str)
@pytest.mark.parametrize("shape", [(1, 256, 256, 64), (1, 1000, 264, 520), (3, 300, 520, 136)], ids=str)
def test_fused_epilogue_with_device_built_tensor_maps(ops, t, family, case, shape):
    """`iree_b200_matmul_bias_*_dtm`: the fused bias / activation / truncation dispatches in
    the form an unmodified IREE CUDA driver launches (FLAG_DEVICE_MAPS). Bit-exact."""
    batch, M, N, K = shape
    case = dict(case)
    if case.get("truncate") and t == "i8":
        case.pop("truncate")
    if case.get("truncate") and case.get("accumulate"):
        case.pop("accumulate")
    if not _tma_ok(t, N, K, case.get("nt", False), out_es=2 if case.get("truncate") else 4):
        pytest.skip("shape does not satisfy the TMA guards")
    flags = (FAMILY[family] | H.MATMUL_BIAS | H.MATMUL_DEVICE_MAPS
             | (H.MATMUL_TRANSPOSE_RHS if case.get("nt") else 0) | (H.MATMUL_RELU if case.get("relu") else 0)
             | (H.MATMUL_TRUNCATE_RESULT if case.get("truncate") else 0))
    info = ops.executable.export_info(ops.select_matmul(t, batch, M, N, K, flags=flags)["export_ordinal"])
    assert info["kernel_name"].startswith("iree_b200_matmul_bias_") and info["kernel_name"].endswith("_dtm")
    assert info["tma_map_count"] == 0 and info["binding_count"] == 4
    out, want = _run(ops, t, batch, M, N, K, family=family, extra_flags=H.MATMUL_DEVICE_MAPS, **case)
    assert np.array_equal(out, want)
