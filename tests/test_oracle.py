"""The CPU oracle against the reference's own fixtures (no GPU).

  * input generator KATs: iree_test_utils_pseudorandom_uint32 (test_utils.c:359-366)
  * matmul: the reference loops (iree-e2e-matmul-test.cc:28-121) vs exact numpy products
  * attention: the 13 golden cases extracted from tests/e2e/linalg_ext_ops/attention*.mlir
"""

import json
import os

import numpy as np
import pytest

import oracle

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "attention_golden.json")


def _minstd(state):
    # uint32 arithmetic: the product wraps mod 2^32 before the modulus (test_utils.c:359-362).
    return ((state * 48271) & 0xFFFFFFFF) % 2147483647


@pytest.mark.parametrize("t,lo,span", [("i8", -2, 5), ("bf16", -2, 5), ("f16", -4, 9)])
def test_generator_matches_reference_recurrence(t, lo, span):
    n, seed = 257, 3
    got = oracle.generate_random_matrix(t, (n,), seed)
    state, want = seed, []
    for _ in range(n):
        state = _minstd(state)
        want.append(state % span + lo)
    vals = got.astype(np.float32) if t == "i8" else oracle.to_f32(t, got)
    assert vals.tolist() == [float(w) for w in want]


@pytest.mark.parametrize("t", ["i8", "f16", "bf16"])
@pytest.mark.parametrize("shape", [(1, 1, 1), (9, 7, 5), (457, 330, 512), (64, 1, 1000)])
@pytest.mark.parametrize("nt", [False, True])
def test_matmul_oracle_exact_on_reference_inputs(t, shape, nt):
    M, N, K = shape
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (N, K) if nt else (K, N), 2)
    acc = oracle.generate_random_matrix("i32" if t == "i8" else "f32", (M, N), 3)
    a = lhs.astype(np.float64) if t == "i8" else oracle.to_f32(t, lhs).astype(np.float64)
    b = rhs.astype(np.float64) if t == "i8" else oracle.to_f32(t, rhs).astype(np.float64)
    b = b.T if nt else b
    want = a @ b
    got = oracle.matmul(t, lhs, rhs, transpose_rhs=nt)
    assert np.array_equal(got.astype(np.float64), want)
    got_acc = oracle.matmul(t, lhs, rhs, acc=acc, transpose_rhs=nt)
    assert np.array_equal(got_acc.astype(np.float64), want + acc.astype(np.float64))
    m, n = M // 2, N // 2
    assert oracle.matmul_element(t, lhs, rhs, m, n, transpose_rhs=nt) == got[m, n]


def test_batch_matmul_is_per_batch_matmul():
    lhs = oracle.generate_random_matrix("bf16", (3, 17, 32), 1)
    rhs = oracle.generate_random_matrix("bf16", (3, 32, 24), 2)
    got = oracle.matmul("bf16", lhs, rhs)
    for b in range(3):
        assert np.array_equal(got[b], oracle.matmul("bf16", lhs[b], rhs[b]))


def test_i8_extremes_wrap_free():
    lhs = np.full((4, 8192), -128, dtype=np.int8)
    rhs = np.full((8192, 4), -128, dtype=np.int8)
    assert (oracle.matmul("i8", lhs, rhs) == 128 * 128 * 8192).all()


def test_float_conversions_round_to_nearest_even():
    x = np.array([1.0, 1.00390625, 1.01171875, -2.5, 3.14159, 65504.0, 1e-8], dtype=np.float32)
    import torch
    tb = torch.from_numpy(x).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    th = torch.from_numpy(x).to(torch.float16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(oracle.f32_to_bf16_bits(x), tb)
    assert np.array_equal(oracle.f32_to_f16_bits(x), th)
    for v, b, h in zip(x, tb, th):
        assert oracle.lib().oracle_f32_to_bf16(float(v)) == b
        assert oracle.lib().oracle_f32_to_f16(float(v)) == h


def test_check_every_sampling_rule():
    # test_utils.c:56-68: smallest stride >= tot/max that does not divide the row length.
    every = oracle.calculate_check_every(8192 * 8192, 8192)
    assert every >= 8192 * 8192 // 10000 and 8192 % every != 0
    assert oracle.calculate_check_every(100, 10) == 1


def _tensor(d):
    a = np.array(d["data"], dtype=np.float64)
    shape = d["shape"]
    if a.size == 1 and int(np.prod(shape)) != 1:
        a = np.full(shape, a.reshape(-1)[0])
    return a.reshape(shape)


def _golden_cases():
    with open(GOLDEN) as f:
        return json.load(f)["cases"]


@pytest.mark.parametrize("case", _golden_cases(), ids=lambda c: c["source"] + ":" + c["name"])
@pytest.mark.parametrize("form", ["online", "online_tile2", "oneshot"])
def test_attention_golden(case, form):
    t = case["query"]["type"]
    q, k, v = (_tensor(case[x]).astype(np.float32) for x in ("query", "key", "value"))
    q, k, v = (oracle.from_f32(t, x) for x in (q, k, v))
    mask = None
    if "mask" in case:
        m = _tensor(case["mask"])
        mask = m.astype(bool) if case["mask"]["type"] == "i1" else m.astype(np.float32)
        mask = np.broadcast_to(mask, (q.shape[0], q.shape[1], k.shape[1])).copy()
    kw = dict(mask=mask, region=case["region"]["kind"], region_param=case["region"]["param"])
    if form == "oneshot":
        got = oracle.attention(t, q, k, v, case["scale"], oneshot=True, **kw)
    else:
        got = oracle.attention(t, q, k, v, case["scale"], k2_tile=2 if form == "online_tile2" else 0,
                               **kw)
    want = _tensor(case["expected"])
    np.testing.assert_allclose(oracle.to_f32(t, got), want, atol=case["atol"], rtol=0)


# ---- the oracle against the REFERENCE's own CPU kernels (oracle/_ref, built from
# ---- /root/reference by oracle/build_ref.sh; skipped where the library is absent)

_needs_ref = pytest.mark.skipif(not oracle.ref_available(),
                                reason="oracle/_ref/libiree_uk_ref.so not built (make -C oracle ref)")


@_needs_ref
@pytest.mark.parametrize("t", ["i8", "f16", "bf16"])
@pytest.mark.parametrize("generic", [False, True], ids=["arch", "generic"])
@pytest.mark.parametrize("shape", [(1, 1, 1), (9, 9, 9), (81, 19, 41), (457, 330, 512),
                                   (513, 128, 55), (1000, 4, 512)], ids=str)
def test_oracle_matches_reference_ukernels(t, generic, shape):
    """iree_uk_pack -> iree_uk_mmt4d -> iree_uk_unpack (what llvm-cpu calls with data tiling +
    ukernels) on the reference generator's inputs: bit-identical to the restated loops."""
    M, K, N = shape
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    got = oracle.ref_matmul(t, lhs, rhs, force_generic=generic)
    assert np.array_equal(got, oracle.matmul(t, lhs, rhs))


@_needs_ref
def test_oracle_matches_reference_ukernels_i8_full_range():
    g = np.random.Generator(np.random.PCG64(0xB200))
    lhs = g.integers(-128, 128, (200, 384), dtype=np.int8)
    rhs = g.integers(-128, 128, (384, 136), dtype=np.int8)
    for generic in (False, True):
        assert np.array_equal(oracle.ref_matmul("i8", lhs, rhs, force_generic=generic),
                              oracle.matmul("i8", lhs, rhs))


@pytest.mark.parametrize("t", ["f8e4m3fn", "f8e5m2"])
def test_f8_decode_known_values(t):
    """OCP 8-bit floats (math.h:550-560): landmarks of both formats."""
    tab = oracle.to_f32(t, np.arange(256, dtype=np.uint8))
    if t == "f8e4m3fn":
        assert tab[0x38] == 1.0 and tab[0x7E] == 448.0 and np.isnan(tab[0x7F]) and np.isnan(tab[0xFF])
        assert tab[0x01] == 2.0 ** -9 and tab[0x08] == 2.0 ** -6 and np.isfinite(tab[:0x7F]).all()
    else:
        assert tab[0x3C] == 1.0 and tab[0x7B] == 57344.0 and np.isinf(tab[0x7C]) and np.isnan(tab[0x7D])
        assert tab[0x01] == 2.0 ** -16 and tab[0x04] == 2.0 ** -14
    assert (tab[128:][np.isfinite(tab[128:])] == -tab[:128][np.isfinite(tab[:128])]).all()
    # The generator's small integers are exactly representable (test_utils.c:391-397).
    ints = np.arange(-2, 3, dtype=np.float32)
    assert (oracle.to_f32(t, oracle.from_f32(t, ints)) == ints).all()


@pytest.mark.parametrize("t", ["f8e4m3fn", "f8e5m2"])
@pytest.mark.parametrize("shape", [(9, 9, 9), (81, 19, 41), (1, 1000, 1000)], ids=str)
def test_f8_matmul_oracle_exact_on_reference_inputs(t, shape):
    M, K, N = shape
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    want = oracle.to_f32(t, lhs).astype(np.float64) @ oracle.to_f32(t, rhs).astype(np.float64)
    assert np.array_equal(oracle.matmul(t, lhs, rhs).astype(np.float64), want)
    got_nt = oracle.matmul(t, lhs, np.ascontiguousarray(rhs.T), transpose_rhs=True)
    assert np.array_equal(got_nt.astype(np.float64), want)


@_needs_ref
@pytest.mark.parametrize("t", ["f8e4m3fn", "f8e5m2", "f16", "bf16"])
def test_float_conversions_match_reference_math_h(t):
    """oracle.c's narrow-float conversions against the reference's own header-only
    implementation (base/internal/math.h, compiled into oracle/_ref): every code decodes to
    the same f32, and ~6M f32 bit patterns (all code values, every midpoint between
    neighbouring codes and its two f32 neighbours, a bit-pattern sweep, specials) encode
    to the same bits."""
    n = 256 if t.startswith("f8") else 65536
    bits = np.arange(n, dtype=np.uint8 if n == 256 else np.uint16)
    mine = oracle.to_f32(t, bits)
    theirs = oracle.decode_reference(t, bits[:: max(1, n // 4096)])
    sub = mine[:: max(1, n // 4096)]
    assert np.array_equal(np.isnan(sub), np.isnan(theirs))       # NaN payload / sign may differ
    assert np.array_equal(sub[~np.isnan(sub)].view(np.uint32), theirs[~np.isnan(theirs)].view(np.uint32))
    finite = np.sort(mine[np.isfinite(mine)].astype(np.float64))
    mids = ((finite[1:] + finite[:-1]) / 2).astype(np.float32)
    probes = np.concatenate([
        mine[np.isfinite(mine)], mids, np.nextafter(mids, np.float32(np.inf)),
        np.nextafter(mids, np.float32(-np.inf)),
        np.arange(0, 2 ** 32, 769, dtype=np.uint64).astype(np.uint32).view(np.float32),
        np.array([np.inf, -np.inf, np.nan, 0.0, -0.0, 1e-45, 3.4e38, -3.4e38, 464.0, 465.0, 61440.0,
                  65504.0, 65520.0], dtype=np.float32)])
    a = oracle.encode_bulk(t, probes)
    b = oracle.encode_bulk(t, probes, reference=True)
    nan_in = np.isnan(probes)
    # One documented difference: math.h:437-441 drops the mantissa when the rounding shift
    # exceeds 23 bits, so it flushes (minsub/2, minsub) to zero where round-to-nearest-even
    # (what compiled arith.truncf does, and what oracle.c does) gives the smallest
    # subnormal. bf16 keeps f32's exponent range and has no such interval.
    minsub = np.float32(finite[finite > 0][0])
    quirk = (np.abs(probes) > minsub / 2) & (np.abs(probes) < minsub) & (t != "bf16")
    same = ~nan_in & ~quirk
    assert np.array_equal(a[same], b[same])
    assert ((a[quirk] & 0x7FFF if n > 256 else a[quirk] & 0x7F) == 1).all()
    assert ((b[quirk] & 0x7FFF if n > 256 else b[quirk] & 0x7F) == 0).all()
    # NaN inputs: both must produce a NaN encoding (payloads may differ).
    assert np.isnan(oracle.to_f32(t, a[nan_in].astype(bits.dtype))).all()
    assert np.isnan(oracle.to_f32(t, b[nan_in].astype(bits.dtype))).all()


def test_matmul_bias_restates_the_unfused_ops():
    """oracle.matmul_bias = linalg.matmul, then addf/addi with bias[n], then maximumf/maxsi 0,
    then truncf — one op (one rounding) at a time."""
    for t, acc_t in (("i8", "i32"), ("bf16", "f32"), ("f16", "f32")):
        lhs = oracle.generate_random_matrix(t, (2, 9, 13), 1)
        rhs = oracle.generate_random_matrix(t, (2, 13, 6), 2)
        bias = oracle.generate_random_matrix(acc_t, (6,), 4)
        mm = oracle.matmul(t, lhs, rhs)
        assert np.array_equal(oracle.matmul_bias(t, lhs, rhs, bias), mm + bias[None, None, :])
        got = oracle.matmul_bias(t, lhs, rhs, bias, relu=True)
        assert np.array_equal(got, np.maximum(mm + bias, 0)) and (got == 0).any()
        if t != "i8":
            assert np.array_equal(oracle.matmul_bias(t, lhs, rhs, bias, relu=True, truncate=True),
                                  oracle.from_f32(t, got))
    # maximumf: NaN propagates, -inf -> 0; K == 0 leaves act(bias); addi wraps.
    l0 = np.zeros((1, 3, 0), dtype=np.uint16)
    r0 = np.zeros((1, 0, 4), dtype=np.uint16)
    b = np.array([1, -2, np.nan, -np.inf], dtype=np.float32)
    got = oracle.matmul_bias("bf16", l0, r0, b, relu=True)
    assert np.array_equal(got[0, 1], np.array([1, 0, np.nan, 0], dtype=np.float32), equal_nan=True)
    one = np.ones((1, 1, 1), dtype=np.int8)
    wrap = oracle.matmul_bias("i8", one, one, np.array([np.iinfo(np.int32).max], dtype=np.int32))
    assert wrap[0, 0, 0] == np.iinfo(np.int32).min


# ---- OCP microscaling (MXFP4): the next row of SURVEY §8(f) starts with its oracle ----------

@pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref not built")
def test_mx_element_and_scale_types_match_reference_math_h():
    """f4E2M1FN and f8E8M0FNU (math.h:606-643): every code decodes like the reference's
    iree_math_*_to_f32, and the scale encoder agrees on a bit-pattern sweep."""
    f4 = np.array([oracle.lib().oracle_f4e2m1fn_to_f32(i) for i in range(16)], dtype=np.float32)
    assert np.array_equal(f4.view(np.uint32),
                          oracle.decode_reference("f4e2m1fn", np.arange(16)).view(np.uint32))
    assert sorted(abs(float(x)) for x in f4[:8]) == [0, 0.5, 1, 1.5, 2, 3, 4, 6]
    e8 = np.array([oracle.lib().oracle_f8e8m0fnu_to_f32(i) for i in range(256)], dtype=np.float32)
    ref = oracle.decode_reference("f8e8m0fnu", np.arange(256))
    assert np.array_equal(e8[:255].view(np.uint32), ref[:255].view(np.uint32))
    assert np.isnan(e8[255]) and np.isnan(ref[255]) and e8[127] == 1.0 and e8[0] == 2.0 ** -127
    # Encoder: every exponent's first / tie / last mantissa, a random sweep, the specials.
    exps = (np.arange(256, dtype=np.uint32) << 23)[:, None]
    mans = np.array([0, 1, 0x3FFFFF, 0x400000, 0x400001, 0x7FFFFF], dtype=np.uint32)[None, :]
    pos = (exps | mans).reshape(-1)
    rng = np.random.Generator(np.random.PCG64(8))
    sweep = np.concatenate([pos, pos | 0x80000000,
                            rng.integers(0, 1 << 32, 200000, dtype=np.uint64).astype(np.uint32)])
    values = sweep.view(np.float32)
    assert np.array_equal(oracle.encode_bulk("f8e8m0fnu", values),
                          oracle.encode_bulk("f8e8m0fnu", values, reference=True))
    for v, want in ((1.0, 127), (1.49, 127), (1.5, 128), (3.0, 129), (0.0, 0), (-4.0, 0),
                    (np.inf, 255), (np.nan, 255), (2.0 ** -127, 0), (3.0e38, 255)):
        assert oracle.lib().oracle_f32_to_f8e8m0fnu(v) == want, v


def test_mx_generator_follows_the_reference_ranges():
    """test_utils.c:376-382 (UINT_8 = random bytes 0..255) and :421-424 (scales from 1..8,
    stored through iree_math_f32_to_f8e8m0fnu, :345-346)."""
    raw = oracle.generate_random_matrix("u8", (64, 64), 1)
    assert raw.dtype == np.uint8 and raw.min() == 0 and raw.max() == 255
    state, want = 1, []
    for _ in range(8):
        state = (state * 48271) % (1 << 32) % 2147483647
        want.append(state % 256)
    assert list(raw.reshape(-1)[:8]) == want
    scales = oracle.generate_random_matrix("f8e8m0fnu", (64, 8), 2)
    assert set(np.unique(scales)) == {127, 128, 129, 130}    # 1, 2, 4, 8
    state = 2
    state = (state * 48271) % (1 << 32) % 2147483647
    first = state % 8 + 1
    assert scales.reshape(-1)[0] == oracle.lib().oracle_f32_to_f8e8m0fnu(float(first))


@pytest.mark.parametrize("shape", [(1, 32, 1), (3, 64, 5), (17, 96, 9), (64, 256, 48)], ids=str)
@pytest.mark.parametrize("accumulate", [False, True], ids=["", "acc"])
def test_matmul_mxfp4_restates_the_scaled_generic(shape, accumulate):
    """generate_code_mx.py:153-172: acc += scaling_extf(l, ls) * scaling_extf(r, rs) over
    (block, element). With the generator's inputs every partial sum is exact in f32, so an
    independent float64 evaluation of the same formula must agree bit for bit."""
    M, K, N = shape
    lhs = oracle.generate_random_matrix("u8", (M, K // 2), 1)
    ls = oracle.generate_random_matrix("f8e8m0fnu", (M, K // 32), 2)
    rhs = oracle.generate_random_matrix("u8", (N, K // 2), 3)
    rs = oracle.generate_random_matrix("f8e8m0fnu", (N, K // 32), 4)
    acc = oracle.generate_random_matrix("f32", (M, N), 5) if accumulate else None
    got = oracle.matmul_mxfp4(lhs, ls, rhs, rs, acc=acc)
    l64 = oracle.mx_decode(lhs, ls).astype(np.float64)
    r64 = oracle.mx_decode(rhs, rs).astype(np.float64)
    want = l64 @ r64.T + (acc.astype(np.float64) if accumulate else 0.0)
    assert got.dtype == np.float32 and np.array_equal(got.astype(np.float64), want)


def test_matmul_mxfp4_packing_and_special_scales():
    # One row, one block: element 0 = 6.0 (low nibble of byte 0), element 1 = -0.5 (high).
    lhs = np.zeros((1, 16), dtype=np.uint8)
    lhs[0, 0] = 0x7 | (0x9 << 4)
    rhs = np.zeros((2, 16), dtype=np.uint8)
    rhs[0, 0] = 0x2            # picks element 0: 1.0
    rhs[1, 0] = 0x2 << 4       # picks element 1: 1.0
    one = np.full((1, 1), 127, dtype=np.uint8)
    got = oracle.matmul_mxfp4(lhs, one, rhs, np.full((2, 1), 127, dtype=np.uint8))
    assert got.tolist() == [[6.0, -0.5]]
    # Scales multiply per (row, block): 2^3 on the lhs row, 2^-2 on rhs row 1.
    got = oracle.matmul_mxfp4(lhs, np.full((1, 1), 130, dtype=np.uint8), rhs,
                              np.array([[127], [125]], dtype=np.uint8))
    assert got.tolist() == [[48.0, -1.0]]
    # 0xFF is NaN and poisons the rows / columns it scales (arith.scaling_extf).
    got = oracle.matmul_mxfp4(lhs, one, rhs, np.array([[127], [255]], dtype=np.uint8))
    assert got[0, 0] == 6.0 and np.isnan(got[0, 1])


# ---- linalg.mmt4d on data-tiled operands ---------------------------------

MMT4D_CASES = [
    # M, K, N (logical), M0, N0, K0
    (37, 50, 29, 16, 16, 2), (64, 64, 64, 8, 8, 1), (5, 3, 7, 16, 16, 1), (100, 40, 33, 8, 32, 2),
    (33, 17, 40, 1, 1, 1), (20, 24, 90, 7, 5, 3), (48, 64, 48, 16, 16, 16),
]


def _mmt4d_inputs(t, case):
    M, K, N, M0, N0, K0 = case
    lhs = oracle.generate_random_matrix(t, (M, K), 1)
    rhs = oracle.generate_random_matrix(t, (K, N), 2)
    l4 = oracle.pack(lhs, M0, K0)
    r4 = oracle.pack(np.ascontiguousarray(rhs.T), N0, K0)
    acc = oracle.generate_random_matrix("i32" if t == "i8" else "f32",
                                        (l4.shape[0], r4.shape[0], M0, N0), 3)
    return lhs, rhs, l4, r4, acc


@pytest.mark.parametrize("t", ["i8", "f16", "bf16"])
@pytest.mark.parametrize("case", MMT4D_CASES, ids=str)
def test_mmt4d_is_the_matmul_of_the_unpacked_operands(t, case):
    """pack -> mmt4d -> unpack == matmul (the identity the reference's e2e_matmul_benchmark.c
    relies on), and the accumulate form adds the packed accumulator."""
    M, K, N = case[:3]
    lhs, rhs, l4, r4, acc = _mmt4d_inputs(t, case)
    out = oracle.mmt4d(t, l4, r4)
    assert out.shape == (l4.shape[0], r4.shape[0], case[3], case[4])
    assert np.array_equal(oracle.unpack(out, M, N), oracle.matmul(t, lhs[None], rhs[None])[0])
    # Padding rows / columns of the tiles hold exact zeros.
    full = oracle.unpack(out, out.shape[0] * case[3], out.shape[1] * case[4])
    assert (full[M:] == 0).all() and (full[:, N:] == 0).all()
    with_acc = oracle.mmt4d(t, l4, r4, acc=acc)
    assert np.array_equal(with_acc, out + acc)      # exact: small integers in f32 / i32


@_needs_ref
@pytest.mark.parametrize("t", ["i8", "f16", "bf16", "f32"])
@pytest.mark.parametrize("generic", [False, True], ids=["arch", "generic"])
@pytest.mark.parametrize("case", MMT4D_CASES, ids=str)
def test_mmt4d_matches_the_reference_microkernel(t, generic, case):
    """Pins oracle.mmt4d on the reference's own iree_uk_mmt4d (oracle/_ref), arch-specific
    tile functions where the tile shape has one (16x16x2, 8x8x1, …) and the generic C tiles
    (mmt4d_tile_generic.c) everywhere."""
    _, _, l4, r4, acc = _mmt4d_inputs(t, case)
    for a in (None, acc):
        assert np.array_equal(oracle.mmt4d(t, l4, r4, acc=a),
                              oracle.ref_mmt4d(t, l4, r4, acc=a, force_generic=generic))


def test_mmt4d_i8_full_range_wraps_like_the_reference():
    rng = np.random.Generator(np.random.PCG64(5))
    l4 = rng.integers(-128, 128, (3, 40, 16, 2), dtype=np.int8)
    r4 = rng.integers(-128, 128, (2, 40, 16, 2), dtype=np.int8)
    acc = np.full((3, 2, 16, 16), np.iinfo(np.int32).max, dtype=np.int32)
    out = oracle.mmt4d("i8", l4, r4, acc=acc)
    want = (np.einsum("akmc,bknc->abmn", l4.astype(np.int64), r4.astype(np.int64)) + acc)
    assert np.array_equal(out, want.astype(np.uint64).astype(np.uint32).view(np.int32).reshape(out.shape))
    if oracle.ref_available():
        assert np.array_equal(out, oracle.ref_mmt4d("i8", l4, r4, acc=acc))


def test_mmt4d_empty_reduction():
    l4 = np.zeros((2, 0, 8, 4), dtype=np.uint16)
    r4 = np.zeros((3, 0, 8, 4), dtype=np.uint16)
    assert (oracle.mmt4d("bf16", l4, r4) == 0).all()
    acc = np.full((2, 3, 8, 8), 2.5, dtype=np.float32)
    assert np.array_equal(oracle.mmt4d("bf16", l4, r4, acc=acc), acc)
