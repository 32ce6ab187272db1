"""Parity of the CUDA attention dispatches with the oracle, through the C-ABI (B200 only).

* the reference's own golden tensors (tests/e2e/linalg_ext_ops/attention*.mlir, atol 1e-4)
  through the generic kernel (f32, masks);
* the tcgen05 prefill kernel and the split-KV decode kernels against the oracle's restated
  llvm-cpu arithmetic on seeded random inputs, tolerance from BASELINE.json north_star
  (1e-3 relative for f16, 1e-2 for bf16), including ragged M / K2;
* at the BASELINE Llama-3-8B shapes (32 heads, D=128, seq 4096 prefill / 8192-KV decode):
  sampled rows against the oracle plus a size-independent property (uniform V => O == V row).
"""

import json
import os

import numpy as np
import pytest

import oracle
from iree_b200 import hal as H

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "attention_golden.json")


def _tensor(d):
    a = np.array(d["data"], dtype=np.float64)
    shape = d["shape"]
    if a.size == 1 and int(np.prod(shape)) != 1:
        a = np.full(shape, a.reshape(-1)[0])
    return a.reshape(shape)


def _golden_cases():
    with open(GOLDEN) as f:
        cases = json.load(f)["cases"]
    # The region is part of the op's body; the matcher only takes the identity region.
    return [c for c in cases if c["region"]["kind"] == "identity"]


def _gpu_attention(ops, t, q, k, v, scale, mask=None, flags=0):
    B, M, K1 = q.shape
    K2, N = k.shape[1], v.shape[2]
    out = np.zeros((B, M, N), dtype=q.dtype)
    mask_kind, mask_ptr = H.ATTENTION_MASK_NONE, 0
    if mask is not None:
        if mask.dtype == np.bool_:
            mask = np.ascontiguousarray(mask, dtype=np.uint8)
            mask_kind = H.ATTENTION_MASK_BOOL
        else:
            mask = np.ascontiguousarray(mask, dtype=np.float32)
            mask_kind = H.ATTENTION_MASK_F32
        mask_ptr = mask.ctypes.data
    q, k, v = (np.ascontiguousarray(x) for x in (q, k, v))
    ops.attention_host(t, B, M, K1, K2, N, scale, q.ctypes.data, k.ctypes.data, v.ctypes.data,
                       out.ctypes.data, mask_ptr=mask_ptr, mask_kind=mask_kind, flags=flags)
    return out


@pytest.mark.parametrize("case", _golden_cases(), ids=lambda c: c["source"] + ":" + c["name"])
def test_reference_golden_vectors(ops, case):
    q, k, v = (_tensor(case[x]).astype(np.float32) for x in ("query", "key", "value"))
    mask = None
    if "mask" in case:
        m = _tensor(case["mask"])
        mask = m.astype(bool) if case["mask"]["type"] == "i1" else m.astype(np.float32)
        mask = np.broadcast_to(mask, (q.shape[0], q.shape[1], k.shape[1])).copy()
    got = _gpu_attention(ops, "f32", q, k, v, case["scale"], mask=mask)
    np.testing.assert_allclose(got, _tensor(case["expected"]), atol=case["atol"], rtol=0)


def _random_qkv(t, B, M, K2, D, seed, N=None):
    g = np.random.Generator(np.random.PCG64(0xB200 + seed))
    N = D if N is None else N
    q = oracle.from_f32(t, np.clip(g.standard_normal((B, M, D)), -4, 4).astype(np.float32))
    k = oracle.from_f32(t, np.clip(g.standard_normal((B, K2, D)), -4, 4).astype(np.float32))
    v = oracle.from_f32(t, np.clip(g.standard_normal((B, K2, N)), -4, 4).astype(np.float32))
    return q, k, v


# Bound on max|got - want| as a fraction of max|want| ITSELF (no absolute floor): the f32
# accumulate is compared after one rounding to the element type (half an ulp = 2^-9 of the
# value for bf16, 2^-12 for f16) on top of the P -> V-type truncation both sides perform in a
# different K2 tile order. north_star's 1e-2 (bf16) / 1e-3 (f16) are relative to the *element*
# for a well-conditioned sum; softmax outputs are averages of ~K2 values of either sign, so
# the bound is stated against the largest output: 2e-2 * max|want| for bf16, 2e-3 for f16
# (CheckOps.td:139 isclose rule with atol = rtol * max|want|), plus a per-row cosine floor.
RTOL_MAX = {"f16": 2e-3, "bf16": 2e-2}
COS_MIN = {"f16": 1.0 - 1e-5, "bf16": 1.0 - 2e-4}
ERRLOG = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out",
                      "attention_parity_errors.jsonl")


def _assert_close(t, got, want, label=None):
    g, w = oracle.to_f32(t, got).astype(np.float64), oracle.to_f32(t, want).astype(np.float64)
    assert np.isfinite(g).all()
    scale = np.abs(w).max()
    err = np.abs(g - w).max()
    g2, w2 = g.reshape(-1, g.shape[-1]), w.reshape(-1, w.shape[-1])
    den = np.linalg.norm(g2, axis=1) * np.linalg.norm(w2, axis=1)
    cos = np.where(den > 0, (g2 * w2).sum(axis=1) / np.where(den > 0, den, 1), 1.0).min()
    if label is not None:
        try:
            os.makedirs(os.path.dirname(ERRLOG), exist_ok=True)
            with open(ERRLOG, "a") as f:
                f.write(json.dumps({"case": label, "type": t, "max_abs_err": float(err),
                                    "max_abs_want": float(scale), "rel_to_max": float(err / scale),
                                    "rms_want": float(np.sqrt((w * w).mean())),
                                    "min_row_cosine": float(cos)}) + "\n")
        except OSError:
            pass
    assert err <= RTOL_MAX[t] * scale, (err, scale, err / scale)
    assert cos >= COS_MIN[t], cos


@pytest.mark.parametrize("t", ["f16", "bf16"])
@pytest.mark.parametrize("exp2", ["default", "mufu", "poly25", "poly50"])
@pytest.mark.parametrize("shape", [(1, 128, 128), (2, 256, 256), (3, 384, 640), (2, 200, 300),
                                   (1, 130, 77), (4, 512, 1024), (1, 17, 129)], ids=str)
def test_prefill_kernel_vs_oracle(ops, t, exp2, shape):
    """Every exp2 split of the prefill kernel (all on the MUFU, 1 or 2 of 4 as a degree-3
    polynomial on the FMA pipe) against the oracle, same tolerance."""
    B, M, K2 = shape
    q, k, v = _random_qkv(t, B, M, K2, 128, seed=M + K2)
    scale = 1.0 / np.sqrt(128.0)
    flag = {"default": 0, "mufu": H.ATTENTION_EXP2_MUFU_ONLY, "poly25": H.ATTENTION_EXP2_POLY_25,
            "poly50": H.ATTENTION_EXP2_POLY_50}[exp2]
    got = _gpu_attention(ops, t, q, k, v, scale, flags=H.ATTENTION_FORCE_PREFILL | flag)
    want = oracle.attention(t, q, k, v, scale, k2_tile=128)
    _assert_close(t, got, want)


@pytest.mark.parametrize("t", ["f16", "bf16"])
def test_prefill_kernel_rescale_path(ops, t):
    """Row maxima that keep growing across K2 blocks force the lazy O correction."""
    B, M, K2 = 1, 128, 1024
    q, k, v = _random_qkv(t, B, M, K2, 128, seed=7)
    kf = oracle.to_f32(t, k)
    kf *= (1.0 + np.arange(K2, dtype=np.float32) / 64.0)[None, :, None]
    k = oracle.from_f32(t, kf)
    scale = 1.0 / np.sqrt(128.0)
    got = _gpu_attention(ops, t, q, k, v, scale, flags=H.ATTENTION_FORCE_PREFILL)
    want = oracle.attention(t, q, k, v, scale, k2_tile=128)
    _assert_close(t, got, want)


@pytest.mark.parametrize("t", ["f16", "bf16"])
@pytest.mark.parametrize("form", ["split", "cluster"])
@pytest.mark.parametrize("shape", [(1, 1, 256), (2, 1, 1000), (4, 3, 4096), (32, 1, 8192),
                                   (1, 16, 77), (3, 2, 5), (2, 1, 8), (5, 1, 263), (1, 1, 32 * 8 * 7 + 1)],
                         ids=str)
def test_decode_kernels_vs_oracle(ops, t, form, shape):
    """Both decode forms: split-KV + merge (two dispatches) and the single-dispatch cluster
    kernel (ragged splits, splits with no rows at all, partial ring stages)."""
    B, M, K2 = shape
    q, k, v = _random_qkv(t, B, M, K2, 128, seed=K2)
    scale = 1.0 / np.sqrt(128.0)
    before = ops.dispatch_count
    got = _gpu_attention(ops, t, q, k, v, scale,
                         flags={"split": H.ATTENTION_FORCE_DECODE_SPLIT,
                                "cluster": H.ATTENTION_FORCE_DECODE_CLUSTER}[form])
    assert ops.dispatch_count - before == {"split": 2, "cluster": 1}[form]
    want = oracle.attention(t, q, k, v, scale, k2_tile=256)
    _assert_close(t, got, want)


@pytest.mark.parametrize("t", ["f16", "bf16"])
def test_decode_cluster_growing_maxima(ops, t):
    """Scores that keep growing along K2 exercise every rescale of the per-half-warp states
    and the two merge levels (shared memory, then distributed shared memory)."""
    B, M, K2 = 20, 1, 2048
    q, k, v = _random_qkv(t, B, M, K2, 128, seed=11)
    kf = oracle.to_f32(t, k) * (0.25 + np.arange(K2, dtype=np.float32) / 512.0)[None, :, None]
    k = oracle.from_f32(t, kf)
    scale = 1.0 / np.sqrt(128.0)
    got = _gpu_attention(ops, t, q, k, v, scale, flags=H.ATTENTION_FORCE_DECODE_CLUSTER)
    want = oracle.attention(t, q, k, v, scale, k2_tile=256)
    _assert_close(t, got, want)


@pytest.mark.parametrize("t", ["f16", "bf16"])
@pytest.mark.parametrize("mask_kind", [None, "bool", "f32"])
@pytest.mark.parametrize("shape", [(2, 5, 16, 40, 24), (1, 33, 64, 300, 64), (1, 4, 130, 129, 200)],
                         ids=str)
def test_generic_kernel_vs_oracle(ops, t, mask_kind, shape):
    B, M, K1, K2, N = shape
    g = np.random.Generator(np.random.PCG64(0xB200 + K2))
    q, k, v = _random_qkv(t, B, M, K2, K1, seed=K1, N=N)
    mask = None
    if mask_kind == "bool":
        mask = g.random((B, M, K2)) > 0.3
        mask[:, :, 0] = True  # no fully masked row (NaN by definition, TileAttention.cpp:141)
    elif mask_kind == "f32":
        mask = g.standard_normal((B, M, K2)).astype(np.float32)
    scale = 1.0 / np.sqrt(float(K1))
    got = _gpu_attention(ops, t, q, k, v, scale, mask=mask, flags=H.ATTENTION_FORCE_GENERIC)
    want = oracle.attention(t, q, k, v, scale, mask=mask, k2_tile=128)
    _assert_close(t, got, want)


def test_selection_guards(ops):
    exe = ops.executable
    names = lambda sel: [exe.export_info(d["export_ordinal"])["kernel_name"] for d in sel["dispatches"]]
    s = ops.select_attention("f16", 32, 4096, 128, 4096, 128, 0.088, flags=H.ATTENTION_NO_TAIL_SPLIT)
    assert names(s) == ["iree_b200_attention_prefill_f16"]
    assert s["dispatches"][0]["workgroup_count"] == [16, 32, 1]
    # 512 equal units on 148 SMs: 444 run whole, the last 68 are cut in two over K2.
    s = ops.select_attention("f16", 32, 4096, 128, 4096, 128, 0.088)
    assert names(s) == ["iree_b200_attention_prefill_split_f16", "iree_b200_attention_prefill_merge_f16"]
    assert s["dispatches"][0]["workgroup_count"] == [444 + 68 * 2, 1, 1]
    assert s["dispatches"][0]["constants"][4:6] == [444, 2]
    s = ops.select_attention("f16", 32, 1, 128, 8192, 128, 0.088)
    assert names(s) == ["iree_b200_attention_decode_cluster_f16"]
    assert s["dispatches"][0]["workgroup_count"] == [8, 1, 32] and s["workspace_bytes"] == 0
    # Few (batch, m) rows: 8 CTAs per row would leave SMs idle -> split-KV + merge.
    s = ops.select_attention("f16", 4, 1, 128, 8192, 128, 0.088)
    assert names(s) == ["iree_b200_attention_decode_split_f16", "iree_b200_attention_decode_merge_f16"]
    assert s["workspace_bytes"] == 4 * s["dispatches"][0]["workgroup_count"][0] * 130 * 4
    assert names(ops.select_attention("f16", 32, 1, 128, 128, 128, 0.088))[0].endswith("split_f16")
    assert names(ops.select_attention("bf16", 2, 64, 64, 64, 64, 0.125)) == \
        ["iree_b200_attention_generic_bf16"]
    assert names(ops.select_attention("f16", 2, 256, 128, 256, 128, 0.1,
                                      mask_kind=H.ATTENTION_MASK_F32)) == \
        ["iree_b200_attention_generic_f16"]
    assert names(ops.select_attention("f16", 2, 256, 128, 256, 128, 0.1, alignment=2)) == \
        ["iree_b200_attention_generic_f16"]
    assert ops.select_attention("f16", 0, 4, 8, 8, 8, 1.0)["dispatches"] == []
    with pytest.raises(H.Status):
        ops.select_attention("i8", 1, 4, 8, 8, 8, 1.0)
    with pytest.raises(H.Status):
        ops.select_attention("f16", 1, 4, 1024, 8, 8, 1.0)
    with pytest.raises(H.Status):
        ops.select_attention("f16", 1, 64, 64, 64, 64, 1.0, flags=H.ATTENTION_FORCE_PREFILL)


@pytest.mark.parametrize("t", ["f16", "bf16"])
def test_llama_prefill_shape_sampled(ops, t):
    """BASELINE configs[3] prefill: 32 heads, D=128, seq 4096."""
    B, M, K2 = 32, 4096, 4096
    q, k, v = _random_qkv(t, B, M, K2, 128, seed=99)
    scale = 1.0 / np.sqrt(128.0)
    got = _gpu_attention(ops, t, q, k, v, scale)
    # Oracle on sampled rows of sampled heads (the full oracle is ~275 GFLOP of scalar code).
    rows = np.array([0, 1, 127, 128, 255, 256, 1000, 2047, 2048, 4095])
    for b in (0, 13, 31):
        want = oracle.attention(t, q[b:b + 1, rows], k[b:b + 1], v[b:b + 1], scale, k2_tile=128)
        _assert_close(t, got[b:b + 1, rows], want, label=f"llama_prefill_32x4096x4096x128 head {b}")
    # Property: softmax weights sum to one, so a V that is constant along K2 comes back.
    vconst = np.broadcast_to(v[:, :1, :], v.shape).copy()
    got2 = _gpu_attention(ops, t, q, k, vconst, scale)
    ref = np.broadcast_to(oracle.to_f32(t, vconst[:, :1, :]), got2.shape)
    # every output row is V's row rounded once more: within one ulp of the element type
    assert np.abs(oracle.to_f32(t, got2) - ref).max() <= {"f16": 2.0 ** -10, "bf16": 2.0 ** -7}[t] * np.abs(ref).max()


@pytest.mark.parametrize("t", ["f16", "bf16"])
def test_llama_decode_shape(ops, t):
    """BASELINE configs[3] decode: seq 1 over an 8K KV cache, 32 heads."""
    B, M, K2 = 32, 1, 8192
    q, k, v = _random_qkv(t, B, M, K2, 128, seed=5)
    scale = 1.0 / np.sqrt(128.0)
    got = _gpu_attention(ops, t, q, k, v, scale)
    want = oracle.attention(t, q, k, v, scale, k2_tile=256)
    _assert_close(t, got, want, label="llama_decode_32x1x8192x128")


@pytest.mark.parametrize("t", ["f16", "bf16"])
@pytest.mark.parametrize("shape", [(4, 512, 1024), (2, 300, 1100), (5, 700, 2048), (1, 256, 4096),
                                   (3, 130, 1025)], ids=str)
def test_prefill_tail_split(ops, t, shape):
    """The tail-split form of the prefill kernel (units of the last, partly filled wave cut
    over K2 + merge dispatch): selected for these shapes, same tolerance against the oracle,
    and within rounding of the unsplit kernel. Ragged M and K2, 2- and 4-way splits."""
    B, M, K2 = shape
    q, k, v = _random_qkv(t, B, M, K2, 128, seed=3 * M + K2)
    scale = 1.0 / np.sqrt(128.0)
    sel = ops.select_attention(t, B, M, 128, K2, 128, scale, flags=H.ATTENTION_FORCE_PREFILL)
    exe = ops.executable
    names = [exe.export_info(d["export_ordinal"])["kernel_name"] for d in sel["dispatches"]]
    assert names == [f"iree_b200_attention_prefill_split_{t}", f"iree_b200_attention_prefill_merge_{t}"]
    split = sel["dispatches"][0]["constants"][5]
    assert split >= 2 and sel["workspace_bytes"] == sel["dispatches"][1]["workgroup_count"][1] * split * 256 * 130 * 4
    got = _gpu_attention(ops, t, q, k, v, scale, flags=H.ATTENTION_FORCE_PREFILL)
    want = oracle.attention(t, q, k, v, scale, k2_tile=128)
    _assert_close(t, got, want)
    whole = _gpu_attention(ops, t, q, k, v, scale,
                           flags=H.ATTENTION_FORCE_PREFILL | H.ATTENTION_NO_TAIL_SPLIT)
    _assert_close(t, got, whole)


@pytest.mark.parametrize("case", [
    # type, B, M, K1, K2, N, mask
    ("f16", 5, 130, 128, 300, 128, None),     # tcgen05 prefill kernel, one entry per slice
    ("bf16", 3, 512, 128, 1025, 128, None),   # slices small enough for the tail split
    ("f32", 4, 33, 24, 50, 40, "bool"),       # generic kernel, the mask is sliced too
    ("f32", 3, 20, 16, 70, 8, "f32"),
], ids=str)
def test_pipelined_host_call(ops, case):
    """iree_b200_ops_attention_host in slices of B on three streams (upload / dispatch /
    download): same result as the whole-problem call, one selection per slice."""
    t, B, M, K1, K2, N, mask_kind = case
    g = np.random.Generator(np.random.PCG64(0xB200 + B + M))
    if t == "f32":
        q, k, v = (g.standard_normal(s).astype(np.float32) for s in ((B, M, K1), (B, K2, K1), (B, K2, N)))
    else:
        q, k, v = _random_qkv(t, B, M, K2, K1, seed=B * M, N=N)
    mask = None
    if mask_kind == "bool":
        mask = g.random((B, M, K2)) > 0.3
        mask[:, :, 0] = True
    elif mask_kind == "f32":
        mask = g.standard_normal((B, M, K2)).astype(np.float32)
    scale = 1.0 / np.sqrt(float(K1))
    whole = _gpu_attention(ops, t, q, k, v, scale, mask=mask)
    before = ops.dispatch_count
    ops.set_host_pipeline(0, 2048, 2048)    # threshold 0: one batch entry per slice
    try:
        sliced = _gpu_attention(ops, t, q, k, v, scale, mask=mask)
    finally:
        ops.set_host_pipeline(16 << 20, 2048, 2048)
    per_slice = len(ops.select_attention(t, 1, M, K1, K2, N, scale,
                                         mask_kind={None: 0, "bool": 1, "f32": 2}[mask_kind])["dispatches"])
    assert ops.dispatch_count - before == B * per_slice
    want = oracle.attention(t, q, k, v, scale, mask=mask, k2_tile=128)
    if t == "f32":
        np.testing.assert_allclose(sliced, want, atol=1e-4, rtol=0)
        assert np.array_equal(sliced, whole)   # same kernel, per-entry arithmetic
    else:
        _assert_close(t, sliced, want)
        _assert_close(t, sliced, whole)


# ---------------------------------------------------------------------------------------
# Masks on the tcgen05 prefill kernel (the op's mask operand, LinalgExtOps.td:967-1071;
# AggregatedOpInterfaceImpl.cpp:239-291): causal computed in the kernel, i1 operand.
# ---------------------------------------------------------------------------------------

def _causal(B, M, K2):
    m = np.arange(M)[:, None]
    k2 = np.arange(K2)[None, :]
    return np.broadcast_to(k2 <= m + (K2 - M), (B, M, K2)).copy()


def _kernel_of(ops, t, B, M, K2, mask_kind, flags=0):
    sel = ops.select_attention(t, B, M, 128, K2, 128, 0.088, mask_kind=mask_kind, flags=flags)
    return [ops.executable.export_info(d["export_ordinal"])["kernel_name"] for d in sel["dispatches"]]


@pytest.mark.parametrize("t", ["f16", "bf16"])
@pytest.mark.parametrize("shape", [(1, 256, 256), (2, 512, 512), (3, 384, 640), (2, 200, 300),
                                   (1, 130, 130), (2, 1024, 1024), (1, 17, 129), (5, 768, 768)],
                         ids=str)
def test_prefill_causal_vs_oracle_with_the_same_mask(ops, t, shape):
    """MASK_CAUSAL (no operand; masked K2 blocks skipped, heaviest units first) must equal the
    oracle run with the explicit i1 mask k2 <= m + (K2 - M)."""
    B, M, K2 = shape
    q, k, v = _random_qkv(t, B, M, K2, 128, seed=3 * M + K2)
    scale = 1.0 / np.sqrt(128.0)
    assert _kernel_of(ops, t, B, M, K2, H.ATTENTION_MASK_CAUSAL) == [f"iree_b200_attention_prefill_{t}"]
    out = np.zeros((B, M, 128), dtype=q.dtype)
    ops.attention_host(t, B, M, 128, K2, 128, scale, q.ctypes.data, k.ctypes.data, v.ctypes.data,
                       out.ctypes.data, mask_kind=H.ATTENTION_MASK_CAUSAL)
    want = oracle.attention(t, q, k, v, scale, mask=_causal(B, M, K2), k2_tile=128)
    _assert_close(t, out, want, label=f"causal prefill {shape}")
    # ... and the generic kernel computes the same mask (mask_kind 3).
    gen = np.zeros_like(out)
    ops.attention_host(t, B, M, 128, K2, 128, scale, q.ctypes.data, k.ctypes.data, v.ctypes.data,
                       gen.ctypes.data, mask_kind=H.ATTENTION_MASK_CAUSAL,
                       flags=H.ATTENTION_FORCE_GENERIC)
    _assert_close(t, gen, want)


@pytest.mark.parametrize("t", ["f16", "bf16"])
@pytest.mark.parametrize("shape", [(1, 256, 256), (2, 384, 640), (2, 200, 304), (1, 130, 144),
                                   (3, 512, 1024)], ids=str)
@pytest.mark.parametrize("pattern", ["causal", "random", "banded"])
def test_prefill_i1_mask_operand_vs_oracle(ops, t, shape, pattern):
    """An i1 mask operand on the tcgen05 kernel (iree_b200_attention_prefill_mask_*)."""
    B, M, K2 = shape
    q, k, v = _random_qkv(t, B, M, K2, 128, seed=M + 5 * K2)
    scale = 1.0 / np.sqrt(128.0)
    g = np.random.Generator(np.random.PCG64(0xB200 + M))
    if pattern == "causal":
        mask = _causal(B, M, K2)
    elif pattern == "random":
        mask = g.random((B, M, K2)) < 0.7
        mask[:, :, 0] = True                       # every row keeps at least one key
    else:
        m, k2 = np.arange(M)[:, None], np.arange(K2)[None, :]
        mask = np.broadcast_to(np.abs(k2 - m * K2 // M) <= 96, (B, M, K2)).copy()
    assert _kernel_of(ops, t, B, M, K2, H.ATTENTION_MASK_BOOL) == [f"iree_b200_attention_prefill_mask_{t}"]
    got = _gpu_attention(ops, t, q, k, v, scale, mask=mask)
    want = oracle.attention(t, q, k, v, scale, mask=mask, k2_tile=128)
    _assert_close(t, got, want, label=f"i1-mask prefill {pattern} {shape}")


def test_mask_selection_guards(ops):
    # K2 not a multiple of 16: the mask rows cannot be read 16 bytes at a time -> generic.
    assert _kernel_of(ops, "f16", 2, 256, 300, H.ATTENTION_MASK_BOOL) == ["iree_b200_attention_generic_f16"]
    # additive f32 masks stay on the generic kernel.
    assert _kernel_of(ops, "f16", 2, 256, 256, H.ATTENTION_MASK_F32) == ["iree_b200_attention_generic_f16"]
    # causal with one query row is no mask at all (decode kernels); with a few rows: generic.
    assert _kernel_of(ops, "f16", 32, 1, 8192, H.ATTENTION_MASK_CAUSAL) == ["iree_b200_attention_decode_cluster_f16"]
    assert _kernel_of(ops, "f16", 2, 8, 512, H.ATTENTION_MASK_CAUSAL) == ["iree_b200_attention_generic_f16"]
    with pytest.raises(H.Status):
        ops.select_attention("f16", 1, 512, 128, 256, 128, 0.088, mask_kind=H.ATTENTION_MASK_CAUSAL)
    # causal: 1-D grid over units, no tail split.
    sel = ops.select_attention("f16", 32, 4096, 128, 4096, 128, 0.088, mask_kind=H.ATTENTION_MASK_CAUSAL)
    assert len(sel["dispatches"]) == 1 and sel["dispatches"][0]["workgroup_count"] == [512, 1, 1]
    assert sel["dispatches"][0]["constants"][4] == 1


@pytest.mark.parametrize("t", ["f16", "bf16"])
def test_llama_prefill_shape_causal_sampled(ops, t):
    """BASELINE configs[3] with the mask real Llama prefill uses: 32 heads x 4096 x 4096,
    sampled rows against the oracle with the explicit causal mask."""
    B, M, K2 = 32, 4096, 4096
    q, k, v = _random_qkv(t, B, M, K2, 128, seed=11)
    scale = 1.0 / np.sqrt(128.0)
    out = np.zeros((B, M, 128), dtype=q.dtype)
    ops.attention_host(t, B, M, 128, K2, 128, scale, q.ctypes.data, k.ctypes.data, v.ctypes.data,
                       out.ctypes.data, mask_kind=H.ATTENTION_MASK_CAUSAL)
    rows = np.array([0, 1, 127, 128, 255, 256, 1000, 2047, 2048, 3000, 4095])
    for b in (0, 13, 31):
        mask = (np.arange(K2)[None, :] <= rows[:, None])[None]
        want = oracle.attention(t, q[b:b + 1, rows], k[b:b + 1], v[b:b + 1], scale, mask=mask,
                                k2_tile=128)
        _assert_close(t, out[b:b + 1, rows], want, label=f"llama causal prefill head {b}")


# ---------------------------------------------------------------------------------------
# `*_dtm`: the tcgen05 prefill kernel without tensor-map parameters (tensormap.replace on the
# device), i.e. the export an unmodified IREE CUDA driver can launch.
# ---------------------------------------------------------------------------------------

@pytest.mark.parametrize("t", ["f16", "bf16"])
@pytest.mark.parametrize("causal", [False, True], ids=["unmasked", "causal"])
@pytest.mark.parametrize("shape", [(1, 256, 256), (3, 384, 640), (2, 200, 300), (5, 768, 768),
                                   (40, 512, 1024)], ids=str)
def test_prefill_with_device_built_tensor_maps_is_the_same_kernel(ops, t, causal, shape):
    """Same bits as the host-fed export (same code after the maps exist), same tolerance
    against the oracle; more CTAs than SMs, so slots are reused within one launch."""
    B, M, K2 = shape
    q, k, v = _random_qkv(t, B, M, K2, 128, seed=5 * M + K2)
    scale = 1.0 / np.sqrt(128.0)
    kind = H.ATTENTION_MASK_CAUSAL if causal else H.ATTENTION_MASK_NONE
    assert _kernel_of(ops, t, B, M, K2, kind, flags=H.ATTENTION_DEVICE_MAPS) == \
        [f"iree_b200_attention_prefill_{t}_dtm"]
    got = np.zeros((B, M, 128), dtype=q.dtype)
    ops.attention_host(t, B, M, 128, K2, 128, scale, q.ctypes.data, k.ctypes.data, v.ctypes.data,
                       got.ctypes.data, mask_kind=kind, flags=H.ATTENTION_DEVICE_MAPS)
    host_fed = np.zeros_like(got)
    ops.attention_host(t, B, M, 128, K2, 128, scale, q.ctypes.data, k.ctypes.data, v.ctypes.data,
                       host_fed.ctypes.data, mask_kind=kind, flags=H.ATTENTION_NO_TAIL_SPLIT)
    assert np.array_equal(got, host_fed)
    want = oracle.attention(t, q, k, v, scale, mask=_causal(B, M, K2) if causal else None, k2_tile=128)
    _assert_close(t, got, want, label=f"dtm prefill {shape}")
