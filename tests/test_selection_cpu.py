"""Kernel selection — the host logic behind iree_b200_ops_select_matmul / _select_attention and
the `count` regions of the match-and-replace spec — checked without a GPU through an offline
(selection-only) ops object (iree_b200_ops_create_offline). The same expectations run against
a real device in tests/test_matmul_gpu.py / test_attention_gpu.py / test_matmul_bias_gpu.py."""

import re

import pytest

from compiler_plugin import gen_spec
from iree_b200 import hal as H


@pytest.fixture(scope="module")
def ops():
    return H.Ops.offline(sm_count=148)


def _name(ops, sel):
    return ops.executable.export_info(sel["export_ordinal"])["kernel_name"]


def test_offline_object_selects_but_never_computes(ops):
    assert ops.executable.export_count > 50
    info = ops.executable.export_info(ops.executable.lookup("iree_b200_matmul_bf16_f32"))
    assert info["block_dims"] == [256, 1, 1] and info["binding_count"] == 3
    for call in (lambda: ops.matmul_host("bf16", 1, 8, 8, 8, 1, 1, 1),
                 lambda: ops.attention_host("f16", 1, 1, 8, 8, 8, 1.0, 1, 1, 1, 1),
                 lambda: ops.matmul_bias_host("bf16", 1, 8, 8, 8, 1, 1, 1, 1),
                 lambda: ops.matmul_mxfp4_host(8, 8, 32, 1, 1, 1, 1, 1),
                 lambda: ops.mmt4d_host("bf16", 1, 1, 1, 8, 8, 1, 1, 1, 1)):
        with pytest.raises(H.Status) as e:
            call()
        assert e.value.name == "FAILED_PRECONDITION"
    with pytest.raises(H.Status):
        H.Ops.offline(sm_count=1)


def test_matmul_selection_guards(ops):
    s = ops.select_matmul("bf16", 1, 8192, 8192, 8192)
    assert _name(ops, s) == "iree_b200_matmul_bf16_f32"
    assert s["workgroup_count"] == [148, 1, 1] and s["constants"] == [8192, 8192, 8192, 1, 0]
    assert _name(ops, ops.select_matmul("bf16", 1, 128, 512, 512)).endswith("_1cta")
    assert "generic" in _name(ops, ops.select_matmul("bf16", 1, 457, 330, 512))   # N pitch 660 B
    assert "generic" in _name(ops, ops.select_matmul("bf16", 1, 512, 512, 512, alignment=2))
    assert _name(ops, ops.select_matmul("i8", 1, 512, 512, 512, flags=H.MATMUL_TRANSPOSE_RHS)) \
        .startswith("iree_b200_matmul_nt_i8")
    assert _name(ops, ops.select_matmul("f8e5m2", 1, 4096, 4096, 4096)) == "iree_b200_matmul_f8e5m2_f32"
    assert ops.select_matmul("bf16", 1, 0, 16, 16)["export_ordinal"] == -1
    with pytest.raises(H.Status):
        ops.select_matmul("bf16", 1, 457, 330, 512, flags=H.MATMUL_FORCE_2CTA)
    with pytest.raises(H.Status):
        ops.select_matmul(7, 1, 8, 8, 8)
    with pytest.raises(H.Status):
        ops.select_matmul("i8", 1, 256, 256, 256, flags=H.MATMUL_TRUNCATE_RESULT)
    # Generic grid: ceil(N/64) x ceil(M/64) x batch, transpose / truncate in the flag word.
    s = ops.select_matmul("f16", 3, 100, 33, 7, flags=H.MATMUL_TRANSPOSE_RHS)
    assert s["workgroup_count"] == [1, 2, 3] and s["constants"][4] == 2


def test_bias_selection(ops):
    B, R, T = H.MATMUL_BIAS, H.MATMUL_RELU, H.MATMUL_TRUNCATE_RESULT
    assert _name(ops, ops.select_matmul("bf16", 1, 8192, 4096, 4096, flags=B)) == \
        "iree_b200_matmul_bias_bf16_f32"
    assert _name(ops, ops.select_matmul("bf16", 1, 8192, 4096, 4096, flags=B | R | T)) == \
        "iree_b200_matmul_bias_bf16_bf16"
    assert _name(ops, ops.select_matmul("f16", 1, 100, 33, 7, flags=B)) == \
        "iree_b200_matmul_generic_bias_f16_f32"
    assert ops.select_matmul("f16", 1, 8192, 4096, 4096, flags=B | R | 1)["constants"][4] == 3
    with pytest.raises(H.Status):
        ops.select_matmul("bf16", 1, 256, 256, 256, flags=R)
    with pytest.raises(H.Status):
        ops.select_matmul("f8e4m3fn", 1, 256, 256, 256, flags=B)


def test_attention_selection(ops):
    names = lambda sel: [ops.executable.export_info(d["export_ordinal"])["kernel_name"]
                         for d in sel["dispatches"]]
    s = ops.select_attention("f16", 32, 4096, 128, 4096, 128, 0.088, flags=H.ATTENTION_NO_TAIL_SPLIT)
    assert names(s) == ["iree_b200_attention_prefill_f16"]
    assert s["dispatches"][0]["workgroup_count"] == [16, 32, 1]
    # 512 equal units on 148 SMs: 444 run whole, the last 68 are cut in two over K2.
    s = ops.select_attention("f16", 32, 4096, 128, 4096, 128, 0.088)
    assert names(s) == ["iree_b200_attention_prefill_split_f16", "iree_b200_attention_prefill_merge_f16"]
    assert s["dispatches"][0]["workgroup_count"] == [444 + 68 * 2, 1, 1]
    assert s["dispatches"][0]["constants"][4:6] == [444, 2]
    assert s["workspace_bytes"] == 68 * 2 * 256 * 130 * 4
    s = ops.select_attention("f16", 32, 1, 128, 8192, 128, 0.088)
    assert names(s) == ["iree_b200_attention_decode_cluster_f16"]
    assert s["dispatches"][0]["workgroup_count"] == [8, 1, 32] and s["workspace_bytes"] == 0
    s = ops.select_attention("f16", 4, 1, 128, 8192, 128, 0.088)
    assert names(s) == ["iree_b200_attention_decode_split_f16", "iree_b200_attention_decode_merge_f16"]
    assert names(ops.select_attention("bf16", 2, 64, 64, 64, 64, 0.125)) == \
        ["iree_b200_attention_generic_bf16"]
    # Masks: an i1 operand with 16-byte readable rows and the computed causal mask run on the
    # tcgen05 prefill kernel; everything else (f32 additive, ragged K2, few rows) is generic.
    mk = lambda kind, B=2, M=256, K2=256: names(ops.select_attention(
        "f16", B, M, 128, K2, 128, 0.1, mask_kind=kind))
    assert mk(H.ATTENTION_MASK_BOOL) == ["iree_b200_attention_prefill_mask_f16"]
    assert mk(H.ATTENTION_MASK_BOOL, K2=300) == ["iree_b200_attention_generic_f16"]
    assert mk(H.ATTENTION_MASK_F32) == ["iree_b200_attention_generic_f16"]
    assert mk(H.ATTENTION_MASK_CAUSAL) == ["iree_b200_attention_prefill_f16"]
    assert mk(H.ATTENTION_MASK_CAUSAL, B=32, M=1, K2=8192) == ["iree_b200_attention_decode_cluster_f16"]
    assert mk(H.ATTENTION_MASK_CAUSAL, M=8, K2=512) == ["iree_b200_attention_generic_f16"]
    s = ops.select_attention("f16", 32, 4096, 128, 4096, 128, 0.088, mask_kind=H.ATTENTION_MASK_CAUSAL)
    assert names(s) == ["iree_b200_attention_prefill_f16"]      # no tail split: units differ
    assert s["dispatches"][0]["workgroup_count"] == [512, 1, 1]  # 1-D, heaviest first
    assert s["dispatches"][0]["constants"][4] == 1               # kAttnFlagCausal
    with pytest.raises(H.Status):                                # K2 < M
        ops.select_attention("f16", 1, 512, 128, 256, 128, 0.1, mask_kind=H.ATTENTION_MASK_CAUSAL)
    assert ops.select_attention("f16", 0, 4, 8, 8, 8, 1.0)["dispatches"] == []
    with pytest.raises(H.Status):
        ops.select_attention("i8", 1, 4, 8, 8, 8, 1.0)
    with pytest.raises(H.Status):
        ops.select_attention("f16", 1, 64, 64, 64, 64, 1.0, flags=H.ATTENTION_FORCE_PREFILL)


def test_selection_scales_with_the_device(ops):
    """The persistent grids and the decode form follow the SM count, not a constant."""
    small = H.Ops.offline(sm_count=64)
    assert small.select_matmul("bf16", 1, 8192, 8192, 8192)["workgroup_count"] == [64, 1, 1]
    # 32 rows x 8 CTAs fill 148 SMs (cluster form) and 64 SMs alike; 4 rows fill neither.
    assert len(small.select_attention("f16", 32, 1, 128, 8192, 128, 0.1)["dispatches"]) == 1
    assert len(small.select_attention("f16", 4, 1, 128, 8192, 128, 0.1)["dispatches"]) == 2
    ops.set_reserved_sms(16)
    try:
        assert ops.select_matmul("bf16", 1, 8192, 8192, 8192)["workgroup_count"] == [132, 1, 1]
    finally:
        ops.set_reserved_sms(0)
    with pytest.raises(H.Status):
        ops.set_reserved_sms(148)


@pytest.mark.parametrize("family,flag,g", [("", H.MATMUL_FORCE_2CTA, 2), ("_1cta", H.MATMUL_FORCE_1CTA, 1)])
def test_spec_count_regions_restate_the_host_selection(ops, family, flag, g):
    """The `count` region the spec gives each tcgen05 matmul entry (evaluated by the IREE
    runtime on the host at dispatch time) is the arithmetic of iree_b200_ops_select_matmul:
    re-evaluate the region's constants over a shape sweep and compare the grids."""
    import os
    text = gen_spec.generate(os.path.join(os.path.dirname(os.path.dirname(
        os.path.abspath(__file__))), "iree_b200", "_build"))
    block = text[text.index(f'hal.dispatch.extern "iree_b200_matmul_bf16_f32{family}"'):]
    block = block[:block.index("hal.return")]
    const = lambda n: int(re.search(r"%%%s = arith.constant (\d+) : index" % n, block).group(1))
    tile_m, tile_n, group, sms = const("tile_m"), const("tile_n"), const("g"), const("sms")
    assert (tile_m, tile_n, group) == (128 * g, 256, g)
    for batch in (1, 3):
        for M in (128, 129, 256, 1000, 4096, 8192):
            for N in (256, 264, 2048, 8192):
                want = min(-(-M // tile_m) * -(-N // tile_n) * batch * group, sms)
                got = ops.select_matmul("bf16", batch, M, N, 512, flags=flag)
                assert got["workgroup_count"] == [want, 1, 1], (batch, M, N)


def test_device_map_exports_are_selected_only_where_they_exist(ops):
    """FLAG_DEVICE_MAPS: the recipe-free twins of the tcgen05 kernels (what a stock IREE CUDA
    driver can launch) — same grids and constants as the host-fed exports."""
    for t, acc in (("bf16", "f32"), ("f16", "f32"), ("i8", "i32")):
        for flags, nt in ((0, ""), (H.MATMUL_TRANSPOSE_RHS, "nt_")):
            big = ops.select_matmul(t, 1, 8192, 8192, 8192, flags=flags | H.MATMUL_DEVICE_MAPS)
            ref = ops.select_matmul(t, 1, 8192, 8192, 8192, flags=flags)
            assert _name(ops, big) == f"iree_b200_matmul_{nt}{t}_{acc}_dtm"
            assert big["workgroup_count"] == ref["workgroup_count"] == [148, 1, 1]
            assert big["constants"] == ref["constants"]
            small = ops.select_matmul(t, 1, 256, 512, 128, flags=flags | H.MATMUL_DEVICE_MAPS)
            assert _name(ops, small) == f"iree_b200_matmul_{nt}{t}_{acc}_1cta_dtm"
            info = ops.executable.export_info(big["export_ordinal"])
            assert info["tma_map_count"] == 0 and info["constant_count"] == 5
    # ... every tcgen05 export has its twin: f8, truncating, bias / activation forms.
    for t, flags, want in (("f8e4m3fn", 0, "iree_b200_matmul_f8e4m3fn_f32_dtm"),
                           ("bf16", H.MATMUL_TRUNCATE_RESULT, "iree_b200_matmul_bf16_bf16_dtm"),
                           ("f16", H.MATMUL_BIAS | H.MATMUL_RELU, "iree_b200_matmul_bias_f16_f32_dtm"),
                           ("bf16", H.MATMUL_BIAS | H.MATMUL_TRUNCATE_RESULT | H.MATMUL_TRANSPOSE_RHS,
                            "iree_b200_matmul_bias_nt_bf16_bf16_dtm")):
        sel = ops.select_matmul(t, 1, 4096, 4096, 4096, flags=flags | H.MATMUL_DEVICE_MAPS)
        ref = ops.select_matmul(t, 1, 4096, 4096, 4096, flags=flags)
        assert _name(ops, sel) == want == _name(ops, ref) + "_dtm"
        assert (sel["workgroup_count"], sel["constants"]) == (ref["workgroup_count"], ref["constants"])
        assert ops.executable.export_info(sel["export_ordinal"])["tma_map_count"] == 0
    for bad in (dict(t="bf16", flags=H.MATMUL_DEVICE_MAPS | H.MATMUL_FORCE_GENERIC),
                dict(t="bf16", flags=H.MATMUL_DEVICE_MAPS, N=1001)):      # fails the TMA guards
        with pytest.raises(H.Status) as e:
            ops.select_matmul(bad["t"], 1, 1024, bad.get("N", 1024), 1024, flags=bad["flags"])
        assert e.value.name == "INVALID_ARGUMENT"
    for t in ("f16", "bf16"):
        for kind, grid in ((H.ATTENTION_MASK_NONE, [16, 32, 1]), (H.ATTENTION_MASK_CAUSAL, [512, 1, 1])):
            sel = ops.select_attention(t, 32, 4096, 128, 4096, 128, 0.088, mask_kind=kind,
                                       flags=H.ATTENTION_DEVICE_MAPS)
            (d,) = sel["dispatches"]
            assert ops.executable.export_info(d["export_ordinal"])["kernel_name"] == \
                f"iree_b200_attention_prefill_{t}_dtm"
            assert d["workgroup_count"] == grid and sel["workspace_bytes"] == 0
            assert d["constants"][4] == (1 if kind == H.ATTENTION_MASK_CAUSAL else 0)
    with pytest.raises(H.Status):
        ops.select_attention("f16", 2, 512, 128, 512, 128, 0.088, mask_kind=H.ATTENTION_MASK_BOOL,
                             flags=H.ATTENTION_DEVICE_MAPS)
