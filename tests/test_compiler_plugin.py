"""The generated match-and-replace spec must agree with the kernel manifest (what the
CUDA target serialises from the export op, CUDATarget.cpp:661-691) and with the host-side
selection logic it mirrors. iree-opt is not available here, so this is a structural check."""

import json
import os
import re

from compiler_plugin import gen_spec
from iree_b200 import build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _spec():
    build.build()
    return gen_spec.generate(os.path.join(ROOT, "iree_b200", "_build"))


def test_spec_is_balanced_and_covers_every_tcgen05_family():
    text = _spec()
    for open_c, close_c in ("{}", "()", "[]"):
        assert text.count(open_c) == text.count(close_c), open_c
    with open(os.path.join(ROOT, "iree_b200", "_build", "exports.json")) as f:
        exports = {e["name"]: e for e in json.load(f)["exports"]}
    dispatched = re.findall(r'hal\.dispatch\.extern "(\w+)"', text)
    # each plain matmul kernel: 2-D + batched entry (10); each bias kernel: with and without relu (10)
    assert len(dispatched) == len(set(dispatched)) + 20
    for name in set(dispatched):
        e = exports[name]
        block = text[text.index(f'hal.dispatch.extern "{name}"'):]
        block = block[:block.index("return %result")]
        assert f"constants = {e['constants']}" in block
        assert block.count("#hal.pipeline.binding<") == len(e["binding_flags"])
        assert block.count("ReadOnly") == sum(f & 1 for f in e["binding_flags"])
        assert f"workgroup_size = [{e['block_dims'][0]} : index" in block
        assert f"workgroup_local_memory = {e['shared_memory']} : index" in block
        assert f'path = "{e["module"]}.cubin"' in block
    for et, acc in (("bf16", "f32"), ("f16", "f32"), ("i8", "i32"), ("f8e4m3fn", "f32"),
                    ("f8e5m2", "f32")):
        assert f"iree_b200_matmul_{et}_{acc}" in dispatched
        assert f"iree_b200_matmul_{et}_{acc}_1cta" in dispatched
    assert "iree_b200_attention_prefill_f16" in dispatched
    for name in ("iree_b200_matmul_bias_bf16_f32", "iree_b200_matmul_bias_bf16_bf16_1cta",
                 "iree_b200_matmul_bias_i8_i32", "iree_b200_matmul_bias_f16_f16"):
        assert name in dispatched
    # relu travels as flags bit 1; the DAG matchers precede the plain contraction matchers
    relu_entry = text[text.index("@iree_b200_matmul_bias_bf16_bf16_relu_entry_point"):]
    assert "%flags = arith.constant 2 : i32" in relu_entry[:relu_entry.index("hal.dispatch.extern")]
    main = text[text.index("@__transform_main"):]
    assert main.index("@match_iree_b200_matmul_bias_bf16_f32 ") < main.index("@match_iree_b200_matmul_bf16_f32 ")
    # every matcher is wired to a rewriter in @__transform_main
    matchers = re.findall(r"named_sequence @(match_\w+)", text)
    for m in matchers:
        assert f"@{m} -> @rewrite_{m[len('match_'):]}" in text
    # SSA names are defined once per function body
    for body in text.split("func.func private")[1:]:
        body = body[:body.index("return %result")]
        top = [l for l in body.splitlines() if re.match(r"    %\w+ = ", l)]
        names = [l.split("=")[0].strip() for l in top]
        assert len(names) == len(set(names)), names


def test_count_regions_match_host_selection():
    """The count regions restate iree_b200_ops_select_*: persistent grid for the GEMMs,
    ceil(M/256) x B for prefill attention."""
    text = _spec()
    assert "arith.constant 148 : index" in text and "arith.constant 256 : index" in text
    two_cta = text[text.index('hal.dispatch.extern "iree_b200_matmul_bf16_f32"'):]
    two_cta = two_cta[:two_cta.index("hal.return")]
    assert "%tile_m = arith.constant 256" in two_cta and "%g = arith.constant 2" in two_cta
    one_cta = text[text.index('hal.dispatch.extern "iree_b200_matmul_bf16_f32_1cta"'):]
    one_cta = one_cta[:one_cta.index("hal.return")]
    assert "%tile_m = arith.constant 128" in one_cta and "%g = arith.constant 1" in one_cta


def test_stock_driver_spec_uses_only_recipe_free_exports():
    """--stock-driver: every dispatched kernel is a `*_dtm` export (no `__tma` recipe, the
    plain custom-dispatch ABI), so the compiled module needs no driver hook."""
    build.build()
    text = gen_spec.generate(os.path.join(ROOT, "iree_b200", "_build"), stock_driver=True)
    for open_c, close_c in ("{}", "()", "[]"):
        assert text.count(open_c) == text.count(close_c), open_c
    with open(os.path.join(ROOT, "iree_b200", "_build", "exports.json")) as f:
        exports = {e["name"]: e for e in json.load(f)["exports"]}
    dispatched = set(re.findall(r'hal\.dispatch\.extern "(\w+)"', text))
    # the host-fed spec's kernels, one for one: 5 types x {2-CTA, 1-CTA} matmul, 5 bias forms x
    # {2-CTA, 1-CTA}, 2 attention
    host_fed = set(re.findall(r'hal\.dispatch\.extern "(\w+)"', _spec()))
    assert dispatched == {n + "_dtm" for n in host_fed} and len(dispatched) == 22
    for name in dispatched:
        e = exports[name]
        assert e["constants"] == 5
        assert e["module"] in ("gemm_dtm_kernels", "gemm_bias_dtm_kernels", "attention_kernels")
        twin = exports[name[:-len("_dtm")]]
        for key in ("block_dims", "shared_memory", "binding_flags"):
            assert e[key] == twin[key], (name, key)
    from iree_b200 import hal as H
    ops = H.Ops.offline(148)
    for name in dispatched:
        info = ops.executable.export_info(ops.executable.lookup(name))
        assert info["tma_map_count"] == 0, name
    matchers = re.findall(r"named_sequence @(match_\w+)", text)
    for m in matchers:
        assert f"@{m} -> @rewrite_{m[len('match_'):]}" in text


def test_known_target_patch_applies_to_the_reference(tmp_path):
    """integration/known_targets_sm100.patch (the sm_100a / b200 entry of SURVEY §8(f) rank 4)
    applies cleanly to the reference's KnownTargets.cpp, and the new WgpDetails initialiser has
    the shape of its sm_120 neighbour (it cannot be compiled here: no LLVM tree)."""
    import shutil
    import subprocess
    rel = "compiler/src/iree/compiler/Codegen/Dialect/GPU/TargetUtils/KnownTargets.cpp"
    src = os.path.join("/root/reference", rel)
    if not os.path.exists(src):
        import pytest
        pytest.skip("no reference checkout")
    dst = tmp_path / rel
    dst.parent.mkdir(parents=True)
    shutil.copy(src, dst)
    patch = os.path.join(ROOT, "integration", "known_targets_sm100.patch")
    subprocess.run(["patch", "-s", "-p1", "-d", str(tmp_path)], stdin=open(patch), check=True)
    text = dst.read_text()
    body = lambda name: text[text.index(f"*get{name}WgpDetails() {{"):].split("\n}\n")[0]
    new, old = body("SM100"), body("SM120")
    assert new.count(",") == old.count(",") and new.count("MMAIntrinsic::") == old.count("MMAIntrinsic::")
    assert "227 * 1024" in new and "b200Chip = {148}" in text
    assert '.Cases({"b200", "sm_100", "sm_100a"}, TargetDetails{sm100Wgp, &b200Chip})' in text
    assert '.Case("b200", "sm_100a")' in text
    for open_c, close_c in ("{}", "()"):
        assert text.count(open_c) == text.count(close_c)
