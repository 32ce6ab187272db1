"""Generates the transform-dialect match-and-replace spec for the B200 dispatch kernels.

  python -m compiler_plugin.gen_spec [--objects-dir DIR] [-o spec.mlir]

Shape of the output: samples/custom_dispatch/vulkan/shaders/example_transform_spec.mlir:20-167
(the reference has no CUDA instance of it). One private entry-point func per kernel wraps a
`hal.dispatch.extern` (HALOps.td:424-560) whose
  * operands are the push constants (i32) followed by the tensors, i.e. the kernel ABI of
    samples/custom_dispatch/cuda/kernels/kernels.cu:7-34;
  * `count` region is the arithmetic of iree_b200_ops_select_matmul / _select_attention
    (iree_b200/csrc/ops/ops.cc);
  * layout / workgroup_size / workgroup_local_memory come from the kernel manifest
    (iree_b200/_build/exports.json, the data CUDATarget.cpp:661-691 serialises);
  * object is the cubin of the kernel's translation unit (one object per external variant,
    CUDATarget.cpp:509-515), resolved via --iree-hal-executable-object-search-path.
Matchers use transform.iree.match.contraction / match.attention
(PreprocessingExtensionsOps.td:175-246,331-410) plus the TMA guards (16-byte row pitches).

The spec cannot be run through iree-opt in this container (no LLVM tree); tests check it
against the manifest (tests/test_compiler_plugin.py).
"""

from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SM_COUNT = 148  # B200; the count region runs on the host at dispatch time

MLIR_TYPE = {"bf16": "bf16", "f16": "f16", "i8": "i8", "f8e4m3fn": "f8E4M3FN", "f8e5m2": "f8E5M2"}
ACC_TYPE = {"bf16": "f32", "f16": "f32", "i8": "i32", "f8e4m3fn": "f32", "f8e5m2": "f32"}
# Elements per 16 bytes: the TMA row-pitch guard.
PITCH = {"bf16": 8, "f16": 8, "i8": 16, "f8e4m3fn": 16, "f8e5m2": 16}


def load_manifest(objects_dir: str) -> dict:
    with open(os.path.join(objects_dir, "exports.json")) as f:
        exports = json.load(f)["exports"]
    return {e["name"]: e for e in exports}


def _layout(e: dict) -> str:
    b = ",\n        ".join("#hal.pipeline.binding<storage_buffer%s>" % (", ReadOnly" if f & 1 else "")
                           for f in e["binding_flags"])
    return ("#hal.pipeline.layout<constants = %d, bindings = [\n        %s\n      ]>"
            % (e["constants"], b))


def _attrs(e: dict) -> str:
    wg = ", ".join("%d : index" % d for d in e["block_dims"])
    s = "workgroup_size = [%s]" % wg
    if e["shared_memory"]:
        s += ", workgroup_local_memory = %d : index" % e["shared_memory"]
    return "attributes {%s}" % s


def _objects(e: dict) -> str:
    return ('objects({\n        #cuda_target ordinal(0) = [\n          #hal.executable.object<{path = '
            '"%s.cubin"}>\n        ]\n      })' % e["module"])


def matmul_entry(e: dict, et: str, batched: bool, cta_group: int) -> str:
    """Entry point: C = acc + A.B with the accumulator tied to the result (flags bit 0)."""
    name = e["name"]
    t, a = MLIR_TYPE[et], ACC_TYPE[et]
    r = "?x?x?" if batched else "?x?"
    lt, ct = f"tensor<{r}x{t}>", f"tensor<{r}x{a}>"
    o = 1 if batched else 0
    tile_m = 128 * cta_group
    fn = f"{name}_{'batch_' if batched else ''}entry_point"
    batch_dim = (f"    %b = tensor.dim %lhs, %c0 : {lt}\n" if batched else "    %b = arith.constant 1 : index\n")
    bd = "%b, " if batched else ""
    return f"""  func.func private @{fn}(%lhs: {lt}, %rhs: {lt}, %acc: {ct}) -> {ct} {{
    %c0 = arith.constant 0 : index
    %c1 = arith.constant 1 : index
    %c2 = arith.constant 2 : index
{batch_dim}    %m = tensor.dim %lhs, %c{o} : {lt}
    %k = tensor.dim %lhs, %c{o + 1} : {lt}
    %n = tensor.dim %rhs, %c{o + 1} : {lt}
    // i32 push constants (kernels.cu:19-21): M, N, K, batch, flags (bit 0: accumulate).
    %m_i32 = arith.index_cast %m : index to i32
    %n_i32 = arith.index_cast %n : index to i32
    %k_i32 = arith.index_cast %k : index to i32
    %b_i32 = arith.index_cast %b : index to i32
    %flags = arith.constant 1 : i32
    %result = hal.dispatch.extern "{name}"[%m, %n, %b](%m_i32, %n_i32, %k_i32, %b_i32, %flags, %lhs, %rhs, %acc)
        : (i32, i32, i32, i32, i32, {lt}{{{bd}%m, %k}}, {lt}{{{bd}%k, %n}}, {ct}{{{bd}%m, %n}}) -> %acc{{{bd}%m, %n}}
      count(%device: !hal.device, %wm: index, %wn: index, %wb: index) -> (index, index, index) {{
        // Persistent grid: min(#tiles * cta_group, #SMs rounded to the cluster size).
        %tile_m = arith.constant {tile_m} : index
        %tile_n = arith.constant 256 : index
        %tm = arith.ceildivui %wm, %tile_m : index
        %tn = arith.ceildivui %wn, %tile_n : index
        %t0 = arith.muli %tm, %tn : index
        %t1 = arith.muli %t0, %wb : index
        %g = arith.constant {cta_group} : index
        %ctas = arith.muli %t1, %g : index
        %sms = arith.constant {SM_COUNT - SM_COUNT % cta_group} : index
        %x = arith.minui %ctas, %sms : index
        %one = arith.constant 1 : index
        hal.return %x, %one, %one : index, index, index
      }}
      layout({_layout(e)})
      {_objects(e)}
      {_attrs(e)}
    return %result : {ct}
  }}
"""


def matmul_matcher(name: str, et: str, batched: bool, big: bool) -> str:
    t, a = MLIR_TYPE[et], ACC_TYPE[et]
    o = 1 if batched else 0
    p = PITCH[et]
    size_guard = ("transform.match.param.cmpi ge %m_dims, %c256 : !transform.param<i64>" if big else
                  "transform.match.param.cmpi lt %m_dims, %c256 : !transform.param<i64>")
    rank = "?x?x?" if batched else "?x?"
    rank_guard = ("transform.iree.match.cast_compatible_type %lhs = tensor<" + rank + "x" + t +
                  "> : !transform.any_value")
    return f"""  transform.named_sequence @match_{name}(%op: !transform.any_op {{transform.readonly}}) -> !transform.any_op {{
    %batch_dims, %m_dims, %n_dims, %k_dims = transform.iree.match.contraction %op,
      lhs_type = {t}, rhs_type = {t}, output_type = {a} : !transform.any_op -> !transform.param<i64>
    %c256 = transform.param.constant 256 : i64 -> !transform.param<i64>
    {size_guard}
    %lhs = transform.get_operand %op[0] : (!transform.any_op) -> !transform.any_value
    %rhs = transform.get_operand %op[1] : (!transform.any_op) -> !transform.any_value
    {rank_guard}
    // TMA guards: 16-byte row pitches of A ([.., M, K]), B ([.., K, N]) and C ([.., M, N]).
    transform.iree.match.dim_is_multiple_of %lhs[{o + 1}], {p} : !transform.any_value
    transform.iree.match.dim_is_multiple_of %rhs[{o + 1}], {p} : !transform.any_value
    transform.yield %op : !transform.any_op
  }}
"""


def matmul_rewriter(name: str, entry: str) -> str:
    return f"""  transform.named_sequence @rewrite_{name}(%op: !transform.any_op {{transform.readonly}}) {{
    %module = transform.util.get_nearest_symbol_table %op : (!transform.any_op) -> !transform.any_op
    %func = transform.util.import_symbol @{entry} into %module : (!transform.any_op) -> !transform.any_op
    %lhs = transform.get_operand %op[0] : (!transform.any_op) -> !transform.any_value
    %rhs = transform.get_operand %op[1] : (!transform.any_op) -> !transform.any_value
    %acc = transform.get_operand %op[2] : (!transform.any_op) -> !transform.any_value
    %out = transform.get_result %op[0] : (!transform.any_op) -> !transform.any_value
    transform.func.cast_and_call %func(%lhs, %rhs, %acc) -> %out before %op {{
        transform.type_conversion.tensor.cast_shape_dynamic_dims
      }} : (!transform.any_op, !transform.any_value, !transform.any_value, !transform.any_value, !transform.any_value, !transform.any_op) -> !transform.any_op
    transform.yield
  }}
"""


def _linear_types(et: str, truncate: bool):
    t, a = MLIR_TYPE[et], ACC_TYPE[et]
    return t, a, (t if truncate else a)


def linear_entry(e: dict, key: str, et: str, relu: bool, truncate: bool, cta_group: int) -> str:
    """Entry point of the fused dispatch out = act(A.B + bias[n]) [truncated]: bindings lhs, rhs,
    bias, out (gemm_bias_kernels.cu); flags bit 1 = relu. %init1 (the matched linalg.fill's
    destination) only carries the shape."""
    name = e["name"]
    t, a, o = _linear_types(et, truncate)
    lt, bt, at, ot = f"tensor<?x?x{t}>", f"tensor<?x{a}>", f"tensor<?x?x{a}>", f"tensor<?x?x{o}>"
    tile_m = 128 * cta_group
    return f"""  func.func private @{key}_entry_point(%lhs: {lt}, %rhs: {lt}, %bias: {bt}, %init1: {at}, %out: {ot}) -> {ot} {{
    %c0 = arith.constant 0 : index
    %c1 = arith.constant 1 : index
    %b = arith.constant 1 : index
    %m = tensor.dim %lhs, %c0 : {lt}
    %k = tensor.dim %lhs, %c1 : {lt}
    %n = tensor.dim %rhs, %c1 : {lt}
    // i32 push constants: M, N, K, batch, flags (bit 0: accumulate, bit 1: relu after the bias).
    %m_i32 = arith.index_cast %m : index to i32
    %n_i32 = arith.index_cast %n : index to i32
    %k_i32 = arith.index_cast %k : index to i32
    %b_i32 = arith.index_cast %b : index to i32
    %flags = arith.constant {2 if relu else 0} : i32
    %result = hal.dispatch.extern "{name}"[%m, %n, %b](%m_i32, %n_i32, %k_i32, %b_i32, %flags, %lhs, %rhs, %bias, %out)
        : (i32, i32, i32, i32, i32, {lt}{{%m, %k}}, {lt}{{%k, %n}}, {bt}{{%n}}, {ot}{{%m, %n}}) -> %out{{%m, %n}}
      count(%device: !hal.device, %wm: index, %wn: index, %wb: index) -> (index, index, index) {{
        // Persistent grid: min(#tiles * cta_group, #SMs rounded to the cluster size).
        %tile_m = arith.constant {tile_m} : index
        %tile_n = arith.constant 256 : index
        %tm = arith.ceildivui %wm, %tile_m : index
        %tn = arith.ceildivui %wn, %tile_n : index
        %t0 = arith.muli %tm, %tn : index
        %t1 = arith.muli %t0, %wb : index
        %g = arith.constant {cta_group} : index
        %ctas = arith.muli %t1, %g : index
        %sms = arith.constant {SM_COUNT - SM_COUNT % cta_group} : index
        %x = arith.minui %ctas, %sms : index
        %one = arith.constant 1 : index
        hal.return %x, %one, %one : index, index, index
      }}
      layout({_layout(e)})
      {_objects(e)}
      {_attrs(e)}
    return %result : {ot}
  }}
"""


def linear_matcher(key: str, et: str, relu: bool, truncate: bool, big: bool) -> str:
    """DAG matcher (transform.iree.match.cast_compatible_dag_from_root,
    PreprocessingExtensionsOps.td:18-59; used like samples/custom_dispatch/cpu/mlp_plugin/
    mlp_spec.mlir:51-72): linalg.fill 0 -> linalg.matmul -> one elementwise linalg.generic
    holding add(bias broadcast along rows) [, max 0] [, truncf] - the producer-consumer group
    dispatch creation puts behind a contraction."""
    t, a, o = _linear_types(et, truncate)
    is_int = a == "i32"
    zero = "0 : i32" if is_int else "0.0 : f32"
    add, mx = ("arith.addi", "arith.maxsi") if is_int else ("arith.addf", "arith.maximumf")
    lt, bt, at, ot = f"tensor<?x?x{t}>", f"tensor<?x{a}>", f"tensor<?x?x{a}>", f"tensor<?x?x{o}>"
    body = [f"            %s = {add} %x, %bv : {a}"]
    last = "%s"
    if relu:
        body.append(f"            %r = {mx} %s, %zero : {a}")
        last = "%r"
    if truncate:
        body.append(f"            %tr = arith.truncf {last} : {a} to {o}")
        last = "%tr"
    body.append(f"            linalg.yield {last} : {o}")
    p = PITCH[et]
    size_guard = ("transform.iree.match.dim_bounds %lhs[0], umin = 256, none : !transform.any_value"
                  if big else
                  "transform.iree.match.dim_bounds %lhs[0], none, umax = 255 : !transform.any_value")
    return f"""  transform.named_sequence @match_{key}(%root: !transform.any_op {{transform.readonly}}) -> (!transform.any_value, !transform.any_value) {{
    %ins, %outs = transform.iree.match.cast_compatible_dag_from_root %root {{
      ^bb0(%a: {lt}, %b: {lt}, %bias: {bt}, %init1: {at}, %init2: {ot}):
        %zero = arith.constant {zero}
        %fill = linalg.fill ins(%zero : {a}) outs(%init1 : {at}) -> {at}
        %mm = linalg.matmul ins(%a, %b : {lt}, {lt}) outs(%fill : {at}) -> {at}
        %res = linalg.generic {{
            indexing_maps = [affine_map<(d0, d1) -> (d0, d1)>, affine_map<(d0, d1) -> (d1)>,
                             affine_map<(d0, d1) -> (d0, d1)>],
            iterator_types = ["parallel", "parallel"]}}
            ins(%mm, %bias : {at}, {bt}) outs(%init2 : {ot}) {{
          ^bb1(%x: {a}, %bv: {a}, %o: {o}):
{chr(10).join(body)}
          }} -> {ot}
      }} : (!transform.any_op) -> (!transform.any_value, !transform.any_value)
    // Guards on the matmul's operands: kernel family by M, 16-byte row pitches for TMA.
    %matmul = transform.get_producer_of_operand %root[0] : (!transform.any_op) -> !transform.any_op
    %lhs = transform.get_operand %matmul[0] : (!transform.any_op) -> !transform.any_value
    %rhs = transform.get_operand %matmul[1] : (!transform.any_op) -> !transform.any_value
    {size_guard}
    transform.iree.match.dim_is_multiple_of %lhs[1], {p} : !transform.any_value
    transform.iree.match.dim_is_multiple_of %rhs[1], {p} : !transform.any_value
    transform.yield %ins, %outs : !transform.any_value, !transform.any_value
  }}
"""


def linear_rewriter(key: str) -> str:
    return f"""  transform.named_sequence @rewrite_{key}(%ins: !transform.any_value {{transform.readonly}},
                                      %out: !transform.any_value {{transform.readonly}}) {{
    %root = transform.get_defining_op %out : (!transform.any_value) -> !transform.any_op
    %module = transform.util.get_nearest_symbol_table %root : (!transform.any_op) -> !transform.any_op
    %func = transform.util.import_symbol @{key}_entry_point into %module if undefined : (!transform.any_op) -> !transform.any_op
    transform.func.cast_and_call %func(%ins) -> %out after %root {{
        transform.type_conversion.tensor.cast_shape_dynamic_dims
      }} : (!transform.any_op, !transform.any_value, !transform.any_value, !transform.any_op) -> !transform.any_op
    transform.yield
  }}
"""


ATTN_MAPS = """#map_query = affine_map<(b, m, n, k1, k2) -> (b, m, k1)>
#map_key = affine_map<(b, m, n, k1, k2) -> (b, k2, k1)>
#map_value = affine_map<(b, m, n, k1, k2) -> (b, k2, n)>
#map_scale = affine_map<(b, m, n, k1, k2) -> ()>
#map_output = affine_map<(b, m, n, k1, k2) -> (b, m, n)>
"""


def attention_entry(e: dict, et: str) -> str:
    """Prefill kernel entry: Q [B,M,128], K/V [B,K2,128], scale scalar -> O [B,M,128]."""
    name = e["name"]
    t = MLIR_TYPE[et]
    qt, kt = f"tensor<?x?x128x{t}>", f"tensor<?x?x128x{t}>"
    return f"""  func.func private @{name}_entry_point(%q: {qt}, %k: {kt}, %v: {kt}, %scale: {t}) -> {qt} {{
    %c0 = arith.constant 0 : index
    %c1 = arith.constant 1 : index
    %b = tensor.dim %q, %c0 : {qt}
    %m = tensor.dim %q, %c1 : {qt}
    %k2 = tensor.dim %k, %c1 : {kt}
    %b_i32 = arith.index_cast %b : index to i32
    %m_i32 = arith.index_cast %m : index to i32
    %k2_i32 = arith.index_cast %k2 : index to i32
    // The scale travels as the f32 bit pattern of the (already {t}-rounded) scalar.
    %scale_f32 = arith.extf %scale : {t} to f32
    %scale_i32 = arith.bitcast %scale_f32 : f32 to i32
    %flags = arith.constant 0 : i32
    %result = hal.dispatch.extern "{name}"[%m, %b](%b_i32, %m_i32, %k2_i32, %scale_i32, %flags, %q, %k, %v)
        : (i32, i32, i32, i32, i32, {qt}{{%b, %m}}, {kt}{{%b, %k2}}, {kt}{{%b, %k2}}) -> {qt}{{%b, %m}}
      count(%device: !hal.device, %wm: index, %wb: index) -> (index, index, index) {{
        // One CTA per two 128-row Q tiles of one batch slice.
        %rows = arith.constant 256 : index
        %x = arith.ceildivui %wm, %rows : index
        %one = arith.constant 1 : index
        hal.return %x, %wb, %one : index, index, index
      }}
      layout({_layout(e)})
      {_objects(e)}
      {_attrs(e)}
    return %result : {qt}
  }}
"""


def attention_matcher(name: str, et: str) -> str:
    t = MLIR_TYPE[et]
    return f"""  transform.named_sequence @match_{name}(%op: !transform.any_op {{transform.readonly}}) -> !transform.any_op {{
    // No mask operand, identity region (the matcher rejects other forms), K1 == N == 128.
    %batch, %m, %n, %k1, %k2 =
      transform.iree.match.attention %op,
        query_type = {t}, key_type = {t}, value_type = {t}, output_type = {t},
        indexing_maps = [#map_query, #map_key, #map_value, #map_scale, #map_output] :
        !transform.any_op -> !transform.param<i64>
    transform.iree.match.dims_equal %k1, [128] : !transform.param<i64>
    transform.iree.match.dims_equal %n, [128] : !transform.param<i64>
    %c16 = transform.param.constant 16 : i64 -> !transform.param<i64>
    transform.match.param.cmpi gt %m, %c16 : !transform.param<i64>
    transform.yield %op : !transform.any_op
  }}
"""


def attention_rewriter(name: str) -> str:
    return f"""  transform.named_sequence @rewrite_{name}(%op: !transform.any_op {{transform.readonly}}) {{
    %module = transform.util.get_nearest_symbol_table %op : (!transform.any_op) -> !transform.any_op
    %func = transform.util.import_symbol @{name}_entry_point into %module : (!transform.any_op) -> !transform.any_op
    %q = transform.get_operand %op[0] : (!transform.any_op) -> !transform.any_value
    %k = transform.get_operand %op[1] : (!transform.any_op) -> !transform.any_value
    %v = transform.get_operand %op[2] : (!transform.any_op) -> !transform.any_value
    %scale = transform.get_operand %op[3] : (!transform.any_op) -> !transform.any_value
    %out = transform.get_result %op[0] : (!transform.any_op) -> !transform.any_value
    transform.func.cast_and_call %func(%q, %k, %v, %scale) -> %out before %op {{
        transform.type_conversion.tensor.cast_shape_dynamic_dims
      }} : (!transform.any_op, !transform.any_value, !transform.any_value, !transform.any_value, !transform.any_value, !transform.any_value, !transform.any_op) -> !transform.any_op
    transform.yield
  }}
"""


def generate(objects_dir: str, stock_driver: bool = False) -> str:
    """|stock_driver|: use the `*_dtm` twins of every export (tensor maps built on the device),
    so the compiled module runs on an UNMODIFIED IREE runtime — no `__tma` recipe, no driver
    hook. Same matchers, same count regions, same layouts."""
    man = load_manifest(objects_dir)
    entries, matchers, rewriters, pairs = [], [], [], []
    dtm = "_dtm" if stock_driver else ""
    for et in ("bf16", "f16", "i8", "f8e4m3fn", "f8e5m2"):
        acc = ACC_TYPE[et]
        for batched in (False, True):
            for big, suffix, g in ((True, "", 2), (False, "_1cta", 1)):
                kernel = f"iree_b200_matmul_{et}_{acc}{suffix}{dtm}"
                key = f"{kernel}{'_batch' if batched else ''}"
                entries.append(matmul_entry(man[kernel], et, batched, g))
                matchers.append(matmul_matcher(key, et, batched, big))
                rewriters.append(matmul_rewriter(
                    key, f"{kernel}_{'batch_' if batched else ''}entry_point"))
                pairs.append(f"@match_{key} -> @rewrite_{key}")
    # Fused consumers first: foreach_match takes the first matcher that succeeds, and the
    # DAG roots (the elementwise generic) are visited independently of the matmul inside.
    linear = []
    for et in ("bf16", "f16", "i8"):
        for truncate in ((False, True) if et != "i8" else (False,)):
            out_t = MLIR_TYPE[et] if truncate else ACC_TYPE[et]
            for relu in (False, True):
                for big, suffix, g in ((True, "", 2), (False, "_1cta", 1)):
                    kernel = f"iree_b200_matmul_bias_{et}_{out_t}{suffix}{dtm}"
                    key = f"{kernel}{'_relu' if relu else ''}"
                    entries.append(linear_entry(man[kernel], key, et, relu, truncate, g))
                    matchers.append(linear_matcher(key, et, relu, truncate, big))
                    rewriters.append(linear_rewriter(key))
                    linear.append(f"@match_{key} -> @rewrite_{key}")
    pairs = linear + pairs
    for et in ("f16", "bf16"):
        kernel = f"iree_b200_attention_prefill_{et}{dtm}"
        entries.append(attention_entry(man[kernel], et))
        matchers.append(attention_matcher(kernel, et))
        rewriters.append(attention_rewriter(kernel))
        pairs.append(f"@match_{kernel} -> @rewrite_{kernel}")
    head = f"""// GENERATED by compiler_plugin/gen_spec.py from {os.path.basename(objects_dir)}/exports.json - do not edit.
// iree-compile --iree-hal-target-backends=cuda --iree-cuda-target=sm_100a \\
//   --iree-preprocessing-transform-spec-filename=<this file> \\
//   --iree-hal-executable-object-search-path=<dir holding the .cubin objects>

#cuda_target = #hal.executable.target<"cuda", "cuda-nvptx-fb", {{target_arch = "sm_100a"}}>
{ATTN_MAPS}
module attributes {{transform.with_named_sequence}} {{
"""
    main = """  transform.named_sequence @__transform_main(%module: !transform.any_op) {
    %funcs = transform.structured.match ops{["func.func"]} in %module : (!transform.any_op) -> !transform.any_op
    transform.foreach %funcs : !transform.any_op {
      ^bb1(%func: !transform.any_op):
        transform.foreach_match in %func
            PAIRS
          : (!transform.any_op) -> (!transform.any_op)
    }
    transform.apply_dce to %module : !transform.any_op
    transform.yield
  }
}
""".replace("PAIRS", ",\n            ".join(pairs))
    return head + "\n".join(entries) + "\n" + "\n".join(matchers) + "\n" + "\n".join(rewriters) + "\n" + main


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--objects-dir", default=os.path.join(ROOT, "iree_b200", "_build"))
    ap.add_argument("-o", "--output", default="-")
    ap.add_argument("--stock-driver", action="store_true",
                    help="use the *_dtm exports: the module runs on an unmodified IREE runtime")
    args = ap.parse_args(argv)
    text = generate(args.objects_dir, stock_driver=args.stock_driver)
    if args.output == "-":
        sys.stdout.write(text)
    else:
        with open(args.output, "w") as f:
            f.write(text)


if __name__ == "__main__":
    main()
