"""Compiler-free IREE modules around the B200 kernels: VM assembly -> `iree-as-module` -> .vmfb.

  python -m iree_b200.vmasm -o module.vmasm [--cda1-only-used]
  iree-as-module --output=module.vmfb module.vmasm            (the reference's assembler)
  iree-run-module --device=cuda --module=module.vmfb --function=matmul_bf16 \\
      --input=8192x8192xbf16=@a.bin --input=8192x8192xbf16=@b.bin --output=@c.bin

There is no iree-compile in this environment (no LLVM), so the host half of what the compiler
would emit for a module whose dispatches were replaced by the transform spec
(compiler_plugin/gen_spec.py --stock-driver) is written here as textual VM assembly
(runtime/src/iree/vm/bytecode/assembler/, syntax as in runtime/src/iree/vm/test/*.vmasm) and
turned into a real .vmfb by the reference's own `iree-as-module` (tools/iree-as-module-main.c).
The module

  * embeds a CDA1 executable (iree_b200/cda1.py) holding the kernels it dispatches, as rodata;
  * `__init`: hal.devices.get(0), hal.executable.create(format "PTXE") once, kept in globals;
  * one exported function per workload: !hal.buffer_view operands in, result out —
    hal.allocator.allocate, hal.command_buffer.create / dispatch / execution_barrier / finalize,
    hal.device.queue.execute with a signal fence, hal.fence.await, hal.buffer_view.create —
    the import signatures of runtime/src/iree/modules/hal/exports.inl:27-89.

Only exports with the plain custom-dispatch ABI are used (`*_dtm` tcgen05 GEMMs and prefill
attention, decode-cluster attention, generic matmul), so the .vmfb runs on the UNMODIFIED
reference runtime: stock iree-run-module / iree-benchmark-module, stock CUDA HAL driver.
"""

from __future__ import annotations

import argparse
import json
import os
import struct
import sys

from . import cda1

HERE = os.path.dirname(os.path.abspath(__file__))
BUILD = os.path.join(HERE, "_build")

# (import name, calling convention) — runtime/src/iree/modules/hal/exports.inl.
IMPORTS = [
    ("hal.devices.get", "0i_r"),
    ("hal.executable.create", "0rIrrr_r"),
    ("hal.executable.lookup.function", "0rr_I"),
    ("hal.buffer_view.buffer", "0r_r"),
    ("hal.device.allocator", "0r_r"),
    ("hal.allocator.allocate", "0rIiiI_r"),
    ("hal.device.queue.alloca", "0rIrrIiiII_r"),
    ("hal.command_buffer.create", "0riiIi_r"),
    ("hal.command_buffer.dispatch", "0rrIiiiICiDCiirIID_v"),
    ("hal.command_buffer.execution_barrier", "0riiI_v"),
    ("hal.command_buffer.finalize", "0r_v"),
    ("hal.fence.create", "0rI_r"),
    ("hal.device.queue.execute", "0rIrrrI_v"),
    ("hal.fence.await", "0iICrD_i"),
    ("hal.buffer_view.create", "0rIIiiCID_r"),
]

# iree/hal/buffer.h, command_buffer.h, buffer_view.h.
MEMORY_TYPE_DEVICE_LOCAL = (1 << 4) | (1 << 5)
BUFFER_USAGE = 0x3 | (0x3 << 10)          # TRANSFER | DISPATCH_STORAGE
COMMAND_BUFFER_MODE_ONE_SHOT = 1
COMMAND_CATEGORY_DISPATCH = 2
STAGE_DISPATCH, STAGE_COMMAND_RETIRE = 1 << 2, 1 << 4
ENCODING_DENSE_ROW_MAJOR = 1
ELEMENT_TYPE = {"bf16": (0x22 << 24) | 16, "f16": (0x21 << 24) | 16, "f32": (0x21 << 24) | 32,
                "i8": (0x10 << 24) | 8, "i32": (0x10 << 24) | 32}
ELEMENT_BYTES = {"bf16": 2, "f16": 2, "f32": 4, "i8": 1, "i32": 4}


class _Func:
    """Straight-line function body with hand-rolled register numbering: i32 registers %iN,
    i64 values in aligned pairs %iN:N+1, refs %rN; nothing is reused."""

    def __init__(self, name, arg_refs=0, result="()"):
        self.name, self.lines = name, []
        self.n_i, self.n_r = 0, arg_refs
        args = ", ".join(f"%r{i}: !vm.ref<?>" for i in range(arg_refs))
        self.head = f"vm.func @{name}({args}) -> {result} {{"

    def i32(self, value):
        r = f"%i{self.n_i}"
        self.n_i += 1
        self.lines.append(f"  {r} = vm.const.i32 {int(value)}")
        return r

    def i64(self, value):
        self.n_i += self.n_i & 1
        r = f"%i{self.n_i}:{self.n_i + 1}"
        self.n_i += 2
        self.lines.append(f"  {r} = vm.const.i64 {int(value)}")
        return r

    def new_i64(self):
        self.n_i += self.n_i & 1
        r = f"%i{self.n_i}:{self.n_i + 1}"
        self.n_i += 2
        return r

    def new_i32(self):
        r = f"%i{self.n_i}"
        self.n_i += 1
        return r

    def new_ref(self):
        r = f"%r{self.n_r}"
        self.n_r += 1
        return r

    def rodata(self, ordinal):
        r = self.new_ref()
        self.lines.append(f"  {r} = vm.const.ref.rodata {ordinal}")
        return r

    def null(self):
        r = self.new_ref()
        self.lines.append(f"  {r} = vm.const.ref.zero")
        return r

    def call(self, target, args, result=None, segments=None):
        op = "vm.call" if segments is None else "vm.call.variadic"
        seg = "" if segments is None else " segments[%s]" % ", ".join(str(s) for s in segments)
        lhs = f"{result} = " if result else ""
        self.lines.append(f"  {lhs}{op} @{target}{seg}({', '.join(args)})")
        return result

    def emit(self, line):
        self.lines.append("  " + line)

    def ret(self, result=None):
        """The VM only sweeps a frame's ref registers on failure (bytecode/dispatch.c:121-127:
        "on success paths the compiler guarantees all refs are explicitly released"): drop every
        ref register — arguments included — and MOVE the result out."""
        others = [f"%r{i}" for i in range(self.n_r) if f"%r{i}" != result]
        if others:
            self.emit("vm.discard.refs " + ", ".join(others))
        self.emit("vm.return" + (f" %R{result[2:]}" if result else ""))

    def text(self, export=True):
        out = [f"vm.export @{self.name}"] if export else []
        return "\n".join(out + [self.head, "^bb0:"] + self.lines + ["}"]) + "\n"


def _hex(data: bytes) -> str:
    return data.hex().upper()


class Workload:
    """One exported function = one dispatch: |inputs| / |output| are (shape, element type);
    bindings are the inputs in order, then the output (all our exports put the result last);
    |kernel|, |grid| and |constants| come from iree_b200_ops_select_* (the same arithmetic the
    transform spec's count regions carry)."""

    def __init__(self, name, kernel, grid, constants, inputs, output):
        self.name, self.kernel, self.grid = name, kernel, [int(g) for g in grid]
        self.constants = [int(c) & 0xFFFFFFFF for c in constants]
        self.inputs, self.output = inputs, output


def default_workloads(sm_count=148):
    """BASELINE.json's single-GPU configurations, through exports the stock driver launches."""
    from . import hal as H
    ops = H.Ops.offline(sm_count)
    name_of = lambda ordinal: ops.executable.export_info(ordinal)["kernel_name"]
    out = []
    for t, acc in (("bf16", "f32"), ("i8", "i32"), ("f16", "f32")):
        M = N = K = 8192
        sel = ops.select_matmul(t, 1, M, N, K, flags=H.MATMUL_DEVICE_MAPS)
        out.append(Workload(f"matmul_{t}", name_of(sel["export_ordinal"]), sel["workgroup_count"],
                            sel["constants"], [((M, K), t), ((K, N), t)], ((M, N), acc)))
    sel = ops.select_matmul("bf16", 64, 2048, 2048, 2048, flags=H.MATMUL_DEVICE_MAPS)
    out.append(Workload("batch_matmul_bf16", name_of(sel["export_ordinal"]), sel["workgroup_count"],
                        sel["constants"], [((64, 2048, 2048), "bf16")] * 2, ((64, 2048, 2048), "f32")))
    # A ragged shape small enough to check element by element (and the 1-CTA family).
    sel = ops.select_matmul("bf16", 1, 1000, 520, 264, flags=H.MATMUL_DEVICE_MAPS)
    out.append(Workload("matmul_bf16_small", name_of(sel["export_ordinal"]), sel["workgroup_count"],
                        sel["constants"], [((1000, 264), "bf16"), ((264, 520), "bf16")],
                        ((1000, 520), "f32")))
    B, D = 32, 128
    for name, M, K2, kind, flags in (("attention_prefill_f16", 4096, 4096, H.ATTENTION_MASK_NONE,
                                      H.ATTENTION_DEVICE_MAPS),
                                     ("attention_prefill_causal_f16", 4096, 4096,
                                      H.ATTENTION_MASK_CAUSAL, H.ATTENTION_DEVICE_MAPS),
                                     ("attention_decode_f16", 1, 8192, H.ATTENTION_MASK_NONE, 0)):
        sel = ops.select_attention("f16", B, M, D, K2, D, 1.0 / D ** 0.5, mask_kind=kind, flags=flags)
        (d,) = sel["dispatches"]
        out.append(Workload(name, name_of(d["export_ordinal"]), d["workgroup_count"], d["constants"],
                            [((B, M, D), "f16"), ((B, K2, D), "f16"), ((B, K2, D), "f16")],
                            ((B, M, D), "f16")))
    # Small ragged cases the tests check element by element against the oracle.
    sel = ops.select_matmul("i8", 1, 1000, 528, 272, flags=H.MATMUL_DEVICE_MAPS)
    out.append(Workload("matmul_i8_small", name_of(sel["export_ordinal"]), sel["workgroup_count"],
                        sel["constants"], [((1000, 272), "i8"), ((272, 528), "i8")], ((1000, 528), "i32")))
    # The bf16 linear layer as ONE dispatch: matmul + bias[n] + relu + truncf.
    fused = H.MATMUL_BIAS | H.MATMUL_RELU | H.MATMUL_TRUNCATE_RESULT | H.MATMUL_DEVICE_MAPS
    for name, (M, K, N) in (("linear_bf16", (8192, 8192, 8192)), ("linear_bf16_small", (1000, 264, 520))):
        sel = ops.select_matmul("bf16", 1, M, N, K, flags=fused)
        out.append(Workload(name, name_of(sel["export_ordinal"]), sel["workgroup_count"],
                            sel["constants"], [((M, K), "bf16"), ((K, N), "bf16"), ((N,), "f32")],
                            ((M, N), "bf16")))
    sel = ops.select_attention("f16", 2, 384, D, 640, D, 1.0 / D ** 0.5, mask_kind=H.ATTENTION_MASK_CAUSAL,
                               flags=H.ATTENTION_DEVICE_MAPS)
    (d,) = sel["dispatches"]
    out.append(Workload("attention_prefill_causal_f16_small", name_of(d["export_ordinal"]),
                        d["workgroup_count"], d["constants"],
                        [((2, 384, D), "f16"), ((2, 640, D), "f16"), ((2, 640, D), "f16")],
                        ((2, 384, D), "f16")))
    return out


def executable_for(workloads, build_dir=BUILD) -> bytes:
    """A CDA1 holding just the cubins the workloads' kernels live in, every export of those."""
    with open(os.path.join(build_dir, "exports.json")) as f:
        exports = json.load(f)["exports"]
    by_name = {e["name"]: e for e in exports}
    modules = []
    for w in workloads:
        m = by_name[w.kernel]["module"]
        if m not in modules:
            modules.append(m)
    images = []
    for m in modules:
        with open(os.path.join(build_dir, m + ".cubin"), "rb") as f:
            images.append(f.read())
    defs = [cda1.ExportDef(kernel_name=e["name"], block_dims=e["block_dims"],
                           block_shared_memory_size=e["shared_memory"],
                           constant_count=e["constants"], binding_flags=e["binding_flags"],
                           module_ordinal=modules.index(e["module"]))
            for e in exports if e["module"] in modules]
    return cda1.serialize(cda1.ExecutableDef(defs, images))


def _function(w, name_rodata, into, result_alloc):
    if True:
        # |into|: destination-passing form `<name>_into(operands..., result) -> result` (the
        # caller owns the result storage: no allocation per call), else the result is allocated.
        f = _Func(w.name + ("_into" if into else ""), arg_refs=len(w.inputs) + int(into),
                  result="(!vm.ref<?>)")
        device, executable = f.new_ref(), f.new_ref()
        f.emit(f"{device} = vm.global.load.ref .refs[0] : !hal.device")
        f.emit(f"{executable} = vm.global.load.ref .refs[1] : !hal.executable")
        function = f.call("hal.executable.lookup.function", [executable, f.rodata(name_rodata)], f.new_i64())
        buffers = [f.call("hal.buffer_view.buffer", [f"%r{j}"], f.new_ref())
                   for j in range(len(w.inputs))]
        any_queue, zero64, zero32 = f.i64(-1), f.i64(0), f.i32(0)
        shape, et = w.output
        out_bytes = ELEMENT_BYTES[et]
        for d in shape:
            out_bytes *= d
        size = f.i64(out_bytes)
        if into:
            allocated = f.null()
            result = f.call("hal.buffer_view.buffer", [f"%r{len(w.inputs)}"], f.new_ref())
        elif result_alloc == "queue":
            allocated = f.call("hal.fence.create", [device, zero64], f.new_ref())
            result = f.call("hal.device.queue.alloca",
                            [device, any_queue, f.null(), allocated, zero64,
                             f.i32(MEMORY_TYPE_DEVICE_LOCAL), f.i32(BUFFER_USAGE), size, zero64],
                            f.new_ref())
        else:
            allocated = f.null()
            allocator = f.call("hal.device.allocator", [device], f.new_ref())
            result = f.call("hal.allocator.allocate",
                            [allocator, any_queue, f.i32(MEMORY_TYPE_DEVICE_LOCAL), f.i32(BUFFER_USAGE),
                             size], f.new_ref())
        cb = f.call("hal.command_buffer.create",
                    [device, f.i32(COMMAND_BUFFER_MODE_ONE_SHOT), f.i32(COMMAND_CATEGORY_DISPATCH),
                     any_queue, zero32], f.new_ref())
        # (cb, executable, function, x, y, z, flags, constants..., (reserved, slot, buffer, offset, length)...)
        args = [cb, executable, function] + [f.i32(g) for g in w.grid] + [zero64]
        # i32 constants as signed literals
        args += [f.i32(c - (1 << 32) if c >= (1 << 31) else c) for c in w.constants]
        sizes = []
        for (s, t) in w.inputs:
            n = ELEMENT_BYTES[t]
            for d in s:
                n *= d
            sizes.append(n)
        for buf, n in list(zip(buffers, sizes)) + [(result, out_bytes)]:
            args += [zero32, zero32, buf, zero64, f.i64(n)]
        f.call("hal.command_buffer.dispatch", args,
               segments=[1] * 7 + [len(w.constants), len(w.inputs) + 1])
        f.call("hal.command_buffer.execution_barrier",
               [cb, f.i32(STAGE_DISPATCH | STAGE_COMMAND_RETIRE),
                f.i32(1 | STAGE_DISPATCH), zero64])
        f.call("hal.command_buffer.finalize", [cb])
        fence = f.call("hal.fence.create", [device, zero64], f.new_ref())
        f.call("hal.device.queue.execute", [device, any_queue, allocated, fence, cb, zero64])
        status = f.call("hal.fence.await", [f.i32(-1), zero64, fence], f.new_i32(), segments=[1, 1, 1])
        f.emit(f"vm.cond_br {status}, ^bb1(), ^bb2()")
        f.emit_label = None
        f.lines.append("^bb1:")
        f.emit(f'vm.fail {status}, "hal.fence.await failed"')
        f.lines.append("^bb2:")
        if into:
            f.ret(f"%r{len(w.inputs)}")
            return f.text()
        view = f.call("hal.buffer_view.create",
                      [result, zero64, size, f.i32(ELEMENT_TYPE[et]), f.i32(ENCODING_DENSE_ROW_MAJOR)]
                      + [f.i64(d) for d in shape], f.new_ref(), segments=[1] * 5 + [len(shape)])
        f.ret(view)
        return f.text()


def generate(workloads=None, build_dir=BUILD, module_name="iree_b200_module",
             result_alloc="queue") -> str:
    """|result_alloc|: "queue" = hal.device.queue.alloca (stream-ordered, from the driver's
    memory pool; what the compiler emits for results on targets with async allocation),
    "allocator" = hal.allocator.allocate (synchronous cuMemAlloc)."""
    workloads = default_workloads() if workloads is None else workloads
    blob = executable_for(workloads, build_dir)
    out = [f"vm.module @{module_name} version 0", "vm.type i32", "vm.type i64", "vm.type !vm.buffer",
           "vm.type !hal.device", "vm.type !hal.executable",
           f"vm.rodata[0] = {_hex(b'PTXE')}", f"vm.rodata[1] = {_hex(blob)}"]
    for i, w in enumerate(workloads):
        out.append(f"vm.rodata[{2 + i}] = {_hex(w.kernel.encode())}")
    out += [f'vm.import @{name} cconv "{cconv}"' for name, cconv in IMPORTS]
    out.append("")

    # Module initialiser: device 0 and the executable, once.
    f = _Func("__init")
    zero = f.i32(0)
    device = f.call("hal.devices.get", [zero], f.new_ref())
    f.emit(f"vm.global.store.ref {device}, .refs[0] : !hal.device")
    any_queue = f.i64(-1)
    executable = f.call("hal.executable.create",
                        [device, any_queue, f.rodata(0), f.rodata(1), f.null()], f.new_ref())
    f.emit(f"vm.global.store.ref {executable}, .refs[1] : !hal.executable")
    f.ret()
    out.append(f.text())

    for i, w in enumerate(workloads):
        for into in (False, True):
            out.append(_function(w, 2 + i, into, result_alloc))
    return "\n".join(out)


# BytecodeModuleDef field slots (runtime/src/iree/schemas/bytecode_module_def.fbs:246-305) that
# hold offsets (the others — version, requirements, bytecode_version — are inline scalars).
_ROOT_OFFSET_FIELDS = {0, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 14, 15}
_DEPENDENCIES_FIELD = 5


def add_dependencies(vmfb: bytes, names=("hal",)) -> bytes:
    """Returns |vmfb| with `dependencies:[ModuleDependencyDef]` listing |names|.

    iree-as-module writes imports but no dependency table, and the stock tools
    (tooling/context_util.c:700-741) register the `hal` module only for modules that declare
    the dependency — iree-compile always does. The root table of the (size-prefixed) FlatBuffer
    sits at the front, so a new root table with the extra slot, its vtable and the dependency
    objects are spliced in right after the 12-byte header; everything else moves by a constant
    and keeps its relative offsets."""
    size, root_rel, ident = struct.unpack_from("<II4s", vmfb, 0)
    if ident != b"IREE" or size + 4 != len(vmfb):
        raise ValueError("not a size-prefixed IREE bytecode module")
    root = 4 + root_rel
    vt = root - struct.unpack_from("<i", vmfb, root)[0]
    vt_len, table_len = struct.unpack_from("<HH", vmfb, vt)
    slots = [struct.unpack_from("<H", vmfb, vt + 4 + 2 * i)[0] for i in range((vt_len - 4) // 2)]
    slots += [0] * (16 - len(slots))
    if slots[_DEPENDENCIES_FIELD]:
        raise ValueError("module already declares dependencies")
    body_start = 12          # what follows the header moves as one block
    # New block, laid out at file offset 12: vtable, root table, vector, tables, strings.
    field_ids = [i for i, off in enumerate(slots) if off] + [_DEPENDENCIES_FIELD]
    field_ids.sort(key=lambda i: (i == _DEPENDENCIES_FIELD, slots[i]))   # old order, new one last
    new_vt_at = 12
    new_vt_len = 4 + 2 * 16
    new_root_at = (new_vt_at + new_vt_len + 7) // 8 * 8
    new_slots, cursor = [0] * 16, 4
    widths = {}
    for i in field_ids:
        widths[i] = 4                     # every present field here is a u32 or a uoffset
        new_slots[i] = cursor
        cursor += 4
    new_table_len = cursor
    vec_at = new_root_at + new_table_len
    dep_tables_at = vec_at + 4 + 4 * len(names)
    # ModuleDependencyDef: vtable {6, 8, name@4}, table {soffset, name uoffset}.
    dep_vt_at = dep_tables_at
    dep_vt = struct.pack("<HHH", 6, 8, 4) + b"\0\0"
    tables_at = dep_vt_at + len(dep_vt)
    strings_at = tables_at + 8 * len(names)
    string_blobs, string_pos = [], []
    pos = strings_at
    for n in names:
        raw = n.encode()
        blob = struct.pack("<I", len(raw)) + raw + b"\0"
        blob += b"\0" * (-len(blob) % 4)
        string_pos.append(pos)
        string_blobs.append(blob)
        pos += len(blob)
    block_end = pos
    shift = (block_end - body_start + 63) // 64 * 64       # keeps every alignment of the body
    out = bytearray(12 + shift) + vmfb[body_start:]
    struct.pack_into("<II4s", out, 0, len(out) - 4, new_root_at - 4, b"IREE")
    struct.pack_into("<HH", out, new_vt_at, new_vt_len, new_table_len)
    for i, off in enumerate(new_slots):
        struct.pack_into("<H", out, new_vt_at + 4 + 2 * i, off)
    struct.pack_into("<i", out, new_root_at, new_root_at - new_vt_at)
    for i in field_ids:
        at = new_root_at + new_slots[i]
        if i == _DEPENDENCIES_FIELD:
            struct.pack_into("<I", out, at, vec_at - at)
            continue
        old_at = root + slots[i]
        value = struct.unpack_from("<I", vmfb, old_at)[0]
        if i in _ROOT_OFFSET_FIELDS:
            value = (old_at + value + shift) - at
        struct.pack_into("<I", out, at, value)
    struct.pack_into("<I", out, vec_at, len(names))
    out[dep_vt_at:dep_vt_at + len(dep_vt)] = dep_vt
    for j in range(len(names)):
        t = tables_at + 8 * j
        struct.pack_into("<I", out, vec_at + 4 + 4 * j, t - (vec_at + 4 + 4 * j))
        struct.pack_into("<iI", out, t, t - dep_vt_at, string_pos[j] - (t + 4))
        out[string_pos[j]:string_pos[j] + len(string_blobs[j])] = string_blobs[j]
    return bytes(out)


def assemble(vmasm_path: str, vmfb_path: str, assembler: str):
    """vmasm -> vmfb with the reference's iree-as-module, then the dependency table."""
    import subprocess
    subprocess.run([assembler, f"--output={vmfb_path}", vmasm_path], check=True)
    with open(vmfb_path, "rb") as f:
        data = f.read()
    with open(vmfb_path, "wb") as f:
        f.write(add_dependencies(data))


MODULE_VMFB = os.path.join(BUILD, "iree_b200_module.vmfb")
DEFAULT_ASSEMBLER = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "iree-rt", "stock",
                                 "iree-as-module")


def build_module(assembler=DEFAULT_ASSEMBLER, force=False):
    """iree_b200/_build/iree_b200_module.vmfb (travels to the GPU box with the other build
    outputs); rebuilt when the kernels' manifest or this file is newer. Returns the path, or
    None when the reference's assembler has not been built (tools/build_ref_runtime.sh)."""
    if not os.path.exists(assembler):
        return None
    sources = [os.path.join(BUILD, "exports.json"), os.path.abspath(__file__), assembler]
    if (not force and os.path.exists(MODULE_VMFB) and
            os.path.getmtime(MODULE_VMFB) >= max(os.path.getmtime(p) for p in sources)):
        return MODULE_VMFB
    vmasm_path = os.path.splitext(MODULE_VMFB)[0] + ".vmasm"
    with open(vmasm_path, "w") as f:
        f.write(generate())
    assemble(vmasm_path, MODULE_VMFB, assembler)
    os.remove(vmasm_path)     # 2x the executable as hex text; regenerate with -o when wanted
    return MODULE_VMFB


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("-o", "--output", required=True)
    ap.add_argument("--build-dir", default=BUILD)
    ap.add_argument("--result-alloc", default="queue", choices=["queue", "allocator"])
    ap.add_argument("--assembler", default="",
                    help="path of iree-as-module: also write <output minus .vmasm>.vmfb")
    args = ap.parse_args(argv)
    text = generate(build_dir=args.build_dir, result_alloc=args.result_alloc)
    with open(args.output, "w") as f:
        f.write(text)
    print(f"wrote {args.output} ({len(text)} bytes)")
    if args.assembler:
        vmfb = os.path.splitext(args.output)[0] + ".vmfb"
        assemble(args.output, vmfb, args.assembler)
        print(f"wrote {vmfb} ({os.path.getsize(vmfb)} bytes)")


if __name__ == "__main__":
    main()
