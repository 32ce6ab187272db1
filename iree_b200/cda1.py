"""Writer/reader for IREE's CUDA executable container ("CDA1").

Compile-side tooling: this is the job `serializeExecutable` does in the reference compiler
for an external (`hal.executable.object`) variant
(compiler/plugins/target/CUDA/CUDATarget.cpp:505-525,627-721): wrap the object bytes
verbatim as `ModuleDef.ptx_image`, emit one `ExportDef` per export ordinal, and prefix the
flatbuffer with the 64-byte `iree_flatbuffer_file_header_t`
(runtime/src/iree/base/internal/flatcc/parsing.h:41-53). Schema:
runtime/src/iree/schemas/cuda_executable_def.fbs. There is no `iree-compile` in this
environment (no LLVM), so the blob is produced here; the run-time side
(iree_b200/csrc/hal/executable.cc) and the reference's
iree_hal_cuda_native_executable_create consume the same bytes.

Differences from what the reference compiler emits: several modules per executable are
used (the schema and the reference loader allow it, native_executable.c:360-400; the
compiler happens to emit one), and no debug info / source files are attached.
"""

from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import List, Sequence

FILE_IDENTIFIER = b"CDA1"
HEADER_SIZE = 64
BINDING_READ_ONLY = 1  # BindingBits.READ_ONLY (cuda_executable_def.fbs:23-26)
BINDING_INDIRECT = 2


@dataclass
class ExportDef:
    kernel_name: str
    block_dims: Sequence[int]
    block_shared_memory_size: int = 0
    constant_count: int = 0
    binding_flags: Sequence[int] = field(default_factory=list)
    module_ordinal: int = 0


@dataclass
class ExecutableDef:
    exports: List[ExportDef]
    modules: List[bytes]  # PTX text or cubin/fatbin bytes, loaded with cuModuleLoadDataEx


class _Builder:
    """Front-to-back flatbuffer writer (offsets are patched once targets are placed)."""

    def __init__(self):
        self.buf = bytearray()

    def align(self, n: int, bias: int = 0):
        while (len(self.buf) + bias) % n:
            self.buf.append(0)

    def pos(self) -> int:
        return len(self.buf)

    def put(self, fmt: str, *values) -> int:
        p = len(self.buf)
        self.buf += struct.pack("<" + fmt, *values)
        return p

    def patch_uoffset(self, field_pos: int, target_pos: int):
        assert target_pos > field_pos
        struct.pack_into("<I", self.buf, field_pos, target_pos - field_pos)

    def vtable(self, table_size: int, field_offsets: Sequence[int]) -> int:
        self.align(2)
        p = self.pos()
        self.put("HH", 4 + 2 * len(field_offsets), table_size)
        for off in field_offsets:
            self.put("H", off)
        return p

    def string(self, data: bytes, data_align: int = 4) -> int:
        # The length prefix sits 4 bytes before the (aligned) character data.
        self.align(max(4, data_align), bias=4)
        p = self.put("I", len(data))
        self.buf += data
        self.buf.append(0)
        return p


def serialize(executable: ExecutableDef) -> bytes:
    """ExecutableDef -> header-prefixed flatbuffer bytes (a `.cda1` file / PTXE blob)."""
    if not executable.modules:
        raise ValueError("executable has no modules")
    for i, e in enumerate(executable.exports):
        if not 0 <= e.module_ordinal < len(executable.modules):
            raise ValueError(f"exports[{i}] module_ordinal out of range")
        if len(e.block_dims) != 3:
            raise ValueError(f"exports[{i}] block_dims must have 3 entries")
        if not e.kernel_name:
            raise ValueError(f"exports[{i}] has an empty kernel name")

    b = _Builder()
    root_field = b.put("I", 0)
    b.buf += FILE_IDENTIFIER

    # ExecutableDef: fields 0 exports, 1 modules (2 source_files omitted).
    root_vt = b.vtable(12, [4, 8, 0])
    b.align(4)
    root = b.put("i", 0)
    struct.pack_into("<i", b.buf, root, root - root_vt)
    exports_field = b.put("I", 0)
    modules_field = b.put("I", 0)
    b.patch_uoffset(root_field, root)

    b.align(4)
    exports_vec = b.put("I", len(executable.exports))
    export_slots = [b.put("I", 0) for _ in executable.exports]
    b.patch_uoffset(exports_field, exports_vec)

    b.align(4)
    modules_vec = b.put("I", len(executable.modules))
    module_slots = [b.put("I", 0) for _ in executable.modules]
    b.patch_uoffset(modules_field, modules_vec)

    # ExportDef tables. Inline layout: soffset, module_ordinal, kernel_name, block_dims
    # (struct, 12 B), block_shared_memory_size, constant_count, binding_flags.
    export_vt = b.vtable(36, [4, 8, 12, 24, 28, 32, 0])
    name_fields, flag_fields = [], []
    for slot, e in zip(export_slots, executable.exports):
        b.align(4)
        t = b.put("i", 0)
        struct.pack_into("<i", b.buf, t, t - export_vt)
        b.put("I", e.module_ordinal)
        name_fields.append(b.put("I", 0))
        b.put("III", *[int(d) for d in e.block_dims])
        b.put("I", int(e.block_shared_memory_size))
        b.put("I", int(e.constant_count))
        flag_fields.append(b.put("I", 0))
        b.patch_uoffset(slot, t)
    for e, name_field, flag_field in zip(executable.exports, name_fields, flag_fields):
        b.patch_uoffset(name_field, b.string(e.kernel_name.encode()))
        b.align(8, bias=4)  # uint64 elements start 8-byte aligned
        v = b.put("I", len(e.binding_flags))
        for f in e.binding_flags:
            b.put("Q", int(f))
        b.patch_uoffset(flag_field, v)

    # ModuleDef tables: soffset, ptx_image.
    module_vt = b.vtable(8, [4])
    image_fields = []
    for slot in module_slots:
        b.align(4)
        t = b.put("i", 0)
        struct.pack_into("<i", b.buf, t, t - module_vt)
        image_fields.append(b.put("I", 0))
        b.patch_uoffset(slot, t)
    for image_field, image in zip(image_fields, executable.modules):
        if not image:
            raise ValueError("module image is empty")
        # 16-byte aligned image bytes (ELF loaders like that; flatbuffers only needs 4).
        b.patch_uoffset(image_field, b.string(bytes(image), data_align=16))
    b.align(16)

    content = bytes(b.buf)
    header = FILE_IDENTIFIER + struct.pack("<IQ", 0, len(content)) + b"\0" * 48
    assert len(header) == HEADER_SIZE
    return header + content


def deserialize(blob: bytes) -> ExecutableDef:
    """Inverse of serialize(); raises ValueError on malformed input (used by tests/tools)."""
    if len(blob) < HEADER_SIZE or blob[:4] != FILE_IDENTIFIER:
        raise ValueError("not a CDA1 blob")
    version, content_size = struct.unpack_from("<IQ", blob, 4)
    if version != 0:
        raise ValueError(f"unsupported version {version}")
    if content_size == 0 or content_size > len(blob) - HEADER_SIZE:
        raise ValueError("content size out of bounds")
    fb = memoryview(blob)[HEADER_SIZE:HEADER_SIZE + content_size]
    if bytes(fb[4:8]) != FILE_IDENTIFIER:
        raise ValueError("flatbuffer identifier mismatch")

    def u16(p): return struct.unpack_from("<H", fb, p)[0]
    def u32(p): return struct.unpack_from("<I", fb, p)[0]
    def i32(p): return struct.unpack_from("<i", fb, p)[0]
    def u64(p): return struct.unpack_from("<Q", fb, p)[0]

    def field_pos(table, index):
        vt = table - i32(table)
        if 4 + 2 * index + 2 > u16(vt):
            return 0
        off = u16(vt + 4 + 2 * index)
        return table + off if off else 0

    def follow(p): return p + u32(p)

    def string(p):
        n = u32(p)
        return bytes(fb[p + 4:p + 4 + n])

    try:
        root = follow(0)
        modules = []
        p = field_pos(root, 1)
        if p:
            vec = follow(p)
            for i in range(u32(vec)):
                t = follow(vec + 4 + 4 * i)
                modules.append(string(follow(field_pos(t, 0))))
        exports = []
        p = field_pos(root, 0)
        if p:
            vec = follow(p)
            for i in range(u32(vec)):
                t = follow(vec + 4 + 4 * i)
                f = field_pos(t, 0)
                module_ordinal = u32(f) if f else 0
                name = string(follow(field_pos(t, 1))).decode()
                f = field_pos(t, 2)
                dims = [u32(f), u32(f + 4), u32(f + 8)] if f else [0, 0, 0]
                f = field_pos(t, 3)
                smem = u32(f) if f else 0
                f = field_pos(t, 4)
                constants = u32(f) if f else 0
                flags = []
                f = field_pos(t, 5)
                if f:
                    v = follow(f)
                    flags = [u64(v + 4 + 8 * j) for j in range(u32(v))]
                exports.append(ExportDef(name, dims, smem, constants, flags, module_ordinal))
    except (struct.error, IndexError) as e:
        raise ValueError(f"malformed CDA1 flatbuffer: {e}") from e
    return ExecutableDef(exports, modules)
