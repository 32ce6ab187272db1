"""CPU-side checks of the drop-in boundary: the library loads without a GPU, exports
every symbol the headers declare, treats executable blobs as untrusted, and refuses to
compute without a device (there is no CPU fallback)."""

import ctypes
import os
import re

import pytest

from iree_b200 import build, cda1, hal

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    for header in ("iree_b200_hal.h", "iree_b200_ops.h"):
        with open(os.path.join(ROOT, "include", header)) as f:
            text = f.read()
        text = re.sub(r"//[^\n]*", "", text)
        for m in re.finditer(r"IREE_B200_API\s[^;{]*?\b(iree_b200_\w+)\s*\(", text, re.S):
            names.add(m.group(1))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(build.build())
    declared = _declared_symbols()
    assert len(declared) >= 55
    missing = [n for n in declared if not hasattr(lib, n)]
    assert not missing, missing
    # ... and the Python mirror binds exactly that set.
    assert sorted(hal.ALL_SYMBOLS) == declared


def test_status_code_strings_match_reference_names():
    lib = hal.lib()
    assert lib.iree_b200_status_code_string(0) == b"OK"
    assert lib.iree_b200_status_code_string(3) == b"INVALID_ARGUMENT"
    assert lib.iree_b200_status_code_string(14) == b"UNAVAILABLE"


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_have_gpu(), reason="checks the no-GPU failure mode")
def test_no_gpu_fails_loudly_no_cpu_fallback():
    with pytest.raises(hal.Status) as e:
        hal.Device(0)
    assert e.value.name == "UNAVAILABLE"


def test_builtin_executable_is_valid_cda1():
    size = ctypes.c_size_t()
    ptr = hal.lib().iree_b200_ops_builtin_executable_data(ctypes.byref(size))
    blob = ctypes.string_at(ptr, size.value)
    with open(build.cda1_path(), "rb") as f:
        assert f.read() == blob
    fmt, inferred = hal.infer_format(blob)
    assert fmt == "PTXE" and inferred == len(blob)
    exe = cda1.deserialize(blob)
    names = [e.kernel_name for e in exe.exports]
    assert "iree_b200_matmul_bf16_f32" in names and "iree_b200_matmul_generic_i8_i32" in names
    assert "iree_b200_attention_prefill_f16" in names and "iree_b200_attention_decode_merge_bf16" in names
    for e in exe.exports:
        assert list(e.block_dims) in ([256, 1, 1], [384, 1, 1], [128, 1, 1], [288, 1, 1], [1024, 1, 1])
        assert e.block_shared_memory_size <= 227 * 1024
        assert e.constant_count <= 64 and len(e.binding_flags) <= 16
    assert all(m[:4] == b"\x7fELF" for m in exe.modules)  # cubins, not PTX (no JIT at load)


def test_cda1_roundtrip_and_untrusted_input():
    exe = cda1.ExecutableDef(
        exports=[cda1.ExportDef("k0", [128, 1, 1], 1024, 2, [1, 0], 0),
                 cda1.ExportDef("k1", [32, 2, 1], 0, 0, [0], 1)],
        modules=[b"// ptx a\n", b"\x7fELFxxxx"])
    blob = cda1.serialize(exe)
    back = cda1.deserialize(blob)
    assert [e.kernel_name for e in back.exports] == ["k0", "k1"]
    assert back.exports[0].block_shared_memory_size == 1024
    assert list(back.exports[1].block_dims) == [32, 2, 1]
    assert back.modules[1] == b"\x7fELFxxxx"
    assert hal.infer_format(blob) == ("PTXE", len(blob))
    # Corrupt magic / truncated / garbage: status, never a crash.
    for bad in (b"XXXX" + blob[4:], blob[:40], b"", os.urandom(256)):
        with pytest.raises(hal.Status) as e:
            hal.infer_format(bad)
        assert e.value.code != 0


def test_cda1_reader_survives_mutations():
    """A bounded run of the mutation fuzzing tools/fuzz_cda1.sh does under ASAN + UBSAN
    (300 k inputs, no finding): every mutated executable is either still well-formed or
    rejected with a status."""
    import random
    with open(build.cda1_path(), "rb") as f:
        exe = cda1.deserialize(f.read())
    exe.modules = [b"\x7fELF" + bytes([i]) * 11 for i in range(len(exe.modules))]
    seed = cda1.serialize(exe)
    assert hal.infer_format(seed) == ("PTXE", len(seed))
    rng = random.Random(0xB200)
    rejected = 0
    for _ in range(3000):
        blob = bytearray(seed)
        kind = rng.randrange(4)
        if kind == 0:
            blob[rng.randrange(64)] ^= 1 << rng.randrange(8)
        elif kind == 1:
            at = rng.randrange(64, len(blob) - 4) & ~3
            blob[at:at + 4] = rng.choice([0, 1, 4, 0x7FFFFFFF, 0x80000000, 0xFFFFFFFF,
                                          rng.getrandbits(32)]).to_bytes(4, "little")
        elif kind == 2:
            del blob[rng.randrange(len(blob)):]
        else:
            at = rng.randrange(64, len(blob))
            for i in range(at, min(len(blob), at + rng.randrange(1, 64))):
                blob[i] = rng.getrandbits(8)
        try:
            fmt, size = hal.infer_format(bytes(blob))
            assert fmt == "PTXE" and size <= len(blob)
        except hal.Status as e:
            assert e.code != 0
            rejected += 1
    assert rejected > 500


def _build_c_smoke(tmp_path):
    import shutil
    import subprocess
    cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else shutil.which("gcc")
    exe = tmp_path / "cabi_smoke"
    lib_dir = os.path.dirname(build.build())
    subprocess.run([cc, "-std=c11", "-Wall", "-Wextra", "-Werror", "-pedantic",
                    "-I", os.path.join(ROOT, "include"), "-o", str(exe),
                    os.path.join(ROOT, "tests", "c", "cabi_smoke.c"), "-L", lib_dir, "-liree_b200",
                    f"-Wl,-rpath,{lib_dir}"], check=True)
    return exe


@pytest.mark.skipif(_have_gpu(), reason="the GPU variant below runs instead")
def test_headers_are_c11_and_the_library_links_from_c(tmp_path):
    """tests/c/cabi_smoke.c: include/*.h under -std=c11 -Wall -Wextra -Werror -pedantic (the
    reference's host language is C), linked against the library, run without a GPU."""
    import subprocess
    r = subprocess.run([str(_build_c_smoke(tmp_path))], capture_output=True, text=True)
    assert r.returncode == 0 and "cabi_smoke ok (no gpu)" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_c_program_creates_a_device(tmp_path):
    import subprocess
    r = subprocess.run([str(_build_c_smoke(tmp_path)), "gpu"], capture_output=True, text=True)
    assert r.returncode == 0 and "cabi_smoke ok (gpu)" in r.stdout, r.stdout + r.stderr
