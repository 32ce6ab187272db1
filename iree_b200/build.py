"""In-tree build of the B200 kernels and the C-ABI library.

  python -m iree_b200.build [--force] [--jobs N]

Steps (all outputs under iree_b200/_build/, git-ignored; they travel to the GPU box):
  1. nvcc -cubin (sm_100a, -lineinfo) for every kernel translation unit;
  2. g++ the manifest tool and run it -> exports.json (block dims, dynamic shared memory,
     constant / binding counts; the data a hal.executable.export op carries);
  3. pack cubins + export table into iree_b200_kernels.cda1 (iree_b200/cda1.py) — the
     `#hal.executable.object` payload, also embedded into the library;
  4. g++ the thin launch layer + ops into libiree_b200.so (libcuda / libnccl are dlopen'd
     at run time; nothing links against them, so the library loads without a GPU).
nvcc cross-compiles without a GPU, so this runs in the CPU-only container.
"""

from __future__ import annotations

import argparse
import concurrent.futures
import hashlib
import json
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "_build")
INCLUDE = os.path.join(ROOT, "include")

KERNEL_UNITS = ["gemm_kernels", "gemm_dtm_kernels", "gemm_bias_kernels", "gemm_bias_dtm_kernels", "generic_kernels", "attention_kernels", "collective_kernels"]
# Kernels that have not run on a GPU yet: packed into their own executable, which the ops layer
# loads only when one of their entry points is called (the validated paths never touch it).
EXPERIMENTAL_UNITS = ["mx_kernels", "mmt4d_kernels"]
HOST_SOURCES = ["hal/device.cc", "hal/executable.cc", "hal/command_buffer.cc", "hal/channel.cc",
                "ops/ops.cc"]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--expt-relaxed-constexpr", "-Xptxas", "-v"]
CXX_FLAGS = ["-O2", "-g", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-Wall", "-Wextra",
             "-Wno-unused-parameter", "-Wno-missing-field-initializers"]

LIB_NAME = "libiree_b200.so"
CDA1_NAME = "iree_b200_kernels.cda1"
EXPERIMENTAL_CDA1_NAME = "iree_b200_pending_kernels.cda1"


def _which(*names):
    for n in names:
        p = shutil.which(n)
        if p:
            return p
    raise RuntimeError(f"none of {names} found on PATH")


def _cuda_home():
    for c in (os.environ.get("CUDA_HOME"), "/usr/local/cuda"):
        if c and os.path.exists(os.path.join(c, "include", "cuda.h")):
            return c
    raise RuntimeError("CUDA toolkit headers not found")


def _cxx():
    # The image exports CXX=/opt/gcc/bin/g++ (a wrapper); prefer the system compiler.
    for c in ("/usr/bin/g++", shutil.which("g++")):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("g++ not found")


def _tree_hash(paths):
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(p.encode())
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def _sources():
    out = []
    for base in (CSRC, INCLUDE):
        for d, _, files in os.walk(base):
            for f in files:
                if f.endswith((".cu", ".cuh", ".h", ".cc", ".inc", ".S")):
                    out.append(os.path.join(d, f))
    out += [os.path.join(HERE, "cda1.py"), os.path.abspath(__file__)]
    return out


def _run(cmd, log_name=None):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log_name:
        with open(os.path.join(OUT, log_name), "w") as f:
            f.write(" ".join(cmd) + "\n" + r.stdout)
    if r.returncode != 0:
        raise RuntimeError(f"command failed: {' '.join(cmd)}\n{r.stdout}")
    return r.stdout


def lib_path() -> str:
    return os.path.join(OUT, LIB_NAME)


def cda1_path() -> str:
    return os.path.join(OUT, CDA1_NAME)


def is_current() -> bool:
    stamp = os.path.join(OUT, "stamp")
    if not (os.path.exists(stamp) and os.path.exists(lib_path()) and os.path.exists(cda1_path())):
        return False
    with open(stamp) as f:
        return f.read().strip() == _tree_hash(_sources())


def build(force: bool = False, jobs: int = 0, verbose: bool = False) -> str:
    """Builds everything if sources changed; returns the library path."""
    if not force and is_current():
        return lib_path()
    os.makedirs(OUT, exist_ok=True)
    nvcc = _which("nvcc", "/usr/local/cuda/bin/nvcc")
    cxx = _cxx()
    cuda_inc = os.path.join(_cuda_home(), "include")
    jobs = jobs or min(8, os.cpu_count() or 1)
    from . import cda1

    def compile_cubin(unit):
        src = os.path.join(CSRC, "kernels", unit + ".cu")
        dst = os.path.join(OUT, unit + ".cubin")
        _run([nvcc, *NVCC_FLAGS, "-I", INCLUDE, "-cubin", "-o", dst, src], unit + ".nvcc.log")
        return dst

    def compile_host(rel):
        src = os.path.join(CSRC, rel)
        dst = os.path.join(OUT, rel.replace("/", "_") + ".o")
        _run([cxx, *CXX_FLAGS, "-I", INCLUDE, "-I", cuda_inc, "-c", "-o", dst, src])
        return dst

    def compile_no_alloc():
        # The one nvcc OBJECT (not cubin) of the library: a static __device__ scratch array and
        # its cudaGetSymbolAddress need the runtime's module registration; cudart is linked
        # statically and loads libcuda lazily, so the library still loads without a GPU.
        src = os.path.join(CSRC, "no_alloc", "attention_decode_no_alloc.cu")
        dst = os.path.join(OUT, "attention_decode_no_alloc.o")
        _run([nvcc, *[f for f in NVCC_FLAGS if f not in ("-Xptxas", "-v")], "-I", INCLUDE,
              "--cudart", "static", "-Xcompiler", "-fPIC,-fvisibility=hidden", "-c", "-o", dst, src],
             "attention_decode_no_alloc.nvcc.log")
        return dst

    def manifest():
        tool = os.path.join(OUT, "manifest_tool")
        _run([cxx, "-O1", "-std=c++17", "-I", INCLUDE, "-o", tool,
              os.path.join(CSRC, "tools", "manifest_tool.cc")])
        text = _run([tool])
        with open(os.path.join(OUT, "exports.json"), "w") as f:
            f.write(text)
        return json.loads(text)

    with concurrent.futures.ThreadPoolExecutor(jobs) as pool:
        cubin_futs = [pool.submit(compile_cubin, u) for u in KERNEL_UNITS]
        experimental_futs = [pool.submit(compile_cubin, u) for u in EXPERIMENTAL_UNITS]
        manifest_fut = pool.submit(manifest)
        host_futs = [pool.submit(compile_host, s) for s in HOST_SOURCES]
        host_futs.append(pool.submit(compile_no_alloc))
        cubins = [f.result() for f in cubin_futs]
        experimental_cubins = [f.result() for f in experimental_futs]
        exports_json = manifest_fut.result()
        objects = [f.result() for f in host_futs]

    # Pack the executable.
    modules = []
    for path in cubins:
        with open(path, "rb") as f:
            modules.append(f.read())
    exports, experimental_exports = [], []
    for e in exports_json["exports"]:
        units, into = ((EXPERIMENTAL_UNITS, experimental_exports)
                       if e["module"] in EXPERIMENTAL_UNITS else (KERNEL_UNITS, exports))
        into.append(cda1.ExportDef(
            kernel_name=e["name"], block_dims=e["block_dims"],
            block_shared_memory_size=e["shared_memory"], constant_count=e["constants"],
            binding_flags=e["binding_flags"], module_ordinal=units.index(e["module"])))
    blob = cda1.serialize(cda1.ExecutableDef(exports, modules))
    with open(cda1_path(), "wb") as f:
        f.write(blob)
    experimental_modules = []
    for path in experimental_cubins:
        with open(path, "rb") as f:
            experimental_modules.append(f.read())
    experimental_path = os.path.join(OUT, EXPERIMENTAL_CDA1_NAME)
    with open(experimental_path, "wb") as f:
        f.write(cda1.serialize(cda1.ExecutableDef(experimental_exports, experimental_modules)))

    # Embed the blob and link.
    blob_s = os.path.join(OUT, "builtin_blob.S")
    with open(blob_s, "w") as f:
        f.write('.section .rodata\n.balign 64\n.global iree_b200_builtin_cda1_start\n'
                'iree_b200_builtin_cda1_start:\n.incbin "%s"\n'
                '.global iree_b200_builtin_cda1_end\niree_b200_builtin_cda1_end:\n'
                '.balign 64\n.global iree_b200_builtin_pending_cda1_start\n'
                'iree_b200_builtin_pending_cda1_start:\n.incbin "%s"\n'
                '.global iree_b200_builtin_pending_cda1_end\niree_b200_builtin_pending_cda1_end:\n'
                '.section .note.GNU-stack,"",@progbits\n' % (cda1_path(), experimental_path))
    blob_o = os.path.join(OUT, "builtin_blob.o")
    _run([cxx, "-c", "-o", blob_o, blob_s])
    _run([cxx, "-shared", "-o", lib_path(), *objects, blob_o,
          "-L", os.path.join(_cuda_home(), "lib64"), "-lcudart_static", "-lrt", "-ldl", "-lpthread"])

    with open(os.path.join(OUT, "stamp"), "w") as f:
        f.write(_tree_hash(_sources()))
    if verbose:
        print(f"built {lib_path()} ({os.path.getsize(lib_path())} bytes), "
              f"{cda1_path()} ({len(blob)} bytes, {len(exports)} exports)")
    return lib_path()


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--jobs", type=int, default=0)
    args = ap.parse_args(argv)
    build(force=args.force, jobs=args.jobs, verbose=True)


if __name__ == "__main__":
    sys.exit(main())
