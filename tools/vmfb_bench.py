#!/usr/bin/env python
"""Our kernels as an IREE module under the reference's stock `iree-benchmark-module` (GPU box).

iree_b200/_build/iree_b200_module.vmfb (iree_b200/vmasm.py + the reference's iree-as-module)
is loaded by the unmodified runtime: VM, HAL module, CUDA HAL driver. Each benchmark iteration
is one call of the module function = record a one-shot command buffer with the dispatch,
queue_execute, wait on the fence — the whole host path of a compiled IREE program, not just
the kernel. `<name>_into` takes the result storage from the caller (destination passing);
`<name>` also allocates the result per call (hal.device.queue.alloca) and so includes the
stock driver's pool allocate / free of the result (256 MiB for the 8192^2 f32 result).
Operands are random files (constant operands toggle few bits, draw less power and read
20-40 % high). One JSON line per function (ms per call as the tool reports it, wall clock, and
the derived rate); raw tool output in gpurun_out/vmfb_<function>.log.
"""
import json
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from iree_b200 import vmasm  # noqa: E402

TOOL = os.path.join(ROOT, "oracle", "_ref", "iree-rt", "stock", "iree-benchmark-module")
OUT = os.path.join(ROOT, "gpurun_out")


def operand_file(tmp, rng, shape, t):
    """Random operand of element type |t| as a raw file iree-benchmark-module reads."""
    path = os.path.join(tmp, "%s_%s.bin" % ("x".join(str(d) for d in shape), t))
    if not os.path.exists(path):
        if t == "i8":
            rng.integers(-128, 128, shape, dtype=np.int8).tofile(path)
        else:
            x = np.clip(rng.standard_normal(shape, dtype=np.float32), -4, 4)
            if t == "bf16":
                (x.view(np.uint32) >> 16).astype(np.uint16).tofile(path)
            elif t == "f16":
                x.astype(np.float16).tofile(path)
            else:
                x.tofile(path)
    return path


def main():
    os.makedirs(OUT, exist_ok=True)
    tmp = tempfile.mkdtemp(prefix="vmfb_")
    rng = np.random.Generator(np.random.PCG64(0xB200))
    work = {"matmul_bf16": 2.0 * 8192 ** 3, "matmul_i8": 2.0 * 8192 ** 3,
            "batch_matmul_bf16": 2.0 * 64 * 2048 ** 3, "linear_bf16": 2.0 * 8192 ** 3,
            "attention_prefill_f16": 4.0 * 32 * 4096 * 4096 * 128,
            "attention_prefill_causal_f16": 4.0 * 32 * 128 * (4096 * 4097 / 2),
            "attention_decode_f16": None}
    for w, into in [(w, into) for w in vmasm.default_workloads() for into in (True, False)]:
        if w.name not in work:
            continue
        function = w.name + ("_into" if into else "")
        cmd = [TOOL, "--device=cuda", f"--module={vmasm.MODULE_VMFB}", f"--function={function}",
               "--benchmark_min_time=1.0s"]
        cmd += ["--input=%sx%s=@%s" % ("x".join(str(d) for d in s), t, operand_file(tmp, rng, s, t))
                for s, t in w.inputs]
        if into:
            cmd.append("--input=%sx%s=0" % ("x".join(str(d) for d in w.output[0]), w.output[1]))
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
        with open(os.path.join(OUT, f"vmfb_{function}.log"), "w") as f:
            f.write("$ " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        line = {"through": "stock iree-benchmark-module --device=cuda (unmodified reference runtime), "
                           "module assembled by the reference's iree-as-module",
                "function": function, "export": w.kernel, "workgroup_count": w.grid}
        m = re.search(r"BM_\S+\s+([0-9.]+) (ns|us|ms)", r.stdout)
        if r.returncode != 0 or not m:
            line["error"] = (r.stderr or r.stdout)[-400:]
        else:
            ms = float(m.group(1)) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0}[m.group(2)]
            line["ms_per_call"] = ms
            if work[w.name]:
                line["tflops"] = work[w.name] / (ms * 1e-3) / 1e12
            else:   # decode: K and V streamed once
                line["gbs"] = 32 * 2 * 8192 * 128 * 2 / (ms * 1e-3) / 1e9
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
