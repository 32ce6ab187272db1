#!/usr/bin/env python
"""Throughput of our exports THROUGH THE REFERENCE HAL (GPU box): the stock tool
`iree-benchmark-executable` (tools/iree-benchmark-executable-main.c:191-300) from the reference
runtime built by tools/build_ref_runtime.sh, loading iree_b200_kernels.cda1 with the reference's
own loader and launching through its stream command buffer.

  stock/  unmodified driver  -> recipe-free exports: the `*_dtm` tcgen05 GEMMs (tensor maps built
          on the device; bf16 / i8 8192^3), decode-cluster attention, generic matmul
  tma/    driver + integration/cuda_driver_tma.patch -> the tcgen05 exports (bf16 / i8 8192^3,
          prefill attention)

Prints one JSON line per case (ms per dispatch as the tool reports it — host wall clock per
batch of dispatches, submission overhead included — and the derived TFLOP/s or GB/s) and
keeps the tool's raw output in gpurun_out/ref_runtime_<case>.log.
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
from iree_b200 import hal as H  # noqa: E402

RT = os.path.join(ROOT, "oracle", "_ref", "iree-rt")
CDA1 = os.path.join(ROOT, "iree_b200", "_build", "iree_b200_kernels.cda1")
OUT = os.path.join(ROOT, "gpurun_out")


def run_tool(flavour, case, ordinal, grid, constants, bindings, batch_size):
    cmd = [os.path.join(RT, flavour, "iree-benchmark-executable"), "--device=cuda",
           "--executable_format=PTXE", f"--executable_file={CDA1}", f"--entry_point={ordinal}",
           "--workgroup_count=%d,%d,%d" % tuple(grid), f"--batch_size={batch_size}",
           "--benchmark_min_time=1.0s"]
    cmd += [f"--constant={int(c)}" for c in constants] + [f"--binding={b}" for b in bindings]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    os.makedirs(OUT, exist_ok=True)
    with open(os.path.join(OUT, f"ref_runtime_{case}.log"), "w") as f:
        f.write("$ " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        return None, (r.stderr or r.stdout)[-400:]
    # Google-benchmark row: "BM_dispatch_XxYxZ/process_time/real_time   123 ns   120 ns  N items_per_second=..."
    m = re.search(r"dispatch_\S+\s+([0-9.]+) (ns|us|ms)", r.stdout)
    if not m:
        return None, r.stdout[-400:]
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0}[m.group(2)]
    return float(m.group(1)) * scale, None


def main():
    ops = H.Ops.offline(148)
    rng = np.random.Generator(np.random.PCG64(0xB200))
    tmp = tempfile.mkdtemp(prefix="refrt_")

    def bf16_file(name, shape):
        x = np.clip(rng.standard_normal(shape, dtype=np.float32), -4, 4)
        bits = (x.view(np.uint32) >> 16).astype(np.uint16)
        p = os.path.join(tmp, name)
        bits.tofile(p)
        return p

    def f16_file(name, shape):
        p = os.path.join(tmp, name)
        np.clip(rng.standard_normal(shape, dtype=np.float32), -4, 4).astype(np.float16).tofile(p)
        return p

    def i8_file(name, shape):
        p = os.path.join(tmp, name)
        rng.integers(-128, 128, shape, dtype=np.int8).tofile(p)
        return p

    cases = []
    M = N = K = 8192
    sel = ops.select_matmul("bf16", 1, M, N, K)
    cases.append(("tma", "matmul_bf16_8192", sel["export_ordinal"], sel["workgroup_count"],
                  sel["constants"],
                  [f"{M}x{K}xbf16=@{bf16_file('a.bin', (M, K))}", f"{K}x{N}xbf16=@{bf16_file('b.bin', (K, N))}",
                   f"{M}x{N}xf32"], 20, 2.0 * M * N * K, None))
    sel = ops.select_matmul("i8", 1, M, N, K)
    cases.append(("tma", "matmul_i8_8192", sel["export_ordinal"], sel["workgroup_count"],
                  sel["constants"],
                  [f"{M}x{K}xi8=@{i8_file('ai.bin', (M, K))}", f"{K}x{N}xi8=@{i8_file('bi.bin', (K, N))}",
                   f"{M}x{N}xi32"], 20, 2.0 * M * N * K, None))
    # The same two GEMMs through the UNMODIFIED driver: device-built tensor maps.
    for t, acc in (("bf16", "f32"), ("i8", "i32")):
        sel = ops.select_matmul(t, 1, M, N, K, flags=H.MATMUL_DEVICE_MAPS)
        a, b = ("a.bin", "b.bin") if t == "bf16" else ("ai.bin", "bi.bin")
        cases.append(("stock", f"matmul_{t}_8192_device_maps", sel["export_ordinal"],
                      sel["workgroup_count"], sel["constants"],
                      [f"{M}x{K}x{t}=@{os.path.join(tmp, a)}", f"{K}x{N}x{t}=@{os.path.join(tmp, b)}",
                       f"{M}x{N}x{acc}"], 20, 2.0 * M * N * K, None))
    B, Mq, K2, D = 32, 4096, 4096, 128
    sel = ops.select_attention("f16", B, Mq, D, K2, D, 1.0 / D ** 0.5, flags=H.ATTENTION_NO_TAIL_SPLIT)
    d = sel["dispatches"][0]
    cases.append(("tma", "attention_prefill_f16", d["export_ordinal"], d["workgroup_count"],
                  d["constants"],
                  [f"{B}x{Mq}x{D}xf16=@{f16_file('q.bin', (B, Mq, D))}",
                   f"{B}x{K2}x{D}xf16=@{f16_file('k.bin', (B, K2, D))}",
                   f"{B}x{K2}x{D}xf16=@{f16_file('v.bin', (B, K2, D))}", f"{B}x{Mq}x{D}xf16"],
                  20, 4.0 * B * Mq * K2 * D, None))
    sel = ops.select_attention("f16", B, Mq, D, K2, D, 1.0 / D ** 0.5, flags=H.ATTENTION_DEVICE_MAPS)
    d = sel["dispatches"][0]
    cases.append(("stock", "attention_prefill_f16_device_maps", d["export_ordinal"], d["workgroup_count"],
                  d["constants"],
                  [f"{B}x{Mq}x{D}xf16=@{os.path.join(tmp, 'q.bin')}",
                   f"{B}x{K2}x{D}xf16=@{os.path.join(tmp, 'k.bin')}",
                   f"{B}x{K2}x{D}xf16=@{os.path.join(tmp, 'v.bin')}", f"{B}x{Mq}x{D}xf16"],
                  20, 4.0 * B * Mq * K2 * D, None))
    B, Mq, K2 = 32, 1, 8192
    sel = ops.select_attention("f16", B, Mq, D, K2, D, 1.0 / D ** 0.5)
    d = sel["dispatches"][0]
    cases.append(("stock", "attention_decode_cluster_f16", d["export_ordinal"], d["workgroup_count"],
                  d["constants"],
                  [f"{B}x{Mq}x{D}xf16=@{f16_file('qd.bin', (B, Mq, D))}",
                   f"{B}x{K2}x{D}xf16=@{f16_file('kd.bin', (B, K2, D))}",
                   f"{B}x{K2}x{D}xf16=@{f16_file('vd.bin', (B, K2, D))}", f"{B}x{Mq}x{D}xf16"],
                  64, None, float(B) * 2 * K2 * D * 2))
    Mg, Kg, Ng = 1000, 1000, 1000
    sel = ops.select_matmul("bf16", 1, Mg, Ng, Kg, flags=H.MATMUL_FORCE_GENERIC)
    cases.append(("stock", "matmul_generic_bf16_1000", sel["export_ordinal"], sel["workgroup_count"],
                  sel["constants"],
                  [f"{Mg}x{Kg}xbf16=@{bf16_file('ag.bin', (Mg, Kg))}",
                   f"{Kg}x{Ng}xbf16=@{bf16_file('bg.bin', (Kg, Ng))}", f"{Mg}x{Ng}xf32"],
                  20, 2.0 * Mg * Ng * Kg, None))

    for flavour, case, ordinal, grid, constants, bindings, batch, flop, nbytes in cases:
        name = ops.executable.export_info(ordinal)["kernel_name"]
        ms, err = run_tool(flavour, case, ordinal, grid, constants, bindings, batch)
        line = {"through": f"reference runtime ({flavour} CUDA HAL driver), iree-benchmark-executable "
                           f"--batch_size={batch}", "case": case, "export": name,
                "workgroup_count": list(grid)}
        if ms is None:
            line["error"] = err
        else:
            line["ms_per_dispatch"] = ms
            if flop:
                line["tflops"] = flop / (ms * 1e-3) / 1e12
            if nbytes:
                line["gbs"] = nbytes / (ms * 1e-3) / 1e9
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
