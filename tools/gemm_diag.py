"""GEMM pipeline diagnostics on a B200: times each tcgen05 matmul export on one shape with
the kernel's diagnostic flag bits (gemm_sm100.cuh: 0x10 no store, 0x20 no load, 0x40 no
MMA, 0x80 L2 eviction hints on the TMA loads, bits 8-15 raster group height) so the slowest pipeline leg can be read off directly. Results under a diagnostic bit
are garbage by construction; flag 0 is the real kernel. Also times torch.matmul (cuBLAS)
on the same operands as the live roofline denominator.

  python tools/gemm_diag.py [--shape 8192x8192x8192] [--types bf16,i8] [--iters 10]
"""

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="8192x8192x8192")
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--types", default="bf16,i8")
    ap.add_argument("--families", default="2cta,1cta")
    ap.add_argument("--forms", default="nn,nt")
    ap.add_argument("--diag", default="0,0x10,0x20,0x40,0x30,0x50")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    import torch
    from iree_b200 import hal

    M, N, K = [int(x) for x in args.shape.split("x")]
    batch = args.batch
    stream = torch.cuda.Stream()
    dev = hal.Device(0, mode=hal.MODE_STREAM, external_stream=stream.cuda_stream)
    ops = hal.Ops(dev)
    exe = ops.executable
    fam_flag = {"2cta": hal.MATMUL_FORCE_2CTA, "1cta": hal.MATMUL_FORCE_1CTA}
    rows = []
    for et in args.types.split(","):
        if et == "i8":
            a = torch.randint(-128, 128, (batch, M, K), dtype=torch.int8, device="cuda")
            b = torch.randint(-128, 128, (batch, K, N), dtype=torch.int8, device="cuda")
            c = torch.empty((batch, M, N), dtype=torch.int32, device="cuda")
        else:
            dt = torch.bfloat16 if et == "bf16" else torch.float16
            a = torch.randn((batch, M, K), device="cuda").to(dt)
            b = torch.randn((batch, K, N), device="cuda").to(dt)
            c = torch.empty((batch, M, N), dtype=torch.float32, device="cuda")
        bufs = [dev.import_buffer(t.data_ptr(), t.numel() * t.element_size()) for t in (a, b, c)]
        flop = 2.0 * batch * M * N * K
        if et != "i8":
            with torch.cuda.stream(stream):
                for _ in range(3):
                    torch.matmul(a, b)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(args.iters):
                    torch.matmul(a, b)
                e1.record(stream)
                e1.synchronize()
            ms = e0.elapsed_time(e1) / args.iters
            rows.append({"type": et, "kernel": "torch.matmul(cuBLAS, bf16 out)", "ms": ms,
                         "tflops": flop / ms / 1e9})
            print(json.dumps(rows[-1]), flush=True)
        for form in args.forms.split(","):
            for fam in args.families.split(","):
                f = fam_flag[fam] | (hal.MATMUL_TRANSPOSE_RHS if form == "nt" else 0)
                sel = ops.select_matmul(et, batch, M, N, K, flags=f)
                name = exe.export_info(sel["export_ordinal"])["kernel_name"]
                for d in args.diag.split(","):
                    diag = int(d, 0)
                    consts = list(sel["constants"])
                    consts[4] |= diag
                    cb = hal.CommandBuffer(dev).begin()
                    cb.dispatch(exe, sel["export_ordinal"], sel["workgroup_count"], consts,
                                [x.ref() for x in bufs])
                    cb.end()
                    with torch.cuda.stream(stream):
                        for _ in range(3):
                            dev.execute(cb)
                        e0 = torch.cuda.Event(enable_timing=True)
                        e1 = torch.cuda.Event(enable_timing=True)
                        e0.record(stream)
                        for _ in range(args.iters):
                            dev.execute(cb)
                        e1.record(stream)
                        e1.synchronize()
                    ms = e0.elapsed_time(e1) / args.iters
                    rows.append({"type": et, "kernel": name, "grid": sel["workgroup_count"][0],
                                 "diag": hex(diag), "ms": ms, "tflops": flop / ms / 1e9})
                    if (diag & 0x70) == 0:
                        # 0x80 (L2 hints) and bits 8-15 (raster group height) keep the result.
                        if diag == 0:
                            want = c.clone()
                        elif "want" in locals():
                            rows[-1]["same_result"] = bool(torch.equal(c, want))
                    print(json.dumps(rows[-1]), flush=True)
    if args.out:
        with open(args.out, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
