"""Times iree_b200_ops_matmul_host (pinned host buffers in and out) on one shape across
pipeline block sizes, next to the raw PCIe copy rates of the box (one direction at a time
and both at once), so the end-to-end number can be read against its floor.

  python tools/host_pipeline_sweep.py [--shape 8192x8192x8192] [--blocks 1024,2048,4096]
"""

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="8192x8192x8192")
    ap.add_argument("--blocks", default="1024,2048,4096")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch
    from iree_b200 import hal

    M, N, K = [int(x) for x in args.shape.split("x")]
    dev = hal.Device(0)
    ops = hal.Ops(dev)
    a = torch.randn(M, K).to(torch.bfloat16).pin_memory()
    b = torch.randn(K, N).to(torch.bfloat16).pin_memory()
    c = torch.empty(M, N, dtype=torch.float32).pin_memory()
    rows = []

    # PCIe: 256 MiB each way, alone and together.
    nbytes = 256 << 20
    h_up = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h_down = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d_up = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    d_down = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    s_up, s_down = torch.cuda.Stream(), torch.cuda.Stream()

    def copies(up, down):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            if up:
                with torch.cuda.stream(s_up):
                    d_up.copy_(h_up, non_blocking=True)
            if down:
                with torch.cuda.stream(s_down):
                    h_down.copy_(d_down, non_blocking=True)
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / 5

    copies(True, True)
    pcie = {"h2d_gbs": nbytes / copies(True, False) / 1e9,
            "d2h_gbs": nbytes / copies(False, True) / 1e9,
            "both_gbs_each": nbytes / copies(True, True) / 1e9}
    rows.append({"pcie": pcie})
    print(json.dumps(rows[-1]), flush=True)

    flop = 2.0 * M * N * K
    for spec in args.blocks.split(","):
        blk = int(spec)
        ops.set_host_pipeline(16 << 20, blk, blk)
        for _ in range(3):
            ops.matmul_host("bf16", 1, M, N, K, a.data_ptr(), b.data_ptr(), c.data_ptr())
        t0 = time.perf_counter()
        for _ in range(args.iters):
            ops.matmul_host("bf16", 1, M, N, K, a.data_ptr(), b.data_ptr(), c.data_ptr())
        ms = (time.perf_counter() - t0) / args.iters * 1e3
        rows.append({"block": spec, "ms": ms, "tflops": flop / ms / 1e9})
        print(json.dumps(rows[-1]), flush=True)
    if args.out:
        with open(args.out, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
