"""Times the kernels that have not had a GPU run yet (the on-demand second executable: the
mmt4d SIMT kernels, the MXFP4 SIMT kernel and the MXFP4 bf16-expansion + tcgen05 route) after
checking each result against the oracle, so the first GPU
call of the next round yields parity AND a number. CUDA events on the dispatch stream, after
warm-up; one JSON line per kernel. Not part of bench.py: nothing here is a headline metric.

  python tools/pending_bench.py [--size 2048] [--iters 10] [--out profiles/pending_r2.json]
"""

import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _time(torch, stream, dev, cb, iters):
    with torch.cuda.stream(stream):
        for _ in range(3):
            dev.execute(cb)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(iters):
            dev.execute(cb)
        e1.record(stream)
        e1.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=2048, help="M = N = K (logical)")
    ap.add_argument("--check-size", type=int, default=256, help="M = N = K of the oracle check")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    import torch
    import oracle
    from iree_b200 import hal

    stream = torch.cuda.Stream()
    dev = hal.Device(0, mode=hal.MODE_STREAM, external_stream=stream.cuda_stream)
    ops = hal.Ops(dev)
    rows = []

    def upload(a):
        b = dev.allocate(max(a.nbytes, 16))
        dev.h2d(a.ctypes.data, b, a.nbytes)
        return b

    # ---- linalg.mmt4d, the AVX-512-BF16 tile shape and an MMA-shaped one -------------------
    for t, (M0, N0, K0) in (("bf16", (16, 16, 2)), ("bf16", (16, 16, 16)), ("i8", (16, 16, 2))):
        for size, check in ((args.check_size, True), (args.size, False)):
            lhs = oracle.generate_random_matrix(t, (size, size), 1)
            rhs = oracle.generate_random_matrix(t, (size, size), 2)
            l4 = oracle.pack(lhs, M0, K0)
            r4 = oracle.pack(np.ascontiguousarray(rhs.T), N0, K0)
            M1, K1, N1 = l4.shape[0], l4.shape[1], r4.shape[0]
            out = np.zeros((M1, N1, M0, N0), dtype=np.int32 if t == "i8" else np.float32)
            bufs = [upload(l4), upload(r4), upload(out)]
            need = ops.mmt4d_workspace_size(t, M1, N1, K1, M0, N0, K0)
            for route in ("simt", "tensor"):
                if route == "tensor" and not need:
                    continue
                cb = hal.CommandBuffer(dev).begin()
                if route == "tensor":
                    ws = dev.allocate(need)
                    ops.record_mmt4d_tensor(cb, t, M1, N1, K1, M0, N0, K0, bufs[0].ref(),
                                            bufs[1].ref(), ws.ref(), bufs[2].ref())
                else:
                    ops.record_mmt4d(cb, t, M1, N1, K1, M0, N0, K0, bufs[0].ref(), bufs[1].ref(),
                                     bufs[2].ref())
                cb.end()
                row = {"kernel": f"mmt4d_{route}_{t}", "tiles": [M0, N0, K0], "size": size}
                if check:
                    dev.execute(cb)
                    dev.wait_idle()
                    dev.d2h(bufs[2], out.ctypes.data, out.nbytes)
                    row["bit_exact_vs_oracle"] = bool(np.array_equal(out, oracle.mmt4d(t, l4, r4)))
                else:
                    ms = _time(torch, stream, dev, cb, args.iters)
                    row.update(ms=ms, tflops=2.0 * size ** 3 / ms / 1e9)
                rows.append(row)
                print(json.dumps(row), flush=True)

    # ---- MXFP4: the SIMT dispatch and the bf16-expansion + tcgen05 route -----------------------
    def _mx_row(route, size, check, cb, bufs, out, lhs, ls, rhs, rs):
        name = "matmul_mxfp4_" + route
        if check:
            dev.execute(cb)
            dev.wait_idle()
            dev.d2h(bufs[4], out.ctypes.data, out.nbytes)
            ok = bool(np.array_equal(out, oracle.matmul_mxfp4(lhs, ls, rhs, rs)))
            rows.append({"kernel": name, "size": size, "bit_exact_vs_oracle": ok})
        else:
            ms = _time(torch, stream, dev, cb, args.iters)
            rows.append({"kernel": name, "size": size, "ms": ms,
                         "tflops": 2.0 * size ** 3 / ms / 1e9})
        print(json.dumps(rows[-1]), flush=True)

    for size, check in ((args.check_size, True), (args.size, False)):
        lhs = oracle.generate_random_matrix("u8", (size, size // 2), 1)
        ls = oracle.generate_random_matrix("f8e8m0fnu", (size, size // 32), 2)
        rhs = oracle.generate_random_matrix("u8", (size, size // 2), 3)
        rs = oracle.generate_random_matrix("f8e8m0fnu", (size, size // 32), 4)
        out = np.zeros((size, size), dtype=np.float32)
        bufs = [upload(a) for a in (lhs, ls, rhs, rs, out)]
        for route in ("simt", "tensor"):
            need = ops.matmul_mxfp4_workspace_size(size, size, size)
            if route == "tensor" and not need:
                continue
            if route == "simt" and not check and size > 4096:
                continue                        # the SIMT kernel is a correctness path
            cb = hal.CommandBuffer(dev).begin()
            refs = [b.ref() for b in bufs]
            if route == "tensor":
                ws = dev.allocate(need)
                ops.record_matmul_mxfp4_tensor(cb, size, size, size, *refs[:4], ws.ref(), refs[4])
            else:
                ops.record_matmul_mxfp4(cb, size, size, size, *refs)
            cb.end()
            _mx_row(route, size, check, cb, bufs, out, lhs, ls, rhs, rs)

    if args.out:
        with open(args.out, "w") as f:
            json.dump(rows, f, indent=1)
    return 0 if all(r.get("bit_exact_vs_oracle", True) for r in rows) else 1


if __name__ == "__main__":
    sys.exit(main())
