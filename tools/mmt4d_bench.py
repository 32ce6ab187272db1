"""linalg.mmt4d on the device, route by route: the tile-direct tcgen05 export (one dispatch), the
relayout route (unpack lhs / rhs -> row-major tcgen05 GEMM -> pack result), the SIMT kernel, and
beside them the row-major NT GEMM of the same logical shape (the rate the direct route should
match). Device-resident operands, CUDA-event timing, rotating operand sets larger than L2; one
JSON line per route.

    python tools/mmt4d_bench.py [--size 8192] [--types bf16,i8] [--tiles 16x16] [--min-ms 200]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from iree_b200 import hal as H


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=8192)
    ap.add_argument("--types", default="bf16,i8")
    ap.add_argument("--tiles", default="16x16,8x32,128x128")
    ap.add_argument("--min-ms", type=float, default=200.0)
    ap.add_argument("--simt", action="store_true", help="also time the SIMT kernel (slow)")
    args = ap.parse_args()
    torch.cuda.init()
    stream = torch.cuda.Stream()
    dev = H.Device(0, mode=H.MODE_STREAM, external_stream=stream.cuda_stream)
    ops = H.Ops(dev)
    S = args.size
    for t in args.types.split(","):
        es = 1 if t == "i8" else 2
        K0 = 16 // es
        tdt = torch.int8 if t == "i8" else (torch.bfloat16 if t == "bf16" else torch.float16)
        sets = max(2, int(256e6 // (2 * S * S * es + S * S * 4)) + 1)
        # Operand bytes only matter for time, not values: random bytes, finite as bf16 / f16.
        def operand():
            x = torch.randint(-3, 4, (S * S,), device="cuda", dtype=torch.int32)
            return x.to(tdt)
        lhs = [operand() for _ in range(sets)]
        rhs = [operand() for _ in range(sets)]
        out = [torch.zeros(S * S, device="cuda", dtype=torch.int32) for _ in range(sets)]
        ws_t = None

        def ref(x):
            return dev.import_buffer(x.data_ptr(), x.numel() * x.element_size())

        routes = [("row-major NT GEMM", None, None)]
        for tile in args.tiles.split(","):
            M0, N0 = (int(v) for v in tile.split("x"))
            routes.append((f"mmt4d {M0}x{N0}x{K0} tile-direct", (M0, N0), 0))
            routes.append((f"mmt4d {M0}x{N0}x{K0} relayout route", (M0, N0), H.MATMUL_MMT4D_RELAYOUT))
            if args.simt:
                routes.append((f"mmt4d {M0}x{N0}x{K0} SIMT", (M0, N0), H.MATMUL_FORCE_GENERIC))
        for name, tile, flags in routes:
            cbs = []
            keep = []
            before = ops.dispatch_count
            for i in range(sets):
                cb = H.CommandBuffer(dev).begin()
                bl, br, bo = ref(lhs[i]), ref(rhs[i]), ref(out[i])
                keep += [bl, br, bo]
                if tile is None:
                    ops.record_matmul(cb, t, 1, S, S, S, bl.ref(), br.ref(), bo.ref(),
                                      flags=H.MATMUL_TRANSPOSE_RHS)
                else:
                    M0, N0 = tile
                    M1, N1, K1 = S // M0, S // N0, S // K0
                    if flags == H.MATMUL_MMT4D_RELAYOUT:
                        need = ops.mmt4d_workspace_size(t, M1, N1, K1, M0, N0, K0)
                        if ws_t is None or ws_t.numel() < need:
                            ws_t = torch.empty(need, device="cuda", dtype=torch.uint8)
                        bw = ref(ws_t)
                        keep.append(bw)
                        ops.record_mmt4d_tensor(cb, t, M1, N1, K1, M0, N0, K0, bl.ref(), br.ref(),
                                                bw.ref(), bo.ref(), flags=flags)
                    else:
                        ops.record_mmt4d(cb, t, M1, N1, K1, M0, N0, K0, bl.ref(), br.ref(), bo.ref(),
                                         flags=flags)
                cb.end()
                cbs.append(cb)
            per_call = (ops.dispatch_count - before) // sets
            with torch.cuda.stream(stream):
                for i in range(3):
                    dev.execute(cbs[i % sets])
            torch.cuda.synchronize()
            def timed(n):   # CUDA events on the stream the dispatches are launched on
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                e0.record(stream)
                for i in range(n):
                    dev.execute(cbs[i % sets])
                e1.record(stream)
                torch.cuda.synchronize()
                return e0.elapsed_time(e1)
            probe = timed(4) / 4
            n = max(8, int(args.min_ms / max(probe, 1e-3)))
            ms = timed(n) / n
            print(json.dumps({"route": name, "type": t, "size": S, "dispatches_per_call": per_call,
                              "ms": round(ms, 4), "tera_per_s": round(2.0 * S ** 3 / ms / 1e9, 1),
                              "calls_timed": n, "operand_sets": sets,
                              "timing": "CUDA events on the launching stream"}),
                  flush=True)


if __name__ == "__main__":
    main()
