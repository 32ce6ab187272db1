"""GPU bring-up: runs every matmul kernel family through the C-ABI host call, one
subprocess per case (a trapped kernel poisons its CUDA context), and compares with an
exact numpy product of the reference generator's small-integer inputs.

  python tools/gpu_bringup.py            # all cases
  python tools/gpu_bringup.py --case i8:2cta:nn:512x512x512   # one case, in-process
"""

import argparse
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

FAMILY_FLAGS = {"generic": 1 << 8, "1cta": 1 << 9, "2cta": 1 << 10, "auto": 0}


def run_case(spec: str) -> int:
    import oracle
    from iree_b200 import hal

    et, family, form, shape = spec.split(":")
    dims = [int(x) for x in shape.split("x")]
    batch = 1
    if len(dims) == 4:
        batch, M, N, K = dims
    else:
        M, N, K = dims
    nt = form == "nt"
    lhs = oracle.generate_random_matrix(et, (batch, M, K), 1)
    rhs = oracle.generate_random_matrix(et, (batch, N, K) if nt else (batch, K, N), 2)
    a = oracle.to_f32(et, lhs) if et != "i8" else lhs.astype(np.float32)
    b = oracle.to_f32(et, rhs) if et != "i8" else rhs.astype(np.float32)
    if nt:
        b = np.swapaxes(b, 1, 2)
    want = np.matmul(a, b)  # exact: |values| <= 4, K <= 8192 -> < 2^24
    out = np.full((batch, M, N), -7, dtype=np.int32 if et == "i8" else np.float32)
    dev = hal.Device(0)
    ops = hal.Ops(dev)
    flags = FAMILY_FLAGS[family] | (hal.MATMUL_TRANSPOSE_RHS if nt else 0)
    t0 = time.time()
    ops.matmul_host(et, batch, M, N, K, lhs.ctypes.data, rhs.ctypes.data, out.ctypes.data, flags)
    dt = time.time() - t0
    got = out.astype(np.float32)
    bad = np.argwhere(got != want)
    if len(bad):
        print(f"FAIL {spec}: {len(bad)}/{got.size} mismatches; first {bad[:4].tolist()} "
              f"got {[float(got[tuple(i)]) for i in bad[:4]]} "
              f"want {[float(want[tuple(i)]) for i in bad[:4]]}")
        rows = np.unique(bad[:, 1])
        cols = np.unique(bad[:, 2])
        print(f"   bad rows: n={len(rows)} [{rows[:8].tolist()}..{rows[-4:].tolist()}] "
              f"bad cols: n={len(cols)} [{cols[:8].tolist()}..{cols[-4:].tolist()}]")
        return 1
    print(f"PASS {spec} ({dt*1e3:.1f} ms host call)")
    return 0


CASES = [
    "bf16:generic:nn:64x64x64",
    "bf16:generic:nn:457x330x512",
    "i8:generic:nt:100x70x33",
    "bf16:1cta:nt:128x256x64",
    "bf16:1cta:nt:128x256x512",
    "bf16:1cta:nt:512x512x512",
    "bf16:1cta:nn:128x256x64",
    "bf16:1cta:nn:512x512x512",
    "f16:1cta:nn:512x128x512",
    "i8:1cta:nt:512x512x512",
    "i8:1cta:nn:512x512x512",
    "bf16:2cta:nt:256x256x64",
    "bf16:2cta:nt:512x512x512",
    "bf16:2cta:nn:512x512x512",
    "i8:2cta:nn:512x512x512",
    "i8:2cta:nt:512x512x512",
    "bf16:1cta:nn:457x328x512",
    "bf16:2cta:nn:1000x520x264",
    "bf16:2cta:nn:4x512x512x256",
    "bf16:auto:nn:4096x4096x4096",
    "i8:auto:nn:4096x4096x4096",
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case")
    ap.add_argument("--timeout", type=int, default=120)
    args = ap.parse_args()
    if args.case:
        sys.exit(run_case(args.case))
    failed = 0
    for spec in CASES:
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--case", spec],
                               stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                               timeout=args.timeout)
            out = r.stdout.strip().splitlines()
            print("\n".join(out[-6:]) if r.returncode else out[-1], flush=True)
            failed += r.returncode != 0
        except subprocess.TimeoutExpired:
            print(f"TIMEOUT {spec}", flush=True)
            failed += 1
    print(f"{len(CASES) - failed}/{len(CASES)} passed")
    sys.exit(1 if failed else 0)


if __name__ == "__main__":
    main()
