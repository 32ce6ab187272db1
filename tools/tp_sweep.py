#!/usr/bin/env python
"""One multi-GPU box visit for the tensor-parallel path (run bare, not under torchrun):
 1. sweeps the NVLS all-reduce kernel's tuning (blocks, threads, unroll, software pipelining)
    with tools/symmetric_check.py --sweep (parity of every variant + CUDA-event timing);
 2. runs bench.py --workload tp_decoder_layer with the best variants that fit the SMs kept
    free of GEMM CTAs, and with NCCL, and prints every JSON line.
Output: gpurun_out/tp_sweep_<N>.jsonl
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def torchrun(n, port, script, *args, env=None, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, script), *args]
    e = dict(os.environ, IREE_B200_COLLECTIVE_WATCHDOG_MS="30000")
    e.update(env or {})
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=e, cwd=ROOT)
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    if r.returncode != 0:
        print(f"[{script} {' '.join(args)}] exit {r.returncode}\n{r.stdout[-800:]}\n{r.stderr[-1500:]}",
              file=sys.stderr)
    return lines


def main():
    import torch
    n = torch.cuda.device_count()
    out = open(os.path.join(ROOT, "gpurun_out", f"tp_sweep_{n}.jsonl"), "w")

    def emit(obj):
        print(json.dumps(obj), flush=True)
        out.write(json.dumps(obj) + "\n")
        out.flush()

    variants = []
    for blocks in (8, 16, 24, 32, 64):
        for threads, unroll, pipe in ((1024, 4, 0), (1024, 2, 1), (512, 4, 0)):
            variants.append((blocks, threads, unroll, pipe))
    lines = torchrun(n, 29581, "tools/symmetric_check.py", "--mib", "64", "--iters", "20", "--sweep",
                     ";".join(",".join(map(str, v)) for v in variants), timeout=900)
    rows = lines[-1]["sweep"] if lines else []
    for row in rows:
        emit({"all_reduce_sweep": row, "world": n})
    port = 29590
    configs = [("nccl", 16, 16, None), ("symmetric", 16, 16, None)]
    ok = [r for r in rows if r["exact"]]
    for cap in (8, 24, 32):
        fit = sorted((r for r in ok if r["blocks"] <= cap), key=lambda r: r["ms_16mib"])
        if fit:
            b = fit[0]
            configs.append(("symmetric", b["blocks"], b["blocks"],
                            f"{b['threads']},{b['unroll']},{b['pipelined']}"))
    seen = set()
    for collective, reserved, blocks, tuning in configs:
        key = (collective, reserved, blocks, tuning)
        if key in seen:
            continue
        seen.add(key)
        port += 1
        env = {"IREE_B200_ALL_REDUCE_TUNING": tuning} if tuning else {}
        for extra in ([], ["--tp-schedule", "staged"], ["--tp-schedule", "fused"]) \
                if collective == "symmetric" and tuning is None else ([],):
            port += 1
            lines = torchrun(n, port, "bench.py", "--gpus", str(n), "--workload", "tp_decoder_layer",
                             "--collective", collective, "--reserved-sms", str(reserved),
                             "--ar-blocks", str(blocks), "--steps", "100", "--warmup", "5",
                             "--e2e-steps", "2", *extra, env=env)
            for l in lines:
                emit({"tp_bench": {"collective": collective, "reserved_sms": reserved, "ar_blocks": blocks,
                                   "tuning": tuning, "schedule": extra[-1] if extra else "dataflow", "value": l["value"],
                                   "ms_per_step": l["ms_per_step"], "e2e": l["e2e"]["value"],
                                   "workload": l["config"]["workload"][-260:]}, "world": n})


if __name__ == "__main__":
    main()
