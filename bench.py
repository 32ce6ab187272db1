#!/usr/bin/env python
"""bench.py — the dense-contraction dispatch path on B200, measured per BASELINE.json.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--workload matmul_bf16|matmul_i8|batch_matmul|attention_prefill|...]

Own arm. One "step" = one pass of the workload's dispatch (one command-buffer execution
through the thin launch layer, include/iree_b200_hal.h) over synthetic inputs already
resident in HBM. At N=1 the default workload is BASELINE.json configs[1]'s bf16 half:
linalg.matmul bf16 x bf16 -> f32, 8192^3. At N>1 every rank runs the same dispatch on its
own batch slice (batch-sharded linalg.batch_matmul, batch = N, no data-path collective):
weak scaling. `value` = aggregate TFLOP/s over all ranks from CUDA-event time (max over
ranks); `e2e` = the same metric through iree_b200_ops_matmul_host with pinned HOST
buffers (H2D of both operands and D2H of the result inside the timed region).

Reference arm (`--impl reference`): the reference's CPU path for the same op on the box's
host cores — the oracle port (oracle/oracle.c; kind "port") — on a bounded sample.

Prints ONE JSON line on rank 0.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "matmul/attention TFLOP/s and % of B200 dense BF16 peak at 1/2/4/8 GPUs"

# name -> (op, element type, batch, M, N, K)   [matmul family]
MATMUL_WORKLOADS = {
    "matmul_bf16": ("matmul", "bf16", 1, 8192, 8192, 8192),
    "matmul_i8": ("matmul", "i8", 1, 8192, 8192, 8192),
    "matmul_f16": ("matmul", "f16", 1, 8192, 8192, 8192),
    "matmul_bf16_4096": ("matmul", "bf16", 1, 4096, 4096, 4096),
    "batch_matmul": ("batch_matmul", "bf16", 64, 2048, 2048, 2048),
}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"source": "measured", "bf16_tflops": p["bf16_tflops"],
                "bf16_tflops_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                "hbm_gbs": p["hbm_gbs"]}
    # /opt/skills/guides/B200_PROFILING.md fallback figures.
    return {"source": "fallback", "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0,
            "hbm_gbs": 6650.0}


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.proc = None
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) >= 9:
                self.rows.append(parts)

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                smax.append(float(r[2]))
                power.append(float(r[3]))
            except ValueError:
                continue
            for name, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_env():
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)),
            int(os.environ.get("LOCAL_RANK", 0)))


def max_over_ranks(value: float, world: int) -> float:
    if world == 1:
        return value
    import torch
    import torch.distributed as dist
    t = torch.tensor([value], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world: int):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def algorithmic(op, et, batch, M, N, K):
    """(flop, bytes) per dispatch: SURVEY.md §8(d) — read-once / write-once operands."""
    es = 1 if et == "i8" else 2
    return 2.0 * batch * M * N * K, float(batch) * (M * K * es + K * N * es + M * N * 4)


def synth_matmul_inputs(torch, et, batch, M, N, K, seed):
    g = torch.Generator(device="cuda")
    g.manual_seed(0xB200 + seed)
    if et == "i8":
        a = torch.randint(-128, 128, (batch, M, K), dtype=torch.int8, device="cuda", generator=g)
        b = torch.randint(-128, 128, (batch, K, N), dtype=torch.int8, device="cuda", generator=g)
        c = torch.empty((batch, M, N), dtype=torch.int32, device="cuda")
    else:
        dt = torch.bfloat16 if et == "bf16" else torch.float16
        a = torch.randn((batch, M, K), device="cuda", generator=g).clamp_(-4, 4).to(dt)
        b = torch.randn((batch, K, N), device="cuda", generator=g).clamp_(-4, 4).to(dt)
        c = torch.empty((batch, M, N), dtype=torch.float32, device="cuda")
    return a, b, c


class MatmulStep:
    """One recorded command buffer holding the workload's dispatch; step() replays it."""

    def __init__(self, torch, hal, dev, ops, et, batch, M, N, K, seed):
        self.shape = (et, batch, M, N, K)
        self.tensors = synth_matmul_inputs(torch, et, batch, M, N, K, seed)
        self.buffers = [dev.import_buffer(t.data_ptr(), t.numel() * t.element_size())
                        for t in self.tensors]
        before = ops.dispatch_count
        self.cb = hal.CommandBuffer(dev).begin()
        ops.record_matmul(self.cb, et, batch, M, N, K, self.buffers[0].ref(),
                          self.buffers[1].ref(), self.buffers[2].ref())
        self.cb.end()
        self.launches_per_step = ops.dispatch_count - before
        self.dev = dev

    def step(self):
        self.dev.execute(self.cb)


def run_ours(args):
    import torch
    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from iree_b200 import hal

    op, et, batch, M, N, K = MATMUL_WORKLOADS[args.workload]
    stream = torch.cuda.Stream()
    dev = hal.Device(local, mode=hal.MODE_STREAM, external_stream=stream.cuda_stream)
    ops = hal.Ops(dev)
    work = MatmulStep(torch, hal, dev, ops, et, batch, M, N, K, seed=rank)
    flop, nbytes = algorithmic(op, et, batch, M, N, K)

    # ---- device-resident timing (CUDA events on the dispatch stream) ----
    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            work.step()
        barrier(world)
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(stream)
        for _ in range(args.steps):
            work.step()
        ev1.record(stream)
        barrier(world)
        clocks = sampler.stop() if rank == 0 else None
    ms_total = max_over_ranks(ev0.elapsed_time(ev1), world)
    ms_per_step = ms_total / args.steps
    value = flop * world / (ms_per_step * 1e-3) / 1e12
    launches = work.launches_per_step * args.steps

    # ---- end to end through the host-buffer C-ABI call ----
    es = 1 if et == "i8" else 2
    host = [torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in work.tensors]
    host[0].copy_(work.tensors[0])
    host[1].copy_(work.tensors[1])
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    for _ in range(2):
        ops.matmul_host(et, batch, M, N, K, host[0].data_ptr(), host[1].data_ptr(),
                        host[2].data_ptr())
    barrier(world)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ops.matmul_host(et, batch, M, N, K, host[0].data_ptr(), host[1].data_ptr(),
                        host[2].data_ptr())
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3, world) / e2e_steps
    barrier(world)
    e2e = {"value": flop * world / (e2e_ms * 1e-3) / 1e12, "unit": "TFLOP/s",
           "ms_per_step": e2e_ms, "steps": e2e_steps,
           "h2d_bytes_per_step": int(batch * (M * K + K * N) * es),
           "d2h_bytes_per_step": int(batch * M * N * 4),
           "api": "iree_b200_ops_matmul_host (pinned host buffers)"}

    if rank != 0:
        return
    peaks = load_peaks()
    per_gpu = value / world
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(args.workload)
    line = {
        "metric": METRIC, "value": value, "unit": "TFLOP/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": {"bf16": "bf16 (f32 accumulate)", "f16": "f16 (f32 accumulate)",
                  "i8": "i8 (i32 accumulate)"}[et],
        "data": "synthetic",
        "config": {"workload": f"linalg.{op} {et} batch={batch} M={M} N={N} K={K} per GPU"
                               + (f", batch-sharded over {world} GPUs (no collective)"
                                  if world > 1 else ""),
                   "l2": "inputs+output (%.0f MiB) exceed the 126 MB L2; no flush" % (nbytes / 2**20),
                   "launch": "stream command buffer replay, one dispatch per step"},
        "roofline": {"bound": "tensor", "achieved": per_gpu, "peak": peaks["bf16_tflops"],
                     "unit": "TFLOP/s", "frac": per_gpu / peaks["bf16_tflops"],
                     "frac_of_sustained": per_gpu / peaks["bf16_tflops_sustained"],
                     "peak_source": peaks["source"] + " cuBLAS bf16 burst", "traffic": traffic,
                     "algorithmic_flop_per_launch": flop, "algorithmic_bytes_per_launch": nbytes},
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(et, threads=os.cpu_count() or 1)
    print(json.dumps(line), flush=True)


def cpu_sample(et, rows, N, K, seed=1):
    """Rows [0, rows) of the 8192^3 problem: a bounded sample of the same workload."""
    import oracle
    lhs = oracle.generate_random_matrix(et, (rows, K), seed)
    rhs = oracle.generate_random_matrix(et, (K, N), seed + 1)
    return lhs, rhs


def cpu_baseline(et, threads):
    """The oracle port (oracle/oracle.c, OpenMP) timed on this box's host cores."""
    import oracle
    rows, N, K = 1024, 4096, 4096
    lhs, rhs = cpu_sample(et, rows, N, K)
    oracle.matmul(et, lhs[:64], rhs)  # warm the thread pool
    t0 = time.perf_counter()
    oracle.matmul(et, lhs, rhs)
    dt = time.perf_counter() - t0
    return {"value": 2.0 * rows * N * K / dt / 1e12, "unit": "TFLOP/s", "cores": threads,
            "kind": "port",
            "sample": f"oracle matmul {et} M={rows} N={N} K={K} (first {rows} rows of "
                      f"BASELINE configs[0] 4096^3), 1 run, {dt:.2f} s"}


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    import oracle
    op, et, batch, M, N, K = MATMUL_WORKLOADS[args.workload]
    threads = os.cpu_count() or 1
    rows = 256
    lhs, rhs = cpu_sample(et, rows, N, K)
    for _ in range(args.warmup):
        oracle.matmul(et, lhs[:32], rhs)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle.matmul(et, lhs, rhs)
    dt = (time.perf_counter() - t0) / args.steps
    value = 2.0 * rows * N * K / dt / 1e12
    sample = (f"oracle port of the reference loops (iree-e2e-matmul-test.cc:28-121), OpenMP, "
              f"{et} M={rows} (of {M}) N={N} K={K} per step")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "TFLOP/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": et, "data": "synthetic",
            "config": {"workload": f"linalg.{op} {et} batch={batch} M={M} N={N} K={K} per GPU",
                       "sample": sample},
            "cpu_baseline": {"value": value, "unit": "TFLOP/s", "cores": threads, "kind": "port",
                             "sample": sample},
            "e2e": {"value": value, "unit": "TFLOP/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser(description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="matmul_bf16", choices=sorted(MATMUL_WORKLOADS))
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    _, world, _ = dist_env()
    if world != args.gpus and world == 1 and args.gpus > 1:
        # Convenience: re-launch under torchrun when called bare with --gpus N.
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
               f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1", "--master-port",
               os.environ.get("MASTER_PORT", "29517"), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
