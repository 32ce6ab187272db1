"""How long does a B200 keep its boost clock on a tensor-bound kernel after idling? Times the
batch_matmul workload (64 x 2048^3 bf16) in consecutive ~200 ms windows for several seconds
after a 3 s idle; prints ms per step, TFLOP/s, SM clock and power per window. Sets the soak
bench.py needs before a "sustained" line (bench.py, time_steps)."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pynvml

import bench
from iree_b200 import hal


def main():
    torch.cuda.init()
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(0)
    stream = torch.cuda.Stream()
    dev = hal.Device(0, mode=hal.MODE_STREAM, external_stream=stream.cuda_stream)
    ops = hal.Ops(dev)
    name = sys.argv[1] if len(sys.argv) > 1 else "batch_matmul"
    work = bench.make_step(torch, hal, dev, ops, name, seed=0)
    flop = work.describe()["flop"]
    with torch.cuda.stream(stream):
        for _ in range(3):
            work.step()
    torch.cuda.synchronize()
    time.sleep(3.0)
    t_start = time.perf_counter()
    for window in range(30):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 250
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(n):
                work.step()
            e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        print(json.dumps({"t_s": round(time.perf_counter() - t_start, 2), "ms_per_step": round(ms, 4),
                          "tflops": round(flop / ms / 1e9, 1),
                          "sm_mhz": pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM),
                          "power_w": pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0}), flush=True)


if __name__ == "__main__":
    main()
