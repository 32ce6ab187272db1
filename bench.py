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
The default workload's line also carries, device-timed in the same invocation:
  N = 1: "also"    — i8 8192^3, batch_matmul 64 x 2048^3, prefill and decode attention, the
                     decoder-layer GEMMs on one GPU (the other single-GPU BASELINE configs);
  N > 1: "sharded" — the configs BASELINE.json shards, STRONG scaling with their own N=1
                     figure and efficiency: batch_matmul 64 x 2048^3 split 64/N (no
                     collective) and the tensor-parallel decoder layer with its all-reduces
                     through iree_b200_channel_t (+ GEMMs alone, collectives alone, overlap).

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
    "matmul_f8": ("matmul", "f8e4m3fn", 1, 8192, 8192, 8192),
    "matmul_bf16_4096": ("matmul", "bf16", 1, 4096, 4096, 4096),
    "batch_matmul": ("batch_matmul", "bf16", 64, 2048, 2048, 2048),
    # The bf16 linear layer as one dispatch: matmul + bias[n] + relu + truncf.
    "linear_bf16": ("matmul", "bf16", 1, 8192, 8192, 8192),
}
FUSED_WORKLOADS = {"linear_bf16"}


# name -> (op, element type, batch*heads, M, K2, head dim)   [attention]
ATTENTION_WORKLOADS = {
    "attention_prefill": ("attention", "f16", 32, 4096, 4096, 128),
    "attention_prefill_bf16": ("attention", "bf16", 32, 4096, 4096, 128),
    "attention_decode": ("attention", "f16", 32, 1, 8192, 128),
    # The same prefill with the mask real Llama prefill carries: causal, computed in the kernel
    # (IREE_B200_ATTENTION_MASK_CAUSAL), masked K2 blocks skipped. Algorithmic FLOP count only
    # the unmasked (m, k2) pairs ("no causal discount unless a mask is present").
    "attention_prefill_causal": ("attention", "f16", 32, 4096, 4096, 128),
    "attention_prefill_causal_bf16": ("attention", "bf16", 32, 4096, 4096, 128),
}
CAUSAL_WORKLOADS = {"attention_prefill_causal", "attention_prefill_causal_bf16"}


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
    """Samples SM clock, power and throttle reasons WHILE the timed region runs: NVML polled
    every 1 ms from a thread (nvidia-smi -lms 100 as the fallback when pynvml is missing)."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index: int):
        self.index = index
        self.rows = []      # (sm_mhz, power_w, reason names)
        self.smax = None
        self.proc = None
        self.thread = None
        self.stop_flag = threading.Event()
        self.source = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML enumerates in PCI order; honour CUDA_VISIBLE_DEVICES when it is a plain list.
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            ids = [v for v in vis.split(",") if v.strip().isdigit()]
            nv_index = int(ids[self.index]) if self.index < len(ids) else self.index
            h = pynvml.nvmlDeviceGetHandleByIndex(nv_index)
            self.smax = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            bits = {"hw_slowdown": pynvml.nvmlClocksEventReasonHwSlowdown,
                    "hw_thermal_slowdown": pynvml.nvmlClocksEventReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": pynvml.nvmlClocksEventReasonSwThermalSlowdown,
                    "sw_power_cap": pynvml.nvmlClocksEventReasonSwPowerCap}

            def poll():
                while not self.stop_flag.is_set():
                    try:
                        sm = float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                        pw = pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0
                        r = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)
                        self.rows.append((sm, pw, [n for n, b in bits.items() if r & b]))
                    except Exception:
                        pass
                    time.sleep(0.001)

            self.thread = threading.Thread(target=poll, daemon=True)
            self.thread.start()
            self.source = "nvml 1 ms"
            return
        except Exception:
            self.thread = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return
        self.source = "nvidia-smi 100 ms"
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm, smax, pw = float(parts[1]), float(parts[2]), float(parts[3])
            except ValueError:
                continue
            self.smax = smax
            self.rows.append((sm, pw, [n for n, v in zip(self.NAMES, parts[5:9])
                                       if v.lower().startswith("active")]))

    def stop(self) -> dict:
        if not self.thread:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no clock source"]}
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        else:
            self.stop_flag.set()
            self.thread.join(timeout=2)
        sm = sorted(r[0] for r in self.rows)
        reasons = sorted({n for r in self.rows for n in r[2]})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.smax,
                "power_w_max": max((r[1] for r in self.rows), default=None),
                "samples": len(sm), "source": self.source, "reasons": reasons}


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
    es = 2 if et in ("f16", "bf16") else 1
    return 2.0 * batch * M * N * K, float(batch) * (M * K * es + K * N * es + M * N * 4)


def synth_matmul_inputs(torch, et, batch, M, N, K, seed):
    g = torch.Generator(device="cuda")
    g.manual_seed(0xB200 + seed)
    if et == "i8":
        a = torch.randint(-128, 128, (batch, M, K), dtype=torch.int8, device="cuda", generator=g)
        b = torch.randint(-128, 128, (batch, K, N), dtype=torch.int8, device="cuda", generator=g)
        c = torch.empty((batch, M, N), dtype=torch.int32, device="cuda")
    elif et.startswith("f8"):
        dt = torch.float8_e4m3fn if et == "f8e4m3fn" else torch.float8_e5m2
        a = torch.randn((batch, M, K), device="cuda", generator=g).clamp_(-4, 4).to(dt)
        b = torch.randn((batch, K, N), device="cuda", generator=g).clamp_(-4, 4).to(dt)
        c = torch.empty((batch, M, N), dtype=torch.float32, device="cuda")
    else:
        dt = torch.bfloat16 if et == "bf16" else torch.float16
        a = torch.randn((batch, M, K), device="cuda", generator=g).clamp_(-4, 4).to(dt)
        b = torch.randn((batch, K, N), device="cuda", generator=g).clamp_(-4, 4).to(dt)
        c = torch.empty((batch, M, N), dtype=torch.float32, device="cuda")
    return a, b, c


class MatmulStep:
    """One recorded command buffer holding the workload's dispatch; step() replays it."""

    def __init__(self, torch, hal, dev, ops, spec, seed, fused=False, mode=None):
        """|mode| (the "also" lines): "device_maps" = the `*_dtm` export, which builds its tensor
        maps on the device (what an unmodified IREE CUDA driver can launch); "mmt4d" = the same
        contraction as linalg.mmt4d on 16x16xK0 data tiles, read directly by the tcgen05 GEMM
        (operand bytes are synthetic, so the row-major tensors double as tiled ones)."""
        op, et, batch, M, N, K = spec
        self.spec = spec
        self.fused = fused
        self.mode = mode
        self.tensors = synth_matmul_inputs(torch, et, batch, M, N, K, seed)
        if fused:   # result in the operand type, plus the bias vector
            a, b, _ = self.tensors
            self.tensors = (a, b, torch.empty((batch, M, N), dtype=a.dtype, device="cuda"),
                            torch.randn(N, device="cuda"))
            self.flags = hal.MATMUL_RELU | hal.MATMUL_TRUNCATE_RESULT
        self.buffers = [dev.import_buffer(t.data_ptr(), t.numel() * t.element_size())
                        for t in self.tensors]
        before = ops.dispatch_count
        self.cb = hal.CommandBuffer(dev).begin()
        if fused:
            ops.record_matmul_bias(self.cb, et, batch, M, N, K, self.buffers[0].ref(),
                                   self.buffers[1].ref(), self.buffers[3].ref(),
                                   self.buffers[2].ref(), flags=self.flags)
        elif mode == "mmt4d":
            k0 = 16 // (2 if et in ("f16", "bf16") else 1)
            ops.record_mmt4d(self.cb, et, M // 16, N // 16, K // k0, 16, 16, k0,
                             self.buffers[0].ref(), self.buffers[1].ref(), self.buffers[2].ref())
        else:
            ops.record_matmul(self.cb, et, batch, M, N, K, self.buffers[0].ref(),
                              self.buffers[1].ref(), self.buffers[2].ref(),
                              flags=hal.MATMUL_DEVICE_MAPS if mode == "device_maps" else 0)
        self.cb.end()
        self.launches_per_step = ops.dispatch_count - before
        self.dev = dev

    def step(self):
        self.dev.execute(self.cb)

    def host_buffers(self, torch):
        host = [torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in self.tensors]
        host[0].copy_(self.tensors[0])
        host[1].copy_(self.tensors[1])
        if self.fused:
            host[3].copy_(self.tensors[3])
        return host

    def host_call(self, ops, host):
        op, et, batch, M, N, K = self.spec
        if self.fused:
            ops.matmul_bias_host(et, batch, M, N, K, host[0].data_ptr(), host[1].data_ptr(),
                                 host[3].data_ptr(), host[2].data_ptr(), flags=self.flags)
            return
        ops.matmul_host(et, batch, M, N, K, host[0].data_ptr(), host[1].data_ptr(),
                        host[2].data_ptr())

    def describe(self):
        op, et, batch, M, N, K = self.spec
        es = 2 if et in ("f16", "bf16") else 1
        flop = 2.0 * batch * M * N * K
        out_es = es if self.fused else 4
        nbytes = float(batch) * (M * K * es + K * N * es + M * N * out_es)
        name = f"linalg.{op}" + (" + bias[n] + relu + truncf (one dispatch)" if self.fused else "")
        if self.fused:
            return {"flop": flop, "bytes": nbytes + N * 4, "bound": "tensor",
                    "workload": f"{name} {et} batch={batch} M={M} N={N} K={K} per GPU",
                    "dtype": f"{et} (f32 accumulate + f32 bias, {et} result)",
                    "h2d": int(batch * (M * K + K * N) * es + N * 4),
                    "d2h": int(batch * M * N * out_es),
                    "api": "iree_b200_ops_matmul_bias_host (pinned host buffers)"}
        if self.mode == "mmt4d":
            k0 = 16 // es
            op = f"mmt4d (tiles 16x16x{k0}, read directly through rank-4 tensor maps)"
        elif self.mode == "device_maps":
            op += " (tensor maps built on the device: the stock-driver ABI)"
        return {"flop": flop, "bytes": nbytes, "bound": "tensor",
                "workload": f"linalg.{op} {et} batch={batch} M={M} N={N} K={K} per GPU",
                "dtype": {"bf16": "bf16 (f32 accumulate)", "f16": "f16 (f32 accumulate)",
                          "i8": "i8 (i32 accumulate)", "f8e4m3fn": "f8E4M3FN (f32 accumulate)",
                          "f8e5m2": "f8E5M2 (f32 accumulate)"}[et],
                "h2d": int(batch * (M * K + K * N) * es), "d2h": int(batch * M * N * 4),
                "api": "iree_b200_ops_matmul_host (pinned host buffers)"}


class MxStep:
    """MXFP4 matmul (f4E2M1FN x f4E2M1FN, f8E8M0FNU scale per 32 elements, f32 result; SURVEY
    §8(f) rank 3, tests/e2e/matmul/generate_code_mx.py) as one dispatch of the block-scaled
    tcgen05 kernel."""

    def __init__(self, torch, hal, dev, ops, spec, seed):
        _, _, _, M, N, K = spec
        self.spec = spec
        g = torch.Generator(device="cuda")
        g.manual_seed(0xB200 + seed)
        rnd = lambda shape, lo, hi: torch.randint(lo, hi, shape, dtype=torch.uint8, device="cuda",
                                                  generator=g)
        self.tensors = [rnd((M, K // 2), 0, 256), rnd((M, K // 32), 120, 135),
                        rnd((N, K // 2), 0, 256), rnd((N, K // 32), 120, 135),
                        torch.empty((M, N), dtype=torch.float32, device="cuda")]
        # Workspace for the scales in the instruction's atom order (two small repack dispatches
        # ahead of the GEMM, every step: the scales of activations change per call).
        self.tensors.append(torch.empty(max(16, ops.matmul_mxfp4_scale_workspace_size(M, N, K)),
                                        dtype=torch.uint8, device="cuda"))
        self.buffers = [dev.import_buffer(t.data_ptr(), t.numel() * t.element_size())
                        for t in self.tensors]
        before = ops.dispatch_count
        self.cb = hal.CommandBuffer(dev).begin()
        b = self.buffers
        ops.record_matmul_mxfp4_tensor(self.cb, M, N, K, b[0].ref(), b[1].ref(), b[2].ref(),
                                       b[3].ref(), b[5].ref(), b[4].ref())
        self.cb.end()
        self.launches_per_step = ops.dispatch_count - before
        self.dev = dev

    def step(self):
        self.dev.execute(self.cb)

    def host_buffers(self, torch):
        host = [torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in self.tensors[:5]]
        for h, t in zip(host[:4], self.tensors[:4]):
            h.copy_(t)
        return host

    def host_call(self, ops, host):
        _, _, _, M, N, K = self.spec
        ops.matmul_mxfp4_host(M, N, K, *(h.data_ptr() for h in host))

    def describe(self):
        _, _, _, M, N, K = self.spec
        operand = (M + N) * (K // 2 + K // 32)
        return {"flop": 2.0 * M * N * K, "bytes": float(operand + M * N * 4), "bound": "tensor",
                "workload": f"MXFP4 matmul (f4E2M1FN, f8E8M0FNU scale / 32 elements) M={M} N={N} K={K} "
                            f"-> f32, tcgen05.mma.kind::mxf4.block_scale (scale repack x2 + GEMM), per GPU",
                "dtype": "f4E2M1FN x f8E8M0FNU block scale (f32 accumulate)",
                "h2d": int(operand), "d2h": int(M * N * 4),
                "api": "iree_b200_ops_matmul_mxfp4_host (pinned host buffers)"}


class AttentionStep:
    """iree_linalg_ext.attention, one recorded command buffer (1 or 2 dispatches)."""
    FLAGS = 0  # --attention-flags (diagnostics: kernel family / exp2 split)

    def __init__(self, torch, hal, dev, ops, spec, seed, causal=False):
        op, et, B, M, K2, D = spec
        self.spec = spec
        self.causal = causal
        self.mask_kind = hal.ATTENTION_MASK_CAUSAL if causal else hal.ATTENTION_MASK_NONE
        g = torch.Generator(device="cuda")
        g.manual_seed(0xB200 + seed)
        dt = torch.bfloat16 if et == "bf16" else torch.float16
        mk = lambda *shape: torch.randn(shape, device="cuda", generator=g).clamp_(-4, 4).to(dt)
        self.tensors = [mk(B, M, D), mk(B, K2, D), mk(B, K2, D),
                        torch.empty((B, M, D), dtype=dt, device="cuda")]
        self.scale = 1.0 / (D ** 0.5)
        sel = ops.select_attention(et, B, M, D, K2, D, self.scale, mask_kind=self.mask_kind,
                                   flags=self.FLAGS)
        self.workspace = torch.empty((max(sel["workspace_bytes"], 16),), dtype=torch.uint8,
                                     device="cuda")
        imp = lambda t: dev.import_buffer(t.data_ptr(), t.numel() * t.element_size())
        self.buffers = [imp(t) for t in self.tensors] + [imp(self.workspace)]
        before = ops.dispatch_count
        self.cb = hal.CommandBuffer(dev).begin()
        ops.record_attention(self.cb, et, B, M, D, K2, D, self.scale, self.buffers[0].ref(),
                             self.buffers[1].ref(), self.buffers[2].ref(), self.buffers[3].ref(),
                             workspace=self.buffers[4].ref(), mask_kind=self.mask_kind,
                             flags=self.FLAGS)
        self.cb.end()
        self.launches_per_step = ops.dispatch_count - before
        self.dev = dev

    def step(self):
        self.dev.execute(self.cb)

    def host_buffers(self, torch):
        host = [torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in self.tensors]
        for h, t in zip(host[:3], self.tensors[:3]):
            h.copy_(t)
        return host

    def host_call(self, ops, host):
        op, et, B, M, K2, D = self.spec
        ops.attention_host(et, B, M, D, K2, D, self.scale, host[0].data_ptr(), host[1].data_ptr(),
                           host[2].data_ptr(), host[3].data_ptr(), mask_kind=self.mask_kind)

    def describe(self):
        op, et, B, M, K2, D = self.spec
        flop = 4.0 * B * M * K2 * D
        if self.causal:   # only the unmasked pairs: row m sees m + (K2 - M) + 1 keys
            flop = 4.0 * B * D * (M * (M + 1) / 2.0 + float(M) * (K2 - M))
        nbytes = float(B) * (2 * M * D + 2 * K2 * D) * 2
        mask = "causal mask (k2 <= m + K2 - M) computed in the kernel" if self.causal else "no mask"
        return {"flop": flop, "bytes": nbytes, "bound": "hbm" if M <= 16 else "tensor",
                "workload": f"iree_linalg_ext.attention {et} batch*heads={B} M={M} K2={K2} "
                            f"K1=N={D} scale=1/sqrt({D}) {mask}, per GPU",
                "dtype": f"{et} (f32 S/softmax/accumulate)",
                "h2d": int(B * (M * D + 2 * K2 * D) * 2), "d2h": int(B * M * D * 2),
                "api": "iree_b200_ops_attention_host (pinned host buffers)"}


class TpLayerStep:
    """BASELINE configs[4]: the GEMMs of one Llama-3-8B decoder layer (hidden 4096, 32 q / 8 kv
    heads x 128, FFN 14336; tokens = 8192), tensor-parallel over the ranks (Megatron split):
    QKV and gate/up column-parallel, O and down row-parallel followed by an all-reduce(sum) of
    the [tokens, 4096] bf16 activation through iree_hal_channel. Recording and schedule live in
    iree_b200/tp_layer.py (row-parallel GEMMs and gate/up in token chunks, staged so that the
    all-reduce of a chunk runs beside the GEMMs of its neighbours). All GEMMs use the
    truncating epilogue (bf16 activations)."""

    TOKENS, HIDDEN, QKV, FFN, CHUNKS = 8192, 4096, 6144, 14336, 4
    STAGED = True
    SCHEDULE = "dataflow"

    def __init__(self, torch, hal, dev, ops, world, rank, channel, seed, region=None):
        from iree_b200 import tp_layer
        self.layer = tp_layer.DecoderLayerGemms(
            torch, hal, dev, ops, world, rank, channel, tokens=self.TOKENS, hidden=self.HIDDEN,
            ffn=self.FFN, chunks=self.CHUNKS, region=region, seed=seed,
            chunk_consumers=self.STAGED, schedule=self.SCHEDULE)
        assert self.layer.dims["qkv"] == self.QKV
        self.t, self.b, self.cb = self.layer.t, self.layer.b, self.layer.cb
        self.world, self.dev, self.channel, self.region = world, dev, channel, region
        self.flop = self.layer.flop
        self.collectives = self.layer.collectives
        self.launches_per_step = self.layer.launches_per_step

    def step(self):
        self.dev.execute(self.cb)

    def host_buffers(self, torch):
        x = torch.empty(self.t["x"].shape, dtype=torch.bfloat16).pin_memory()
        x.copy_(self.t["x"])
        return [x, torch.empty(self.t["d"].shape, dtype=torch.bfloat16).pin_memory()]

    def host_call(self, ops, host):
        n = host[0].numel() * 2
        self.dev.h2d(host[0].data_ptr(), self.b["x"], n)
        self.dev.execute(self.cb)
        self.dev.d2h(self.b["d"], host[1].data_ptr(), n)
        self.dev.wait_idle()

    def describe(self):
        T, H, QKV, FFN, w = self.TOKENS, self.HIDDEN, self.QKV, self.FFN, self.world
        weights = 2.0 * (H * QKV + H * H + H * 2 * FFN + FFN * H) / w
        acts = 2.0 * T * (H + QKV / w + H / w + H + H + 2 * FFN / w + FFN / w + H)
        return {"flop": self.flop, "bytes": weights + acts, "bound": "tensor",
                "workload": f"Llama-3-8B decoder-layer GEMMs (QKV/O/gate+up/down), tokens={T}, "
                            f"bf16 -> bf16 (f32 accumulate), tensor-parallel over {w} GPU(s), "
                            f"{self.collectives} all-reduce(sum, bf16, chunks of {self.CHUNKS} x [.,{H}]) per "
                            f"step via iree_hal_channel ({self.layer.collective_path()}), "
                            f"schedule {self.SCHEDULE if w > 1 else 'sequence'}: collectives on "
                            f"the collective stream beside the GEMMs that do not consume them",
                "dtype": "bf16 (f32 accumulate)", "h2d": int(T * H * 2), "d2h": int(T * H * 2),
                "api": "h2d(x) + iree_b200_device_queue_execute(layer command buffer) + d2h(y)"}


MX_WORKLOADS = {"matmul_mxfp4": ("matmul", "mxfp4", 1, 8192, 8192, 8192)}


def make_step(torch, hal, dev, ops, workload, seed):
    if workload in MX_WORKLOADS:
        return MxStep(torch, hal, dev, ops, MX_WORKLOADS[workload], seed)
    if workload in MATMUL_WORKLOADS:
        return MatmulStep(torch, hal, dev, ops, MATMUL_WORKLOADS[workload], seed,
                          fused=workload in FUSED_WORKLOADS)
    return AttentionStep(torch, hal, dev, ops, ATTENTION_WORKLOADS[workload], seed,
                         causal=workload in CAUSAL_WORKLOADS)


SOAK_MS = 1000.0   # --extra-soak-ms


def time_steps(torch, works, steps, warmup, stream, world, sampler=None, min_ms=0.0):
    """|warmup| untimed steps, then exactly |steps| between two CUDA events on |stream|,
    bracketed by barrier + synchronize; returns ms per step, max over ranks. |min_ms| > 0
    (the extra lines only, never the headline): |steps| is raised so that the timed region
    lasts at least that long, and the GPU is first soaked with SOAK_MS of the same work,
    untimed — a 20 ms burst runs at 1965 MHz, before the power limiter has settled, and reads
    30-40 % above what the same kernel sustains; from a COLD GPU (the first line after the
    host-link copies) even a 200 ms region stayed at the boost clock on some boxes
    (batch_matmul 0.63 ms cold vs 0.92-0.94 ms warm), which made T(1) / (N T(N)) depend on
    the order of the lines."""
    sets = len(works)
    with torch.cuda.stream(stream):
        for i in range(warmup):
            works[i % sets].step()
        if min_ms > 0:
            barrier(world)
            p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            p0.record(stream)
            for i in range(5):
                works[i % sets].step()
            p1.record(stream)
            barrier(world)
            pilot = max_over_ranks(p0.elapsed_time(p1), world) / 5
            steps = int(max(steps, min(5000, -(-min_ms // max(pilot, 1e-3)))))
            for i in range(int(min(20000, SOAK_MS // max(pilot, 1e-3)))):
                works[i % sets].step()
        barrier(world)
        if sampler is not None:
            sampler.start()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(stream)
        for i in range(steps):
            works[i % sets].step()
        ev1.record(stream)
        barrier(world)
    ms = max_over_ranks(ev0.elapsed_time(ev1), world) / steps
    return (ms, steps) if min_ms > 0 else ms


def soak(torch, work, stream, world, seconds):
    """Runs |work| back to back for |seconds| of wall clock, untimed. After seconds of
    PCIe-bound phases (e2e, host-link copies) a B200 ran the FIRST tensor-bound line at the
    boost clock for more than 1.4 s before the power limiter pulled it to the sustained clock
    (batch_matmul 64 x 2048^3: 0.64 ms per step there, 0.93-0.97 ms on every later line and in
    every isolated run — tools/power_budget_probe.py); with T(1) measured first that read as a
    strong-scaling efficiency of 0.65 for a split that is 1.00. MEASURED_PEAKS.json's sustained
    figure is "back to back for 4 s": the extra lines start from the same state."""
    if seconds <= 0:
        return
    barrier(world)
    t0 = time.perf_counter()
    with torch.cuda.stream(stream):
        while time.perf_counter() - t0 < seconds:
            for _ in range(50):
                work.step()
            stream.synchronize()
    barrier(world)


def rotating_sets(make, info_bytes):
    """Input sets cycled through between timed steps so that the working set exceeds L2."""
    first = make(0)
    sets = max(1, min(4, int(-(-2 * 126e6 // info_bytes(first)))))
    return [first] + [make(i) for i in range(1, sets)]


# A timed region shorter than this runs before the power limiter settles (1965 MHz) and is
# compared with the BURST cuBLAS figure; a longer one runs power-capped (~1.4 GHz) and is
# compared with the SUSTAINED figure (MEASURED_PEAKS.json holds both: best of 10 single
# launches / back to back for 4 s). Both fractions are always printed.
BURST_REGION_MS = 50.0


def roofline_of(info, per_gpu_tflops, ms_per_step, peaks, steps=1):
    if info["bound"] == "tensor":
        region_ms = ms_per_step * steps
        regime = "burst" if region_ms < BURST_REGION_MS else "sustained"
        peak = peaks["bf16_tflops"] if regime == "burst" else peaks["bf16_tflops_sustained"]
        return {"bound": "tensor", "achieved": per_gpu_tflops, "peak": peak,
                "unit": "TFLOP/s", "frac": per_gpu_tflops / peak,
                "frac_of_burst": per_gpu_tflops / peaks["bf16_tflops"],
                "frac_of_sustained": per_gpu_tflops / peaks["bf16_tflops_sustained"],
                "frac_of_nominal_2250": per_gpu_tflops / 2250.0,
                "peak_source": "%s cuBLAS bf16 %s (timed region %.0f ms)"
                               % (peaks["source"], regime, region_ms)}
    gbs = info["bytes"] / (ms_per_step * 1e-3) / 1e9
    return {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
            "frac": gbs / peaks["hbm_gbs"], "frac_of_nominal_8000": gbs / 8000.0,
            "peak_source": peaks["source"] + " copy"}


def host_link_line(torch, world):
    """The PCIe floor under `e2e`: H2D alone, D2H alone and both at once, 256 MiB pinned
    buffers, every rank at the same time (aggregate = sum over ranks of bytes / slowest
    rank's time). `e2e` moves its operands over this link inside the timed region, so its
    scaling with N is bounded by how the aggregate scales — a property of the box (how many
    GPUs share a host bridge / NUMA node), not of the kernels."""
    n = 256 << 20
    host_a = torch.empty(n, dtype=torch.uint8).pin_memory()
    host_b = torch.empty(n, dtype=torch.uint8).pin_memory()
    dev_a = torch.empty(n, dtype=torch.uint8, device="cuda")
    dev_b = torch.empty(n, dtype=torch.uint8, device="cuda")
    up, down = torch.cuda.Stream(), torch.cuda.Stream()
    out = {}

    def timed(fn, bytes_moved, reps=4):
        fn()
        barrier(world)
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        ms = max_over_ranks((time.perf_counter() - t0) * 1e3, world) / reps
        return bytes_moved * world / (ms * 1e-3) / 1e9

    def h2d():
        with torch.cuda.stream(up):
            dev_a.copy_(host_a, non_blocking=True)

    def d2h():
        with torch.cuda.stream(down):
            host_b.copy_(dev_b, non_blocking=True)

    def both():
        h2d()
        d2h()

    def sync(fn):
        def run():
            fn()
            torch.cuda.synchronize()
        return run

    out["h2d_gbs_aggregate"] = timed(sync(h2d), n)
    out["d2h_gbs_aggregate"] = timed(sync(d2h), n)
    out["duplex_gbs_aggregate_each_way"] = timed(sync(both), n)
    out["ranks"] = world
    out["what"] = "256 MiB pinned copies, all ranks concurrently; GB/s summed over ranks"
    return out


def make_tp_device(torch, hal, args, world, rank, local):
    """Device + ops + channel (+ symmetric region) for the tensor-parallel layer: two streams
    (overlap_collectives), a few SMs kept free of persistent GEMM CTAs for the collective
    kernels, the channel created by iree_b200_channel_create through the DEFAULT provider
    (rank / count from the launcher's environment, id through IREE_B200_CHANNEL_ID_FILE)."""
    stream = torch.cuda.Stream()
    overlap = 0
    if world > 1:
        overlap = {"fused": 1, "staged": 2, "dataflow": 3}[args.tp_schedule]
    dev = hal.Device(local, mode=hal.MODE_STREAM, external_stream=stream.cuda_stream,
                     overlap_collectives=overlap)
    ops = hal.Ops(dev)
    channel = region = None
    if world > 1:
        import tempfile
        os.environ.setdefault("IREE_B200_CHANNEL_ID_FILE", os.path.join(
            tempfile.gettempdir(), "iree_b200_bench_id." + os.environ.get("MASTER_PORT", "0")))
        channel = hal.Channel(dev)
        assert channel.rank_and_count() == (rank, world)
        ops.set_reserved_sms(args.reserved_sms)
        if args.collective == "symmetric":
            from iree_b200 import symmetric
            region = symmetric.SymmetricRegion(
                2 * TpLayerStep.TOKENS * TpLayerStep.HIDDEN * 2, torch.device("cuda", local))
            blocks = args.ar_blocks or (args.reserved_sms if args.reserved_sms > 0 else 64)
            region.attach(channel, ops.executable, max_blocks=max(1, min(64, blocks)))
    if args.tp_chunk_rows:
        TpLayerStep.CHUNKS = [int(r) for r in args.tp_chunk_rows.split(",")]
    elif args.tp_chunks:
        TpLayerStep.CHUNKS = args.tp_chunks
    elif world >= 4:
        # All-reduce-bound: a short first chunk starts the link early, two large ones keep the
        # per-collective cost (two rank meetings + ramp, ~20 us) low
        # (profiles/tp_chunk_experiments_n8_r2.jsonl: 0.559 ms vs 0.586 for 4 equal chunks).
        T = TpLayerStep.TOKENS
        TpLayerStep.CHUNKS = [T // 8, 7 * T // 16, T - T // 8 - 7 * T // 16]
    else:
        TpLayerStep.CHUNKS = 4
    TpLayerStep.STAGED = args.tp_schedule != "fused"
    TpLayerStep.SCHEDULE = args.tp_schedule
    return stream, dev, ops, channel, region


def sharded_lines(torch, hal, args, world, rank, local, peaks):
    """N > 1: the two configurations BASELINE.json shards, STRONG scaling, each
    with its own N=1 figure measured in this same invocation (every rank runs the unsharded
    workload on its own GPU at the same time; the slowest rank counts) so that the efficiency
    is T(1) / (N * T(N)) on one box, one run:
      batch_matmul       configs[2]: 64 x 2048^3 bf16, batch split 64/N, no collective
      tp_decoder_layer   configs[4]: Llama-3-8B layer GEMMs, Megatron split, all-reduce through
                         iree_b200_channel_t, + the step's GEMMs alone and collectives alone"""
    out = {"host_link": host_link_line(torch, world)}
    steps, warmup = max(10, min(args.steps, 50)), max(3, min(args.warmup, 5))
    MIN = max(args.extra_min_ms, 1.0)   # every extra line is timed over at least this long (sustained)
    stream = torch.cuda.Stream()
    dev = hal.Device(local, mode=hal.MODE_STREAM, external_stream=stream.cuda_stream)
    ops = hal.Ops(dev)
    # ---- configs[2]
    op, et, B, M, N, K = MATMUL_WORKLOADS["batch_matmul"]
    if B % world == 0:   # (every N the driver runs divides 64)
        full = [MatmulStep(torch, hal, dev, ops, (op, et, B, M, N, K), seed=rank)]
        soak(torch, full[0], stream, world, args.presoak_s)
        t1, _ = time_steps(torch, full, steps, warmup, stream, world, min_ms=MIN)
        del full
        torch.cuda.empty_cache()
        part = rotating_sets(lambda i: MatmulStep(torch, hal, dev, ops, (op, et, B // world, M, N, K),
                                                  seed=rank + 100 * i),
                             lambda w: w.describe()["bytes"])
        tn, steps_n = time_steps(torch, part, steps, warmup, stream, world, min_ms=MIN)
        flop = 2.0 * B * M * N * K
        out["batch_matmul"] = {
            "workload": f"linalg.batch_matmul bf16 batch={B} M=N=K={M}, batch split {B // world} per GPU "
                        f"over {world} GPUs, no collective", "scaling": "strong",
            "value": flop / (tn * 1e-3) / 1e12, "unit": "TFLOP/s", "ms_per_step": tn,
            "n1_ms_per_step": t1, "n1_value": flop / (t1 * 1e-3) / 1e12,
            "efficiency": t1 / (world * tn), "steps": steps_n, "timed_region_ms": tn * steps_n,
            "roofline_frac": flop / world / (tn * 1e-3) / 1e12 / peaks["bf16_tflops"],
            "roofline_frac_of_sustained": flop / world / (tn * 1e-3) / 1e12 / peaks["bf16_tflops_sustained"]}
        del part
        torch.cuda.empty_cache()
    # ---- configs[3]: attention, heads (= batch entries) split 32/N, no collective
    try:
        aop, aet, AB, AM, AK2, AD = ATTENTION_WORKLOADS["attention_prefill"]
        if AB % world == 0:
            full = [AttentionStep(torch, hal, dev, ops, (aop, aet, AB, AM, AK2, AD), seed=rank)]
            t1, _ = time_steps(torch, full, steps, warmup, stream, world, min_ms=MIN)
            del full
            torch.cuda.empty_cache()
            part = rotating_sets(lambda i: AttentionStep(torch, hal, dev, ops,
                                                         (aop, aet, AB // world, AM, AK2, AD),
                                                         seed=rank + 100 * i),
                                 lambda w: w.describe()["bytes"])
            tn, steps_n = time_steps(torch, part, steps, warmup, stream, world, min_ms=MIN)
            aflop = 4.0 * AB * AM * AK2 * AD
            out["attention_prefill"] = {
                "workload": f"iree_linalg_ext.attention {aet} batch*heads={AB} M={AM} K2={AK2} K1=N={AD}, "
                            f"heads split {AB // world} per GPU over {world} GPUs, no collective",
                "scaling": "strong", "value": aflop / (tn * 1e-3) / 1e12, "unit": "TFLOP/s",
                "ms_per_step": tn, "n1_ms_per_step": t1, "n1_value": aflop / (t1 * 1e-3) / 1e12,
                "efficiency": t1 / (world * tn), "steps": steps_n, "timed_region_ms": tn * steps_n,
                "launches_per_step": part[0].launches_per_step,
                "note": f"{AB // world * ((AM + 255) // 256)} CTA-units of work per GPU on 148 SMs: "
                        "below one full wave the tail split cuts units over K2 (+ a merge dispatch)"}
            del part
            torch.cuda.empty_cache()
    except Exception as e:  # noqa: BLE001 - an extra line must never cost the headline
        out["attention_prefill"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    # ---- configs[4]
    solo = TpLayerStep(torch, hal, dev, ops, 1, 0, None, seed=rank)
    t1, _ = time_steps(torch, [solo], steps, warmup, stream, world, min_ms=MIN)
    layer_flop = solo.flop
    del solo
    torch.cuda.empty_cache()
    tstream, tdev, tops, channel, region = make_tp_device(torch, hal, args, world, rank, local)
    step = TpLayerStep(torch, hal, tdev, tops, world, rank, channel, seed=rank, region=region)
    tn, steps_n = time_steps(torch, [step], steps, warmup, tstream, world, min_ms=MIN)
    info = step.describe()
    # The step's two halves alone: GEMMs without collectives, collectives without GEMMs.
    from iree_b200 import tp_layer
    kw = dict(tokens=TpLayerStep.TOKENS, hidden=TpLayerStep.HIDDEN, ffn=TpLayerStep.FFN,
              chunks=TpLayerStep.CHUNKS, region=region, seed=rank, chunk_consumers=TpLayerStep.STAGED,
              schedule=TpLayerStep.SCHEDULE)
    gemms = tp_layer.DecoderLayerGemms(torch, hal, tdev, tops, world, rank, channel,
                                       only="gemms", **kw)
    t_gemm, _ = time_steps(torch, [gemms], steps, warmup, tstream, world, min_ms=MIN)
    colls = tp_layer.DecoderLayerGemms(torch, hal, tdev, tops, world, rank, channel,
                                       only="collectives", **kw)
    t_coll, _ = time_steps(torch, [colls], steps, warmup, tstream, world, min_ms=MIN / 2)
    payload = 2.0 * TpLayerStep.TOKENS * TpLayerStep.HIDDEN * 2   # two [tokens, hidden] bf16
    out["tp_decoder_layer"] = {
        "workload": info["workload"], "scaling": "strong",
        "value": layer_flop / (tn * 1e-3) / 1e12, "unit": "TFLOP/s", "ms_per_step": tn,
        "n1_ms_per_step": t1, "n1_value": layer_flop / (t1 * 1e-3) / 1e12,
        "efficiency": t1 / (world * tn), "steps": steps_n, "timed_region_ms": tn * steps_n,
        "gemm_ms": t_gemm, "collective_ms": t_coll,
        "overlap_frac": max(0.0, (t_gemm + t_coll - tn) / max(1e-9, min(t_gemm, t_coll))),
        "collective": {"what": f"{step.collectives} x all_reduce(sum, bf16) = {payload / 2**20:.0f} MiB "
                               f"per step, {args.collective}",
                       "algbw_gbs": payload / (t_coll * 1e-3) / 1e9,
                       "nvlink_bound_ms": payload * (1.0 + 1.0 / world) / 900e9 * 1e3,
                       "frac_of_nvlink_bound": payload * (1.0 + 1.0 / world) / 900e9 * 1e3 / t_coll},
        "limiter": "collective" if t_coll > t_gemm else "gemm",
        "launches_per_step": step.launches_per_step}
    return out


ALSO_WORKLOADS = ["matmul_i8", "batch_matmul", "attention_prefill", "attention_prefill_causal",
                  "attention_decode"]


def also_lines(torch, hal, dev, ops, stream, args, peaks):
    """N = 1: the other single-GPU configurations of BASELINE.json, driver-runnable, as short
    device-timed runs under the "also" key (value + roofline fraction each)."""
    out = {"host_link": host_link_line(torch, 1)}
    steps, warmup = max(10, min(args.steps, 50)), max(3, min(args.warmup, 5))
    for name in ALSO_WORKLOADS:
        works = rotating_sets(lambda i: make_step(torch, hal, dev, ops, name, seed=100 * i),
                              lambda w: w.describe()["bytes"])
        if name == ALSO_WORKLOADS[0]:
            soak(torch, works[0], stream, 1, args.presoak_s)
        ms, steps_n = time_steps(torch, works, steps, warmup, stream, 1, min_ms=max(args.extra_min_ms, 1.0))
        info = works[0].describe()
        tflops = info["flop"] / (ms * 1e-3) / 1e12
        roof = roofline_of(info, tflops, ms, peaks, steps_n)
        out[name] = {"workload": info["workload"], "value": tflops, "unit": "TFLOP/s",
                     "ms_per_step": ms, "steps": steps_n, "timed_region_ms": ms * steps_n,
                     "dtype": info["dtype"],
                     "launches_per_step": works[0].launches_per_step,
                     "roofline": {k: roof[k] for k in ("bound", "achieved", "peak", "unit", "frac",
                                                        "peak_source")}}
        del works
        torch.cuda.empty_cache()
    # The headline contraction through the two other doors of the boundary: the export an
    # unmodified IREE CUDA driver launches, and linalg.mmt4d on data-tiled operands.
    for name, mode in (("matmul_bf16_device_maps", "device_maps"), ("mmt4d_bf16", "mmt4d")):
        works = rotating_sets(lambda i: MatmulStep(torch, hal, dev, ops, MATMUL_WORKLOADS["matmul_bf16"],
                                                   100 * i, mode=mode),
                              lambda w: w.describe()["bytes"])
        ms, steps_n = time_steps(torch, works, steps, warmup, stream, 1, min_ms=max(args.extra_min_ms, 1.0))
        info = works[0].describe()
        tflops = info["flop"] / (ms * 1e-3) / 1e12
        roof = roofline_of(info, tflops, ms, peaks, steps_n)
        out[name] = {"workload": info["workload"], "value": tflops, "unit": "TFLOP/s",
                     "ms_per_step": ms, "steps": steps_n, "timed_region_ms": ms * steps_n,
                     "dtype": info["dtype"], "launches_per_step": works[0].launches_per_step,
                     "roofline": {k: roof[k] for k in ("bound", "achieved", "peak", "unit", "frac",
                                                        "peak_source")}}
        del works
        torch.cuda.empty_cache()
    solo = TpLayerStep(torch, hal, dev, ops, 1, 0, None, seed=0)
    ms, steps_n = time_steps(torch, [solo], steps, warmup, stream, 1, min_ms=max(args.extra_min_ms, 1.0))
    tflops = solo.flop / (ms * 1e-3) / 1e12
    out["tp_decoder_layer_n1"] = {"workload": solo.describe()["workload"], "value": tflops,
                                  "unit": "TFLOP/s", "ms_per_step": ms, "steps": steps_n,
                                  "roofline": {"bound": "tensor",
                                               "frac": tflops / peaks["bf16_tflops_sustained"],
                                               "peak": peaks["bf16_tflops_sustained"],
                                               "peak_source": peaks["source"] + " cuBLAS bf16 sustained"}}
    return out


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

    peaks = load_peaks()
    AttentionStep.FLAGS = args.attention_flags
    tp = args.workload == "tp_decoder_layer"
    if tp:
        stream, dev, ops, channel, region = make_tp_device(torch, hal, args, world, rank, local)
        first = TpLayerStep(torch, hal, dev, ops, world, rank, channel, seed=rank, region=region)
        works = [first]
    else:
        stream = torch.cuda.Stream()
        dev = hal.Device(local, mode=hal.MODE_STREAM, external_stream=stream.cuda_stream)
        ops = hal.Ops(dev)
        # Rotating input sets: the working set cycled through between timed steps exceeds L2.
        works = rotating_sets(lambda i: make_step(torch, hal, dev, ops, args.workload,
                                                  seed=rank + 100 * i),
                              lambda w: w.describe()["bytes"])
        first = works[0]
    info = first.describe()
    sets = len(works)
    flop, nbytes = info["flop"], info["bytes"]

    # ---- device-resident timing (CUDA events on the dispatch stream) ----
    sampler = ClockSampler(local) if rank == 0 else None
    ms_per_step = time_steps(torch, works, args.steps, args.warmup, stream, world, sampler)
    clocks = sampler.stop() if rank == 0 else None
    # Weak scaling: every rank's |flop|; the TP layer: |flop| is this rank's share (strong).
    value = flop * world / (ms_per_step * 1e-3) / 1e12
    launches = first.launches_per_step * args.steps

    # ---- end to end through the host-buffer C-ABI call ----
    host = first.host_buffers(torch)
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    for _ in range(2):
        first.host_call(ops, host)
    barrier(world)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        first.host_call(ops, host)
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3, world) / e2e_steps
    barrier(world)
    e2e = {"value": flop * world / (e2e_ms * 1e-3) / 1e12, "unit": "TFLOP/s",
           "ms_per_step": e2e_ms, "steps": e2e_steps,
           "h2d_bytes_per_step": info["h2d"], "d2h_bytes_per_step": info["d2h"],
           "api": info["api"]}
    del host

    extra = {}
    if args.workload == "matmul_bf16" and not args.no_extra:
        if world > 1:
            del works
            torch.cuda.empty_cache()
            extra["sharded"] = sharded_lines(torch, hal, args, world, rank, local, peaks)
        else:
            extra["also"] = also_lines(torch, hal, dev, ops, stream, args, peaks)

    if rank != 0:
        return
    if tp:
        info = first.describe()  # refreshed: how many collectives went which way
    per_gpu = value / world
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(args.workload)
    roofline = roofline_of(info, per_gpu, ms_per_step, peaks, args.steps)
    roofline.update({"traffic": traffic, "algorithmic_flop_per_launch": flop,
                     "algorithmic_bytes_per_launch": nbytes})
    line = {
        "metric": METRIC, "value": value, "unit": "TFLOP/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "strong" if tp else "weak", "vs_baseline": None,
        "dtype": info["dtype"], "data": "synthetic",
        "config": {"workload": workload_string(args.workload, world, info),
                   "l2": "%d rotating input sets of %.0f MiB each (> 126 MB L2 between reuses); "
                         "no flush" % (sets, nbytes / 2**20),
                   "launch": "stream command buffer replay, %d dispatch(es) per step"
                             % first.launches_per_step},
        "roofline": roofline, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
    }
    line.update(extra)
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args.workload, threads=use_all_host_threads())
    print(json.dumps(line), flush=True)


def workload_string(workload, world, info=None):
    """config.workload — identical in both arms (the reference arm times a bounded sample of
    this same workload; its own `sample` key says which)."""
    if info is None:
        if workload in MX_WORKLOADS:
            _, _, _, M, N, K = MX_WORKLOADS[workload]
            text = (f"MXFP4 matmul (f4E2M1FN, f8E8M0FNU scale / 32 elements) M={M} N={N} K={K} "
                    f"-> f32, tcgen05.mma.kind::mxf4.block_scale (scale repack x2 + GEMM), per GPU")
        elif workload in MATMUL_WORKLOADS:
            op, et, batch, M, N, K = MATMUL_WORKLOADS[workload]
            name = f"linalg.{op}" + (" + bias[n] + relu + truncf (one dispatch)"
                                     if workload in FUSED_WORKLOADS else "")
            text = f"{name} {et} batch={batch} M={M} N={N} K={K} per GPU"
        elif workload in ATTENTION_WORKLOADS:
            op, et, B, M, K2, D = ATTENTION_WORKLOADS[workload]
            mask = ("causal mask (k2 <= m + K2 - M) computed in the kernel"
                    if workload in CAUSAL_WORKLOADS else "no mask")
            text = (f"iree_linalg_ext.attention {et} batch*heads={B} M={M} K2={K2} "
                    f"K1=N={D} scale=1/sqrt({D}) {mask}, per GPU")
        else:
            text = "Llama-3-8B decoder-layer GEMMs (QKV/O/gate+up/down), tokens=8192"
    else:
        text = info["workload"]
    if world > 1 and workload != "tp_decoder_layer":
        text += f", batch-sharded over {world} GPUs (no collective)"
    return text


def cpu_work(workload, rows):
    """A bounded sample of |workload| for the CPU arm: returns (callable, flop, description).

    matmul: the first |rows| rows of the LHS against the full RHS. attention: the first
    |rows| query rows of one head against that head's full K/V. Inputs come from the
    reference generator (test_utils.c:359-420) / a seeded PCG as in the parity tests."""
    import numpy as np
    import oracle
    if workload in MX_WORKLOADS:
        _, _, _, M, N, K = MX_WORKLOADS[workload]
        rows = min(rows // 16, M)     # the scalar port decodes every element: a small slab
        lhs = oracle.generate_random_matrix("u8", (rows, K // 2), 1)
        ls = oracle.generate_random_matrix("f8e8m0fnu", (rows, K // 32), 2)
        rhs = oracle.generate_random_matrix("u8", (N, K // 2), 3)
        rs = oracle.generate_random_matrix("f8e8m0fnu", (N, K // 32), 4)
        desc = (f"oracle port of the MX reference (generate_code_mx.py:153-172), OpenMP, "
                f"M={rows} (of {M}) N={N} K={K}")
        return (lambda: oracle.matmul_mxfp4(lhs, ls, rhs, rs)), 2.0 * rows * N * K, desc, "port"
    if workload in MATMUL_WORKLOADS or workload == "tp_decoder_layer":
        # The TP layer's CPU sample is its QKV projection (same loops, same dtype).
        op, et, batch, M, N, K = MATMUL_WORKLOADS.get(
            workload, ("matmul", "bf16", 1, TpLayerStep.TOKENS, TpLayerStep.QKV, TpLayerStep.HIDDEN))
        rows = min(rows, M)
        lhs = oracle.generate_random_matrix(et, (rows, K), 1)
        rhs = oracle.generate_random_matrix(et, (K, N), 2)
        if workload in FUSED_WORKLOADS:
            bias = oracle.generate_random_matrix("f32", (N,), 4)
            mm = ((lambda: oracle.ref_matmul(et, lhs, rhs)) if oracle.ref_available()
                  else (lambda: oracle.matmul(et, lhs, rhs)))
            desc = (("reference ukernel path iree_uk_pack+mmt4d+unpack" if oracle.ref_available()
                     else "oracle port of the reference loops")
                    + f", then numpy bias + relu + truncf, {et} M={rows} (of {batch}x{M}) N={N} K={K}")
            return ((lambda: oracle.from_f32(et, np.maximum(mm() + bias, np.float32(0)))),
                    2.0 * rows * N * K, desc, "reference" if oracle.ref_available() else "port")
        if oracle.ref_available() and not et.startswith("f8"):
            # The reference's own CPU kernels (oracle/_ref, compiled from /root/reference).
            desc = (f"reference ukernel path iree_uk_pack+mmt4d+unpack (not the full llvm-cpu "
                    f"pipeline), OpenMP over M tiles, {et} M={rows} (of {batch}x{M}) N={N} K={K}")
            return (lambda: oracle.ref_matmul(et, lhs, rhs)), 2.0 * rows * N * K, desc, "reference"
        desc = (f"oracle port of the reference loops (iree-e2e-matmul-test.cc:28-121), OpenMP, "
                f"{et} M={rows} (of {batch}x{M}) N={N} K={K}")
        return (lambda: oracle.matmul(et, lhs, rhs)), 2.0 * rows * N * K, desc, "port"
    op, et, B, M, K2, D = ATTENTION_WORKLOADS[workload]
    rows = min(rows, M)
    heads = 1 if M > 1 else min(B, 8)
    g = np.random.Generator(np.random.PCG64(0xB200))
    mk = lambda *shape: oracle.from_f32(et, np.clip(g.standard_normal(shape), -4, 4).astype(np.float32))
    q, k, v = mk(heads, rows, D), mk(heads, K2, D), mk(heads, K2, D)
    scale = 1.0 / (D ** 0.5)
    desc = (f"oracle port of the llvm-cpu online-attention arithmetic "
            f"(AggregatedOpInterfaceImpl.cpp:323-632), OpenMP, {et} heads={heads} (of {B}) "
            f"M={rows} (of {M}) K2={K2} D={D}")
    if workload in CAUSAL_WORKLOADS:
        # The LAST |rows| query rows (the ones with the most unmasked keys), explicit i1 mask:
        # the oracle, like the reference's decomposition, evaluates masked entries too, so its
        # FLOP are counted like the GPU's (unmasked pairs) to keep the two arms comparable.
        first = M - rows
        mask = (np.arange(K2)[None, :] <= (first + np.arange(rows))[:, None] + (K2 - M))[None]
        mask = np.broadcast_to(mask, (heads, rows, K2)).copy()
        return ((lambda: oracle.attention(et, q, k, v, scale, mask=mask, k2_tile=128)),
                4.0 * heads * D * float(mask[0].sum()), desc + ", last rows, explicit causal i1 mask",
                "port")
    return ((lambda: oracle.attention(et, q, k, v, scale, k2_tile=128)),
            4.0 * heads * rows * K2 * D, desc, "port")


def oracle_ref() -> bool:
    import oracle
    return oracle.ref_available()


def use_all_host_threads() -> int:
    """Makes the OpenMP runtime behind the CPU arms (oracle/liboracle.so, oracle/_ref) use every
    core this process may run on, and returns that count. torchrun exports OMP_NUM_THREADS=1
    to its workers, which would otherwise turn the N>1 reference arm into a 1-thread run."""
    import ctypes
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(n)
    try:
        gomp = ctypes.CDLL("libgomp.so.1")
        gomp.omp_set_num_threads(ctypes.c_int(n))
        n = int(gomp.omp_get_max_threads())
    except OSError:
        pass
    return n


# Both CPU arms (cpu_baseline in the own arm's line, and --impl reference) time the SAME
# bounded sample: the first CPU_SAMPLE_ROWS rows of the LHS / of one head's queries against
# the full RHS / K,V. Each call includes the reference path's packing of both operands
# (oracle/ref_glue.c: pack -> mmt4d -> unpack, as tools/e2e_matmul_benchmark.c does).
CPU_SAMPLE_ROWS = 2048


def cpu_baseline(workload, threads):
    """The oracle port (oracle/oracle.c, OpenMP) timed on this box's host cores."""
    fn, flop, desc, kind = cpu_work(workload, rows=CPU_SAMPLE_ROWS)
    fn()  # warm the thread pool / page in
    t0 = time.perf_counter()
    runs = 0
    while runs < 1 or (time.perf_counter() - t0 < 5.0 and runs < 20):
        fn()
        runs += 1
    dt = (time.perf_counter() - t0) / runs
    return {"value": flop / dt / 1e12, "unit": "TFLOP/s", "cores": threads, "kind": kind,
            "sample": f"{desc}, {runs} run(s), {dt:.2f} s each"}


CPU_DTYPES = {"bf16": "bf16 (f32 accumulate)", "f16": "f16 (f32 accumulate)",
              "i8": "i8 (i32 accumulate)", "f8e4m3fn": "f8E4M3FN (f32 accumulate)",
              "f8e5m2": "f8E5M2 (f32 accumulate)"}


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    threads = use_all_host_threads()
    fn, flop, desc, kind = cpu_work(args.workload, rows=CPU_SAMPLE_ROWS)
    for _ in range(max(1, min(args.warmup, 2))):
        fn()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn()
    dt = (time.perf_counter() - t0) / args.steps
    value = flop / dt / 1e12
    spec = (MATMUL_WORKLOADS.get(args.workload) or ATTENTION_WORKLOADS.get(args.workload) or
            MX_WORKLOADS.get(args.workload) or
            ("matmul", "bf16", 1, TpLayerStep.TOKENS, TpLayerStep.QKV, TpLayerStep.HIDDEN))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "TFLOP/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "strong" if args.workload == "tp_decoder_layer" else "weak",
            "vs_baseline": None, "dtype": CPU_DTYPES.get(spec[1], spec[1]), "data": "synthetic",
            "config": {"workload": workload_string(args.workload, world),
                       "sample": desc + " per step"},
            "cpu_baseline": {"value": value, "unit": "TFLOP/s", "cores": threads, "kind": kind,
                             "sample": desc + " per step"},
            "e2e": {"value": value, "unit": "TFLOP/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser(description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="matmul_bf16", choices=sorted(list(MATMUL_WORKLOADS) + list(ATTENTION_WORKLOADS) + list(MX_WORKLOADS)) + ["tp_decoder_layer"])
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--reserved-sms", type=int, default=8,
                    help="tp_decoder_layer, N>1: SMs the GEMM grids leave to the NCCL kernels")
    ap.add_argument("--collective", default="symmetric", choices=["symmetric", "nccl"],
                    help="tp_decoder_layer, N>1: all-reduce through the symmetric-region NVLink "
                         "kernels (default) or through NCCL")
    ap.add_argument("--attention-flags", type=lambda v: int(v, 0), default=0,
                    help="IREE_B200_ATTENTION_FLAG_* bits for the attention workloads")
    ap.add_argument("--tp-schedule", default="dataflow", choices=["dataflow", "staged", "fused"],
                    help="tp_decoder_layer, N>1: dataflow = dependencies derived from binding ranges "
                         "(overlap_collectives 3); staged = barrier-separated stages, barriers join "
                         "the streams (2); fused = GEMM chunk, all-reduce, ... with one persistent "
                         "all-reduce launch per row-parallel GEMM (1)")
    ap.add_argument("--extra-min-ms", type=float, default=200.0,
                    help="minimum timed region of each `also` / `sharded` line")
    ap.add_argument("--extra-soak-ms", type=float, default=1000.0,
                    help="untimed soak of the same work before every extra line's timed region")
    ap.add_argument("--presoak-s", type=float, default=4.0,
                    help="seconds of untimed tensor-bound work before the first extra line "
                         "(burns the boost budget a B200 accumulates while idle)")
    ap.add_argument("--no-extra", action="store_true",
                    help="skip the `also` (N=1) / `sharded` (N>1) lines of the default workload")
    ap.add_argument("--ar-blocks", type=int, default=0,
                    help="tp_decoder_layer: CTAs per symmetric all-reduce (0 = the reserved SMs)")
    ap.add_argument("--tp-chunk-rows", default="",
                    help="tp_decoder_layer: explicit token rows per chunk, e.g. 1024,3584,3584")
    ap.add_argument("--tp-chunks", type=int, default=0,
                    help="tp_decoder_layer: equal token chunks per row-parallel GEMM (0 = auto)")
    args = ap.parse_args()
    global SOAK_MS
    SOAK_MS = max(0.0, args.extra_soak_ms)
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
        return
    try:
        run_ours(args)
    finally:
        # Every rank tears the process group down (rank 0 alone prints the line).
        if world > 1:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.destroy_process_group()


if __name__ == "__main__":
    main()
