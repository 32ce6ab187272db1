"""The GEMMs of one Llama-style decoder layer, tensor-parallel over the ranks of one box
(BASELINE.json configs[4]): what an IREE module sharded the Megatron
way records per rank —

    qkv = x · Wqkv[:, slice]                  column-parallel, no exchange
    o   = all_reduce_sum(attn[:, slice] · Wo[slice, :])      row-parallel
    gu  = o · Wgu[:, slice]                   column-parallel
    d   = all_reduce_sum(act[:, slice] · Wd[slice, :])       row-parallel

— as ONE command buffer through the C-ABI: `iree_b200_ops_record_matmul` dispatches and
`iree_b200_command_buffer_collective(ALL_REDUCE, SUM)` on an `iree_b200_channel_t`
(flow.collective.all_reduce → hal.command_buffer.collective → nccl_channel.c:360-380 in the
reference). Row-parallel GEMMs are recorded in token chunks, each followed by the
all-reduce of its rows, and the column-parallel consumer (gate/up) is recorded over the
same chunks: rows are independent, so chunk c of `gu` needs only chunk c of `o`, and the
launch layer (command_buffer.cc) lets it start while later chunks are still being reduced.

Host-side orchestration only (shapes, slices, recording order); shared by bench.py and the
multi-rank parity tests. `attn` and `act` are inputs (the attention and the SiLU·mul between
the GEMMs are other dispatches).
"""

from __future__ import annotations

from . import sharding


class DecoderLayerGemms:
    def __init__(self, torch, hal, dev, ops, world, rank, channel, *, tokens=8192, hidden=4096,
                 q_heads=32, kv_heads=8, head_dim=128, ffn=14336, chunks=4, region=None,
                 init=None, seed=0, chunk_consumers=True, only=None, schedule="staged"):
        """|init(name, shape) -> torch tensor (bf16, on the device)| fills an operand with the
        FULL (unsharded) shape; the rank's slice is taken here. Default: N(0, 0.05) noise.
        All GEMMs use the truncating epilogue: bf16 activations, bf16 all-reduce (f32
        accumulate), as a bf16 model's module does. |only| = "gemms" / "collectives" records
        the same schedule with the other kind left out (the bench's breakdown of a step).
        |schedule| (with chunk_consumers): "staged" = barrier-separated stages balanced by
        list scheduling (for a launch layer whose barrier joins its streams:
        overlap_collectives 2); "dataflow" = every command its own barrier group, in the order
        each in-order stream should see them (for overlap_collectives 3, where the launch
        layer derives the true dependencies from the binding ranges)."""
        T, H, FFN = tokens, hidden, ffn
        QKV = (q_heads + 2 * kv_heads) * head_dim
        plan = {g.name: g for g in sharding.decoder_layer_plan(
            world, rank, tokens=T, hidden=H, q_heads=q_heads, kv_heads=kv_heads,
            head_dim=head_dim, ffn=FFN)}
        self.plan = plan
        self.dims = dict(tokens=T, hidden=H, qkv=QKV, ffn=FFN)
        self.world, self.rank, self.dev, self.channel, self.region = world, rank, dev, channel, region
        self.chunks = chunks
        out_dt, es_out = torch.bfloat16, 2
        if init is None:
            g = torch.Generator(device="cuda")
            g.manual_seed(0xB200 + seed)
            init = lambda name, shape: (torch.randn(shape, device="cuda", generator=g) * 0.05
                                        ).to(torch.bfloat16)
        qo, ho, fo = plan["qkv"].n_offset, plan["o"].k_offset, plan["down"].k_offset
        qn, hn, fn = plan["qkv"].N, plan["o"].K, plan["down"].K
        full = {
            "x": (T, H), "wqkv": (H, QKV), "attn": (T, H), "wo": (H, H),
            "wgu": (H, 2 * FFN), "act": (T, FFN), "wd": (FFN, H),
        }
        take = {
            "x": lambda t: t, "wqkv": lambda t: t[:, qo:qo + qn], "attn": lambda t: t[:, ho:ho + hn],
            "wo": lambda t: t[ho:ho + hn], "wgu": lambda t: t[:, 2 * fo:2 * fo + 2 * fn],
            "act": lambda t: t[:, fo:fo + fn], "wd": lambda t: t[fo:fo + fn],
        }
        self.t = {k: take[k](init(k, shape)).contiguous() for k, shape in full.items()}
        out = lambda *shape: torch.empty(shape, dtype=out_dt, device="cuda")
        self.t.update({"qkv": out(T, qn), "o": out(T, H), "gu": out(T, 2 * fn), "d": out(T, H)})
        if region is not None:
            # The all-reduced activations live in the symmetric region (every rank maps it):
            # their all-reduces run on the NVLink kernels instead of NCCL.
            self.t["o"] = region.view(0, (T, H), out_dt)
            self.t["d"] = region.view(T * H * es_out, (T, H), out_dt)
        self.b = {k: dev.import_buffer(v.data_ptr(), v.numel() * v.element_size())
                  for k, v in self.t.items()}
        self.flop = 2.0 * T * (H * QKV + H * H + H * 2 * FFN + FFN * H) / world
        flags = hal.MATMUL_TRUNCATE_RESULT
        ar = hal.collective_op(hal.ALL_REDUCE, hal.REDUCTION_SUM, "bf16")
        rows_of = sharding.token_chunks(T, chunks)
        chunks = len(rows_of)
        self.chunks = chunks
        before = ops.dispatch_count
        b = self.b
        self.collectives = 0

        def gemm(lhs, rhs, dst, n, k, chunk=None):
            r0, rows = (0, T) if chunk is None else rows_of[chunk]
            return lambda cb: ops.record_matmul(
                cb, "bf16", 1, rows, n, k, b[lhs].ref(r0 * k * 2, rows * k * 2), b[rhs].ref(),
                b[dst].ref(r0 * n * es_out, rows * n * es_out), flags=flags)

        def all_reduce(dst, chunk):
            r0, rows = rows_of[chunk]
            ref = b[dst].ref(r0 * H * es_out, rows * H * es_out)

            def record(cb):
                cb.collective(channel, ar, 0, ref, ref, rows * H)
                self.collectives += 1
            return record

        # Tasks: name -> (kind, relative cost, dependencies, recorder). Costs are GEMM flop in
        # units of one `o` chunk (5.8 us at TP-8 with 4 chunks, shrinking with 1/world); the
        # all-reduce of a chunk is link time, independent of the TP degree: ~8 such units at
        # TP-8 (47 us measured), i.e. about |world| units.
        ar_cost = float(world)
        tasks = {"qkv": ("gemm", H * qn / (H * hn), [], gemm("x", "wqkv", "qkv", qn, H))}
        single = world == 1 or not chunk_consumers
        for c in range(chunks):
            tasks[f"o{c}"] = ("gemm", 1.0, [], gemm("attn", "wo", "o", H, hn, c))
            if world > 1:
                tasks[f"ar_o{c}"] = ("collective", ar_cost, [f"o{c}"], all_reduce("o", c))
            o_done = f"ar_o{c}" if world > 1 else f"o{c}"
            if not single:
                tasks[f"gu{c}"] = ("gemm", H * 2 * fn / (H * hn), [o_done],
                                   gemm("o", "wgu", "gu", 2 * fn, H, c))
            gu_done = f"gu{c}" if not single else "gu"
            tasks[f"d{c}"] = ("gemm", fn * H / (H * hn), [gu_done], gemm("act", "wd", "d", H, fn, c))
            if world > 1:
                tasks[f"ar_d{c}"] = ("collective", ar_cost, [f"d{c}"], all_reduce("d", c))
        if single:
            o_all = [(f"ar_o{c}" if world > 1 else f"o{c}") for c in range(chunks)]
            tasks["gu"] = ("gemm", chunks * H * 2 * fn / (H * hn), o_all,
                           gemm("o", "wgu", "gu", 2 * fn, H))
        if schedule == "dataflow" and chunk_consumers and world > 1:
            # GEMM stream: the producers of the first reduce, then chunk by chunk what each
            # finished reduce unblocks; collective stream: the reduces in the order their
            # consumers need them. One command per barrier group: the launch layer orders
            # exactly the pairs that share memory.
            order = [f"o{c}" for c in range(chunks)] + ["qkv"]
            for c in range(chunks):
                order += [f"ar_o{c}"]
            for c in range(chunks):
                order += [f"gu{c}", f"d{c}", f"ar_d{c}"]
            assert sorted(order) == sorted(tasks)
            self.stages = [[n] for n in order]
        else:
            self.stages = self.schedule(tasks, software_pipeline=chunk_consumers and world > 1)
        cb = hal.CommandBuffer(dev).begin()
        for stage in self.stages:
            # The collective first: the launch layer issues it on the collective stream behind
            # everything already queued, and the stage's GEMMs then run beside it.
            for name in sorted(stage, key=lambda n: tasks[n][0] != "collective"):
                if only is None or only == tasks[name][0] + "s":
                    tasks[name][3](cb)
            cb.barrier()
        cb.end()
        self.cb = cb
        self.launches_per_step = ops.dispatch_count - before

    @staticmethod
    def schedule(tasks, software_pipeline=True):
        """Stages separated by execution barriers (the only ordering a HAL command buffer
        has): every task of a stage may run concurrently, and depends only on tasks of earlier
        stages. Greedy list scheduling in dependency order: at most one all-reduce per stage,
        beside it as much independent GEMM work as about covers it — so the reduce of chunk c
        runs under the GEMMs of the chunks around it instead of on the critical path.
        |software_pipeline| False gives the plain sequence (one stage per dependency level of
        the original recording order)."""
        done, stages = set(), []
        order = list(tasks)
        if not software_pipeline:
            return [[n] for n in order]   # recording order is already a valid sequence
        # Priority = longest path to the end of the layer (critical-path list scheduling).
        users = {n: [m for m in order if n in tasks[m][2]] for n in order}
        rank = {}
        for n in reversed(order):
            rank[n] = tasks[n][1] + max((rank[m] for m in users[n]), default=0.0)
        while len(done) < len(tasks):
            ready = sorted((n for n in order if n not in done and all(d in done for d in tasks[n][2])),
                           key=lambda n: -rank[n])
            assert ready, "dependency cycle"
            stage = []
            collectives = [n for n in ready if tasks[n][0] == "collective"]
            budget = float("inf")
            if collectives:
                stage.append(collectives[0])
                budget = 1.4 * tasks[collectives[0]][1]
            spent = 0.0
            for n in (n for n in ready if tasks[n][0] == "gemm"):
                cost = tasks[n][1]
                if spent and spent + cost > budget:
                    continue
                stage.append(n)
                spent += cost
            stages.append(stage)
            done.update(stage)
        return stages

    def step(self):
        self.dev.execute(self.cb)

    def collective_path(self):
        if self.world == 1:
            return "none at N=1"
        if self.region is None:
            return "NCCL"
        counts = self.channel.collective_counts()
        kind = "NVSwitch multimem.ld_reduce/st kernels" if self.region.multicast_base else \
            "peer load/store kernels"
        return (f"symmetric region, {kind}: {counts['symmetric']} issued by our kernels, "
                f"{counts['nccl']} by NCCL so far")
