"""Multi-GPU check + timing of the symmetric-region all-reduce (run under torchrun):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29541 tools/symmetric_check.py [--mib 64]

Parity: in-place all_reduce(sum) recorded through iree_b200_command_buffer_collective on a
channel with an attached symmetric region (NVLS kernels, then the peer load/store kernels)
against the exact expected sum (small-integer inputs, so every summation order gives the
same bits) — f16 / bf16 / f32, ragged sizes, sub-ranges, back-to-back launches (flag reuse).
Timing: the same collective through NCCL (a channel with no region) and through both kernel
families, CUDA events, max over ranks. Rank 0 prints one JSON line.
"""

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=64)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--max-blocks", type=int, default=0, help="CTAs per collective (0 = default 16)")
    ap.add_argument("--sweep", default="",
                    help="timing only: ';'-separated 'blocks,threads,unroll,pipelined' variants of "
                         "the NVLS kernel (IREE_B200_ALL_REDUCE_TUNING per channel)")
    args = ap.parse_args()
    os.environ.setdefault("IREE_B200_COLLECTIVE_WATCHDOG_MS", "30000")
    import torch
    import torch.distributed as dist
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from iree_b200 import hal, symmetric

    stream = torch.cuda.Stream()
    dev = hal.Device(local, mode=hal.MODE_STREAM, external_stream=stream.cuda_stream,
                     overlap_collectives=True)
    ops = hal.Ops(dev)

    def make_channel():
        ids = [hal.Channel.create_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        return hal.Channel(dev, rank=rank, count=world, id_bytes=ids[0])

    data_bytes = args.mib << 20
    if args.sweep:
        return sweep(args, torch, dist, hal, symmetric, dev, ops, stream, make_channel, rank, world,
                     local, data_bytes)
    ch_nvls, ch_p2p, ch_nccl = make_channel(), make_channel(), make_channel()
    region = symmetric.SymmetricRegion(data_bytes, torch.device("cuda", local), flag_blocks=2)
    have_mc = region.multicast_base != 0
    if have_mc:
        region.attach(ch_nvls, ops.executable, flag_block=0, use_multicast=True,
                      max_blocks=args.max_blocks)
    region.attach(ch_p2p, ops.executable, flag_block=1, use_multicast=False,
                  max_blocks=args.max_blocks)
    buf = dev.import_buffer(region.local_base, region.data_bytes)
    plain = torch.empty(data_bytes, dtype=torch.uint8, device="cuda")
    plain_buf = dev.import_buffer(plain.data_ptr(), data_bytes)

    def all_reduce(channel, buffer, et, byte_offset, count, repeat=1):
        es = 4 if et == "f32" else 2
        op = hal.collective_op(hal.ALL_REDUCE, hal.REDUCTION_SUM, et)
        cb = hal.CommandBuffer(dev).begin()
        for _ in range(repeat):
            ref = buffer.ref(byte_offset, count * es)
            cb.collective(channel, op, 0, ref, ref, count)
            cb.barrier()
        cb.end()
        return cb

    failures = []
    dtypes = {"f16": torch.float16, "bf16": torch.bfloat16, "f32": torch.float32}
    families = ([("nvls", ch_nvls)] if have_mc else []) + [("p2p", ch_p2p)]
    with torch.cuda.stream(stream):
        for fam, ch in families:
            for et, dt in dtypes.items():
                es = dt.itemsize
                for byte_offset, count in ((0, 8 * 1024), (4096, (1 << 20) + 8 * 3),
                                           (16, 8), (1 << 20, (3 << 20) // es)):
                    if byte_offset + count * es > data_bytes:
                        continue
                    view = region.view(byte_offset, (count,), dt)
                    idx = torch.arange(count, device="cuda")
                    # Values in [-3, 3] * (rank + 1): sums stay small integers -> exact.
                    mine = (((idx * 7 + 3) % 7) - 3).to(torch.float32) * (rank + 1)
                    view.copy_(mine.to(dt))
                    guard_lo = region.view(0, (byte_offset // es,), dt).clone() if byte_offset else None
                    dist.barrier()
                    before = ch.collective_counts()["symmetric"]
                    dev.execute(all_reduce(ch, buf, et, byte_offset, count))
                    dev.wait_idle()
                    want = (((idx * 7 + 3) % 7) - 3).to(torch.float32) * (world * (world + 1) // 2)
                    ok = torch.equal(view.to(torch.float32), want)
                    ok = ok and ch.collective_counts()["symmetric"] == before + 1
                    if guard_lo is not None:
                        ok = ok and torch.equal(region.view(0, (byte_offset // es,), dt), guard_lo)
                    if not ok:
                        failures.append(f"{fam} {et} offset {byte_offset} count {count}")
                    dist.barrier()
            # Back-to-back launches reuse the flag words: x -> W*x -> W*W*x.
            view = region.view(0, (1 << 16,), torch.float32)
            view.fill_(1.0)
            dist.barrier()
            dev.execute(all_reduce(ch, buf, "f32", 0, 1 << 16, repeat=3))
            dev.wait_idle()
            if not bool((view == float(world ** 3)).all()):
                failures.append(f"{fam} back-to-back")
            dist.barrier()
            # Range fusion: producer (fill), all-reduce, producer, all-reduce, ... between two
            # barriers becomes one persistent launch; range c is released behind producer c.
            import struct
            chunks, chunk_elems = 4, 1 << 20
            region.tensor[:chunks * chunk_elems * 4].zero_()
            dist.barrier()
            op = hal.collective_op(hal.ALL_REDUCE, hal.REDUCTION_SUM, "f32")
            cb = hal.CommandBuffer(dev).begin()
            for c in range(chunks):
                ref = buf.ref(c * chunk_elems * 4, chunk_elems * 4)
                cb.fill(ref, struct.pack("<f", float((rank + 1) * (c + 1))))
                cb.collective(ch, op, 0, ref, ref, chunk_elems)
            cb.barrier()
            cb.end()
            for rep in range(2):   # twice: the ready word keeps counting up across replays
                before = ch.collective_counts()["symmetric"]
                dev.execute(cb)
                dev.wait_idle()
                got = region.view(0, (chunks, chunk_elems), torch.float32)
                want = torch.arange(1, chunks + 1, device="cuda", dtype=torch.float32) \
                    * (world * (world + 1) // 2)
                if not (torch.equal(got, want[:, None].expand(chunks, chunk_elems)) and
                        ch.collective_counts()["symmetric"] == before + chunks):
                    failures.append(f"{fam} fused ranges (replay {rep})")
                dist.barrier()
        # A range outside the region (or an unaligned one) must fall back to NCCL.
        t = torch.ones(1024, dtype=torch.float32, device="cuda")
        tb = dev.import_buffer(t.data_ptr(), 4096)
        before = ch_p2p.collective_counts()
        dev.execute(all_reduce(ch_p2p, tb, "f32", 0, 1024))
        dev.wait_idle()
        after = ch_p2p.collective_counts()
        if not (bool((t == float(world)).all()) and after["nccl"] == before["nccl"] + 1):
            failures.append("fallback to NCCL outside the region")

        # ---- timing ----
        def time_it(channel, buffer, nbytes=data_bytes):
            count = nbytes // 2
            cb = all_reduce(channel, buffer, "bf16", 0, count)
            for _ in range(3):
                dev.execute(cb)
            dev.wait_idle()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(args.iters):
                dev.execute(cb)
            e1.record(stream)
            dev.wait_idle()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / args.iters], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())

        region.tensor.zero_()
        plain.zero_()
        times = {"nccl": time_it(ch_nccl, plain_buf)}
        if have_mc:
            times["nvls"] = time_it(ch_nvls, buf)
        times["p2p"] = time_it(ch_p2p, buf)
        # The tensor-parallel layer's chunk size ([2048, 4096] bf16).
        chunk = min(16 << 20, data_bytes)
        times_chunk = {"nccl": time_it(ch_nccl, plain_buf, chunk)}
        if have_mc:
            times_chunk["nvls"] = time_it(ch_nvls, buf, chunk)
        times_chunk["p2p"] = time_it(ch_p2p, buf, chunk)
    flag_ok = True
    flags = region.tensor[region.data_bytes:]
    torch.cuda.synchronize()
    dist.barrier()
    flag_ok = bool((flags == 0).all())
    if not flag_ok:
        failures.append("flag words not back to zero")
    bad = torch.tensor([len(failures)], device="cuda")
    dist.all_reduce(bad)
    if rank == 0:
        print(json.dumps({
            "world": world, "multicast": have_mc, "failures": failures,
            "failing_ranks_total": int(bad.item()), "all_reduce_mib": args.mib, "dtype": "bf16",
            "ms": times,
            "algbw_gbs": {k: data_bytes / (v * 1e-3) / 1e9 for k, v in times.items()},
            "max_blocks": args.max_blocks or 16,
            "ms_16mib": times_chunk}), flush=True)
    else:
        for f in failures:
            print(f"rank {rank}: FAILED {f}", flush=True)
    dist.destroy_process_group()
    sys.exit(1 if failures else 0)


def sweep(args, torch, dist, hal, symmetric, dev, ops, stream, make_channel, rank, world, local,
          data_bytes):
    variants = [tuple(int(x) for x in v.split(",")) for v in args.sweep.split(";") if v]
    region = symmetric.SymmetricRegion(data_bytes, torch.device("cuda", local),
                                       flag_blocks=len(variants))
    assert region.multicast_base, "the sweep is of the NVLS kernel"
    buf = dev.import_buffer(region.local_base, region.data_bytes)
    region.tensor[:data_bytes].zero_()
    op = hal.collective_op(hal.ALL_REDUCE, hal.REDUCTION_SUM, "bf16")
    results = []
    with torch.cuda.stream(stream):
        for i, (blocks, threads, unroll, pipelined) in enumerate(variants):
            os.environ["IREE_B200_ALL_REDUCE_TUNING"] = f"{threads},{unroll},{pipelined}"
            ch = make_channel()
            region.attach(ch, ops.executable, flag_block=i, use_multicast=True, max_blocks=blocks)
            row = {"blocks": blocks, "threads": threads, "unroll": unroll, "pipelined": pipelined}
            # Parity first: x -> W*x on a ragged size.
            count = (1 << 20) + 8 * 3
            view = region.view(0, (count,), torch.bfloat16)
            view.fill_(1.0)
            dist.barrier()
            cb = hal.CommandBuffer(dev).begin()
            cb.collective(ch, op, 0, buf.ref(0, count * 2), buf.ref(0, count * 2), count)
            cb.barrier()
            cb.end()
            dev.execute(cb)
            dev.wait_idle()
            row["exact"] = bool((view == float(world)).all()) and ch.collective_status() == 0
            region.tensor[:data_bytes].zero_()
            for nbytes, key in ((data_bytes, "ms"), (min(16 << 20, data_bytes), "ms_16mib")):
                count = nbytes // 2
                cb = hal.CommandBuffer(dev).begin()
                cb.collective(ch, op, 0, buf.ref(0, nbytes), buf.ref(0, nbytes), count)
                cb.barrier()
                cb.end()
                for _ in range(3):
                    dev.execute(cb)
                dev.wait_idle()
                dist.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(args.iters):
                    dev.execute(cb)
                e1.record(stream)
                dev.wait_idle()
                torch.cuda.synchronize()
                t = torch.tensor([e0.elapsed_time(e1) / args.iters], device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                row[key] = float(t.item())
            row["algbw_gbs"] = data_bytes / (row["ms"] * 1e-3) / 1e9
            results.append(row)
    if rank == 0:
        print(json.dumps({"world": world, "all_reduce_mib": args.mib, "dtype": "bf16",
                          "sweep": results}), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if all(r["exact"] for r in results) else 1)


if __name__ == "__main__":
    main()
