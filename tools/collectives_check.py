"""Every collective kind of iree_hal_collective_op_t through iree_b200_command_buffer_collective
on real ranks — what the reference's tests/e2e/collectives/collectives_test.py:46-93 checks per
rank count (all_reduce, all_gather, all_to_all, reduce_scatter, broadcast, reduce, send / recv,
send_recv), in stream mode and captured into a CUgraph. Run under torchrun:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29555 tools/collectives_check.py

Inputs are small integers (rank-dependent), so every reduction is exact in f32 / bf16 / i32.
Rank 0 prints one JSON line; exit status 1 on any mismatch on any rank.
"""

import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from iree_b200 import hal

    failures, ran = [], []
    n = 4096                                  # elements per rank
    stream = torch.cuda.Stream()
    dtypes = {"f32": torch.float32, "bf16": torch.bfloat16, "i32": torch.int32}
    with torch.cuda.stream(stream):
        for mode_name, mode in (("stream", hal.MODE_STREAM), ("graph", hal.MODE_GRAPH)):
            dev = hal.Device(local, mode=mode, external_stream=stream.cuda_stream)
            ids = [hal.Channel.create_id() if rank == 0 else None]
            dist.broadcast_object_list(ids, src=0)
            ch = hal.Channel(dev, rank=rank, count=world, id_bytes=ids[0])
            assert ch.rank_and_count() == (rank, world)

            def run(kind, et, send, recv, count, param=0, reduction=hal.REDUCTION_SUM):
                op = hal.collective_op(kind, reduction, et)
                keep = [None if t is None else dev.import_buffer(t.data_ptr(), t.numel() * t.element_size())
                        for t in (send, recv)]     # alive until recorded (the command buffer retains them)
                refs = [hal.NULL_REF if b is None else b.ref() for b in keep]
                cb = hal.CommandBuffer(dev).begin()
                cb.collective(ch, op, param, refs[0], refs[1], count)
                cb.barrier()
                cb.end()
                dist.barrier()
                dev.execute(cb)
                dev.execute(cb)        # twice: a graph replays, a stream command buffer re-issues
                dev.wait_idle()
                torch.cuda.synchronize()

            def check(label, got, want):
                ran.append(f"{mode_name}/{label}")
                if not torch.equal(got, want):
                    failures.append(f"{mode_name}/{label}: {int((got != want).sum())} of {got.numel()} differ")

            idx = torch.arange(n, device="cuda")
            for et, dt in dtypes.items():
                base = ((idx * 5 + 1) % 7 - 3).to(dt)            # values in [-3, 3]
                mine = base * (rank + 1)
                total = base * (world * (world + 1) // 2)
                # all_reduce(sum), out of place
                out = torch.zeros(n, dtype=dt, device="cuda")
                run(hal.ALL_REDUCE, et, mine, out, n)
                check(f"all_reduce sum {et}", out, total)
                # all_reduce(max): the largest rank's multiple where base > 0, rank 0's otherwise
                out.zero_()
                run(hal.ALL_REDUCE, et, mine, out, n, reduction=hal.REDUCTION_MAX)
                check(f"all_reduce max {et}", out, torch.where(base > 0, base * world, base))
                # all_gather: rank r's block lands at block r
                out = torch.zeros(n * world, dtype=dt, device="cuda")
                run(hal.ALL_GATHER, et, mine, out, n)
                want = torch.cat([base * (r + 1) for r in range(world)])
                check(f"all_gather {et}", out, want)
                # reduce_scatter: every rank contributes world blocks, gets the sum of its block
                send = torch.cat([base * (rank + 1) + b for b in range(world)]) if et != "bf16" else \
                    torch.cat([base * (rank + 1) for b in range(world)])
                out = torch.zeros(n, dtype=dt, device="cuda")
                run(hal.REDUCE_SCATTER, et, send, out, n)
                want = total + (rank * world if et != "bf16" else 0)
                check(f"reduce_scatter {et}", out, want.to(dt))
                # broadcast from the last rank
                out = torch.zeros(n, dtype=dt, device="cuda")
                run(hal.BROADCAST, et, mine, out, n, param=world - 1)
                check(f"broadcast {et}", out, base * world)
                # reduce to rank 0 (other ranks' recv is untouched)
                out = torch.full((n,), 9, dtype=dt, device="cuda")
                run(hal.REDUCE, et, mine, out, n, param=0)
                check(f"reduce {et}", out, total if rank == 0 else torch.full_like(out, 9))
                # all_to_all: block r of rank s goes to block s of rank r
                per = n // world
                send = torch.cat([torch.full((per,), (rank + 1) * 10 + r, device="cuda").to(dt)
                                  for r in range(world)])
                out = torch.zeros(per * world, dtype=dt, device="cuda")
                run(hal.ALL_TO_ALL, et, send, out, per * world)
                want = torch.cat([torch.full((per,), (s + 1) * 10 + rank, device="cuda").to(dt)
                                  for s in range(world)])
                check(f"all_to_all {et}", out, want)
            # send_recv ring: send to rank+1, receive from rank-1 (param: target | source << 16)
            dt = torch.float32
            mine = torch.full((n,), float(rank + 1), device="cuda")
            out = torch.zeros(n, device="cuda")
            nxt, prv = (rank + 1) % world, (rank - 1) % world
            run(hal.SEND_RECV, "f32", mine, out, n, param=(nxt & 0xffff) | (prv << 16))
            check("send_recv ring", out, torch.full((n,), float(prv + 1), device="cuda"))
            # send / recv pair: even ranks send to rank+1 (when it exists)
            if world >= 2 and not (world % 2 and rank == world - 1):
                if rank % 2 == 0:
                    run(hal.SEND, "f32", mine, None, n, param=rank + 1)
                else:
                    out.zero_()
                    run(hal.RECV, "f32", None, out, n, param=rank - 1)
                    check("send / recv", out, torch.full((n,), float(rank), device="cuda"))
            else:
                dist.barrier()
                dist.barrier()
            dist.barrier()
    bad = torch.tensor([len(failures)], device="cuda")
    dist.all_reduce(bad)
    if rank == 0:
        print(json.dumps({"world": world, "cases": len(ran), "failures": failures,
                          "failing_ranks_total": int(bad.item())}), flush=True)
    else:
        for f in failures:
            print(f"rank {rank}: FAILED {f}", flush=True)
    dist.destroy_process_group()
    sys.exit(1 if int(bad.item()) else 0)


if __name__ == "__main__":
    main()
