"""World-size-2 CPU (gloo) tests of the N>1 host logic: sharding arithmetic, the channel
provider's rank discovery / id exchange, and that sharded execution + all-reduce reproduces
the unsharded op (with the oracle standing in for the device kernels)."""

import os
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from iree_b200 import hal, sharding


def test_contiguous_shard_covers_exactly():
    for extent in (0, 1, 7, 64, 4097):
        for world in (1, 2, 3, 8):
            spans = [sharding.contiguous_shard(extent, world, r) for r in range(world)]
            assert spans[0][0] == 0 and sum(l for _, l in spans) == extent
            for (o0, l0), (o1, _) in zip(spans, spans[1:]):
                assert o0 + l0 == o1
    with pytest.raises(ValueError):
        sharding.contiguous_shard(8, 2, 2)


def test_decoder_layer_plan_llama3_8b():
    total = 0.0
    for rank in range(8):
        plan = sharding.decoder_layer_plan(8, rank)
        assert [(g.name, g.M, g.N, g.K, g.parallel) for g in plan] == [
            ("qkv", 8192, 768, 4096, "column"), ("o", 8192, 4096, 512, "row"),
            ("gate_up", 8192, 3584, 4096, "column"), ("down", 8192, 4096, 1792, "row")]
        total += sum(g.flop for g in plan)
    assert abs(total - 3573.4e9) / 3573.4e9 < 1e-4       # 3573.4 GFLOP per layer
    with pytest.raises(ValueError):
        sharding.decoder_layer_plan(5, 0)
    assert sharding.token_chunks(8192, 4) == [(0, 2048), (2048, 2048), (4096, 2048), (6144, 2048)]


def _worker(rank, world, port, tmp, results):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), IREE_B200_CHANNEL_ID_FILE=os.path.join(tmp, "id"))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. default channel provider: rank/count from the launcher's environment.
        assert hal.default_rank_and_count() == (rank, world)
        # 2. id exchange without MPI: rank 0 publishes 128 bytes, the others receive them.
        ident = bytes(range(128)) if rank == 0 else bytes(128)
        if rank != 0:
            # Let rank 0 get there first: it then finds only the planted leftovers (a stale
            # hello AND a matching stale ack of "rank 1") and must keep waiting for the real
            # rank 1 — acknowledgements echo rank 0's per-exchange nonce.
            import time
            time.sleep(1.0)
        assert hal.exchange_default_id(rank, ident) == bytes(range(128))
        # ... a second channel in the same run gets ITS id (per-process sequence number), and
        # neither exchange is fooled by the stale files the test planted (a leftover
        # id file must never be taken for the current id).
        ident2 = bytes(reversed(range(128))) if rank == 0 else bytes(128)
        assert hal.exchange_default_id(rank, ident2) == bytes(reversed(range(128)))
        if rank == 0:
            left = [f for f in os.listdir(tmp) if f.startswith("id.") and ".tmp." not in f
                    and not f.startswith("id.7.")]
            assert left == [], left   # rank 0 removes the files of a finished exchange
        # 3. row-parallel GEMM + all_reduce(sum) == the full GEMM (oracle as the kernel).
        M, N, K = 48, 40, 64
        lhs = oracle.generate_random_matrix("bf16", (M, K), 1)
        rhs = oracle.generate_random_matrix("bf16", (K, N), 2)
        ko, kl = sharding.contiguous_shard(K, world, rank)
        part = oracle.matmul("bf16", np.ascontiguousarray(lhs[:, ko:ko + kl]),
                             np.ascontiguousarray(rhs[ko:ko + kl]))
        t = torch.from_numpy(part.copy())
        dist.all_reduce(t)
        full = oracle.matmul("bf16", lhs, rhs)
        assert np.array_equal(t.numpy(), full)            # exact: small-integer inputs
        gathered = [torch.empty_like(torch.from_numpy(part)) for _ in range(world)]
        dist.all_gather(gathered, torch.from_numpy(part.copy()))
        assert np.array_equal(oracle.all_reduce_sum_f32(np.stack([g.numpy() for g in gathered])),
                              full)
        # 4. column-parallel GEMM + all_gather == the full GEMM.
        no, nl = sharding.contiguous_shard(N, world, rank)
        col = oracle.matmul("bf16", lhs, np.ascontiguousarray(rhs[:, no:no + nl]))
        cols = [torch.empty((M, sharding.contiguous_shard(N, world, r)[1])) for r in range(world)]
        dist.all_gather(cols, torch.from_numpy(col.copy()))
        assert np.array_equal(np.concatenate([c.numpy() for c in cols], axis=1), full)
        # 5. batch-sharded batch_matmul: slices are independent.
        B = 5
        a = oracle.generate_random_matrix("i8", (B, 8, 16), 3)
        b = oracle.generate_random_matrix("i8", (B, 16, 8), 4)
        bo, bl = sharding.contiguous_shard(B, world, rank)
        mine = oracle.matmul("i8", a[bo:bo + bl], b[bo:bo + bl])
        assert np.array_equal(mine, oracle.matmul("i8", a, b)[bo:bo + bl])
        results[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo():
    world = 2
    with tempfile.TemporaryDirectory() as tmp, mp.Manager() as manager:
        results = manager.dict()
        import socket
        with socket.socket() as sock:      # a port the kernel says is free right now
            sock.bind(("127.0.0.1", 0))
            port = sock.getsockname()[1]
        # Leftovers of an "earlier run": a well-formed id file for exchanges #0 and #1 with
        # someone else's id and tokens, a stale hello and a stale ack of rank 1.
        import struct
        for seq in (0, 1, 7):
            stale = struct.pack("<4I", 0x31444942, seq, world, 0) + struct.pack("<2Q", 0, 0x1234567) \
                + bytes([0xEE]) * 128
            with open(os.path.join(tmp, f"id.{seq}.id"), "wb") as f:
                f.write(stale)
            with open(os.path.join(tmp, f"id.{seq}.hello.1"), "wb") as f:
                f.write(struct.pack("<Q", 0x1234567))
            with open(os.path.join(tmp, f"id.{seq}.ack.1"), "wb") as f:   # {token, rank 0's nonce}
                f.write(struct.pack("<2Q", 0x1234567, 0x7654321) if seq else struct.pack("<Q", 0x1234567))
        mp.spawn(_worker, args=(world, port, tmp, results), nprocs=world, join=True)
        assert dict(results) == {0: "ok", 1: "ok"}
