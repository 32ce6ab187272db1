"""Multi-GPU parity of the tensor-parallel decoder-layer GEMMs and of collectives issued
through iree_b200_channel_t — the per-rank-count checks of the reference's
tests/e2e/collectives/collectives_test.py (SingleRank / TwoRanks / FourRanks), for this path.
Run under torchrun:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29551 tools/tp_layer_check.py

Every rank builds the SAME unsharded operands with the reference generator
(test_utils.c:359-420: small integers), takes its Megatron slice and runs the layer
(iree_b200/tp_layer.py) through the C-ABI. Dimensions are chosen so that every partial sum
and every sum is an integer of magnitude <= 256 — exactly representable in bf16 — which the
script asserts; the sharded result must then EQUAL the oracle's unsharded
`truncf(matmul(...))` bit for bit on every rank, for every collective path:

  nccl/stream        ncclAllReduce on the dispatch stream (the reference's arrangement)
  nccl/overlap       ncclAllReduce on the collective stream, staged schedule
  */dataflow         one command per barrier group; the launch layer orders only commands that
                     share memory (overlap_collectives 3), NCCL and symmetric kernels
  nccl/graph         CUgraph command buffer, NCCL captured into child graphs (2+ ranks: the
                     reference has UNIMPLEMENTED here, graph_command_buffer.c:291-327)
  symmetric/overlap  our NVLink all-reduce kernels (multimem when the box has NVLS, else
                     peer loads/stores), staged schedule
  symmetric/fused    the unstaged recording order with range fusion (one persistent launch
                     per row-parallel GEMM)
  symmetric/graph    our kernels as plain kernel nodes of the CUgraph

The channel is created through the DEFAULT provider (rank / count from the launcher's
environment, id through IREE_B200_CHANNEL_ID_FILE) for the nccl paths, with an explicit id for
the others. Rank 0 prints one JSON line; exit status 1 on any mismatch.
"""

import argparse
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--id-file", default=os.path.join(tempfile.gettempdir(), "iree_b200_tp_check_id"))
    args = ap.parse_args()
    os.environ.setdefault("IREE_B200_COLLECTIVE_WATCHDOG_MS", "30000")
    os.environ["IREE_B200_CHANNEL_ID_FILE"] = args.id_file + "." + os.environ.get("MASTER_PORT", "0")
    import numpy as np
    import torch
    import torch.distributed as dist
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import oracle
    from iree_b200 import hal, symmetric, tp_layer

    dims = dict(tokens=1024, hidden=256, q_heads=4, kv_heads=2, head_dim=32, ffn=256, chunks=4)
    T, H, FFN = dims["tokens"], dims["hidden"], dims["ffn"]
    QKV = (dims["q_heads"] + 2 * dims["kv_heads"]) * dims["head_dim"]
    shapes = {"x": (T, H), "wqkv": (H, QKV), "attn": (T, H), "wo": (H, H), "wgu": (H, 2 * FFN),
              "act": (T, FFN), "wd": (FFN, H)}
    host = {k: oracle.generate_random_matrix("bf16", s, i + 1) for i, (k, s) in enumerate(shapes.items())}

    def init(name, shape):
        assert host[name].shape == tuple(shape)
        return torch.from_numpy(host[name].view(np.int16)).cuda().view(torch.bfloat16)

    # The unsharded layer on the CPU oracle (each GEMM followed by arith.truncf).
    mm = lambda a, b: oracle.from_f32("bf16", oracle.matmul("bf16", a, b))
    want = {"qkv": mm(host["x"], host["wqkv"]), "o": mm(host["attn"], host["wo"]),
            "d": mm(host["act"], host["wd"])}
    want["gu"] = mm(want["o"], host["wgu"])
    # Exactness precondition: what gets rounded to bf16 on the sharded path — every rank's
    # partial sum over its K slice and the all-reduced sum (accumulated in f32) — is an
    # integer with |v| <= 256, so rounding never changes a value.
    from iree_b200 import sharding
    for k, a, w in (("o", host["attn"], host["wo"]), ("d", host["act"], host["wd"])):
        for r in range(world):
            k0, kn = sharding.contiguous_shard(a.shape[1], world, r)
            part = oracle.matmul("bf16", np.ascontiguousarray(a[:, k0:k0 + kn]),
                                 np.ascontiguousarray(w[k0:k0 + kn]))
            assert np.abs(part).max() <= 256, f"test inputs: a partial sum of {k} reaches {np.abs(part).max()}"
        assert np.abs(oracle.matmul("bf16", a, w)).max() <= 256
    failures, ran = [], []
    stream = torch.cuda.Stream()

    def check(label, layer):
        torch.cuda.synchronize()
        plan = layer.plan
        got = {k: layer.t[k].view(torch.int16).cpu().numpy().view(np.uint16) for k in ("qkv", "o", "gu", "d")}
        qo, qn = plan["qkv"].n_offset, plan["qkv"].N
        go, gn = plan["gate_up"].n_offset, plan["gate_up"].N
        exp = {"qkv": want["qkv"][:, qo:qo + qn], "o": want["o"], "gu": want["gu"][:, go:go + gn],
               "d": want["d"]}
        for k in exp:
            if not np.array_equal(got[k], exp[k]):
                bad = int((got[k] != exp[k]).sum())
                failures.append(f"{label}: {k} differs from the oracle in {bad} of {exp[k].size} elements")
        ran.append(label)

    def default_channel(dev):
        return hal.Channel(dev)       # rank, count and id all from the default provider

    def explicit_channel(dev):
        ids = [hal.Channel.create_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        return hal.Channel(dev, rank=rank, count=world, id_bytes=ids[0])

    configs = [("nccl/stream", hal.MODE_STREAM, 0, False, True),
               ("nccl/overlap", hal.MODE_STREAM, 2, False, True),
               ("nccl/dataflow", hal.MODE_STREAM, 3, False, "dataflow"),
               ("nccl/graph", hal.MODE_GRAPH, 0, False, True),
               ("symmetric/overlap", hal.MODE_STREAM, 2, True, True),
               ("symmetric/dataflow", hal.MODE_STREAM, 3, True, "dataflow"),
               ("symmetric/fused", hal.MODE_STREAM, 1, True, False),
               ("symmetric/graph", hal.MODE_GRAPH, 0, True, True)]
    have_mc = None
    with torch.cuda.stream(stream):
        for label, mode, overlap, use_region, staged in configs:
            dev = hal.Device(local, mode=mode, external_stream=stream.cuda_stream,
                             overlap_collectives=overlap)
            ops = hal.Ops(dev)
            channel = (explicit_channel if use_region else default_channel)(dev)
            assert channel.rank_and_count() == (rank, world)
            region = None
            if use_region:
                region = symmetric.SymmetricRegion(2 * T * H * 2, torch.device("cuda", local))
                have_mc = region.multicast_base != 0
                region.attach(channel, ops.executable, max_blocks=16)
            for rep in range(2):   # twice: flag words / ready sequence / graph replay
                layer = tp_layer.DecoderLayerGemms(
                    torch, hal, dev, ops, world, rank, channel, region=region, init=init,
                    chunk_consumers=bool(staged), schedule="dataflow" if staged == "dataflow" else "staged",
                    **dims) if rep == 0 else layer
                for k in ("qkv", "o", "gu", "d"):
                    layer.t[k].fill_(float("nan"))
                dist.barrier()
                layer.step()
                dev.wait_idle()
                check(f"{label} (run {rep})", layer)
                dist.barrier()
            counts = channel.collective_counts()
            expected = "symmetric" if use_region else "nccl"
            # Stream mode counts per execution; a graph counts its collective nodes once, when
            # the CUgraph is built.
            n = layer.collectives * (1 if mode == hal.MODE_GRAPH else 2)
            if counts[expected] != n or counts["symmetric" if not use_region else "nccl"]:
                failures.append(f"{label}: collectives went {counts}, expected {n} via {expected}")
            if use_region and channel.collective_status():
                failures.append(f"{label}: collective status 0x{channel.collective_status():x}")
    bad = torch.tensor([len(failures)], device="cuda")
    dist.all_reduce(bad)
    if rank == 0:
        print(json.dumps({"world": world, "multicast": have_mc, "dims": dims, "paths": ran,
                          "failures": failures, "failing_ranks_total": int(bad.item())}), flush=True)
    else:
        for f in failures:
            print(f"rank {rank}: FAILED {f}", flush=True)
    dist.destroy_process_group()
    sys.exit(1 if int(bad.item()) else 0)


if __name__ == "__main__":
    main()
