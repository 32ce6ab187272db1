"""Host-side partitioning of the dense-contraction workloads over the GPUs of one box
(SURVEY.md §8(e)). Pure arithmetic: shared by bench.py and the world-size-2 gloo tests.

* batch_matmul / attention: the batch (or head) dimension is split in contiguous slices, no
  data-path collective.
* decoder-layer GEMMs: Megatron split. Column-parallel weights are sliced along N (each
  rank produces a column slab of the activation, no exchange); row-parallel weights are
  sliced along K (each rank produces a full-shape partial sum) and followed by an
  all_reduce(sum) through iree_hal_channel (flow.collective.all_reduce, FlowOps.td:1788-1941
  → hal.command_buffer.collective → ncclAllReduce, nccl_channel.c:360-380).
"""

from __future__ import annotations

from dataclasses import dataclass


def contiguous_shard(extent: int, world: int, rank: int):
    """(offset, length) of rank's slice; the first (extent % world) ranks get one more."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(extent, world)
    offset = rank * base + min(rank, extra)
    return offset, base + (1 if rank < extra else 0)


@dataclass(frozen=True)
class Gemm:
    name: str
    M: int
    N: int          # per-rank N
    K: int          # per-rank K
    parallel: str   # "column" | "row"
    n_offset: int   # slice of the full weight this rank owns
    k_offset: int

    @property
    def flop(self) -> float:
        return 2.0 * self.M * self.N * self.K


def decoder_layer_plan(world: int, rank: int, tokens=8192, hidden=4096, q_heads=32, kv_heads=8,
                       head_dim=128, ffn=14336):
    """The four GEMMs of one Llama-style decoder layer on |rank| of a TP group of |world|."""
    qkv = (q_heads + 2 * kv_heads) * head_dim
    for extent, what in ((qkv, "QKV width"), (hidden, "hidden"), (ffn, "FFN")):
        if extent % world:
            raise ValueError(f"{what} {extent} is not divisible by the TP degree {world}")
    shard = lambda extent: contiguous_shard(extent, world, rank)
    qo, qn = shard(qkv)
    ho, hn = shard(hidden)
    fo, fn = shard(ffn)
    return [
        Gemm("qkv", tokens, qn, hidden, "column", qo, 0),
        Gemm("o", tokens, hidden, hn, "row", 0, ho),
        Gemm("gate_up", tokens, 2 * fn, hidden, "column", 2 * fo, 0),
        Gemm("down", tokens, hidden, fn, "row", 0, fo),
    ]


def token_chunks(tokens: int, chunks):
    """Row chunks of a row-parallel GEMM; chunk c's all-reduce overlaps chunk c+1's GEMM.
    |chunks|: a count (equal chunks) or a list of row counts summing to |tokens| (a short
    first chunk starts the link early; few, larger chunks keep the per-collective cost low)."""
    if isinstance(chunks, int):
        return [contiguous_shard(tokens, chunks, c) for c in range(chunks)]
    rows = [int(r) for r in chunks]
    if sum(rows) != tokens or any(r <= 0 for r in rows):
        raise ValueError(f"chunk rows {rows} do not partition {tokens} tokens")
    out, at = [], 0
    for r in rows:
        out.append((at, r))
        at += r
    return out
