"""sharding.py — how the paged-attention path spreads over the GPUs of one NVSwitch box.

One process per GPU (torch.distributed: NCCL on GPUs, gloo in the CPU tests).  The units of the path are
independent — attention for (sequence s, kv head h) touches only that pair's K/V blocks — so there are exactly
two partitionings and no collective inside attention:

  A. sequence-batch sharding   sequences -> ranks, balanced by context length; every rank keeps its own block
                               manager and caches; no data-path collective at all.
  B. KV-head sharding          (70B-class GQA)  rank g holds kv heads [g*Hkv/G, (g+1)*Hkv/G) of EVERY block and
                               computes the query heads that map to them; block table / seq_lens are replicated;
                               one all-gather of the F32 outputs afterwards (the reference has no counterpart:
                               its multi-GPU path is cudaMemcpyPeerAsync row splits, ggml-cuda.cu:598,1592).

Nothing here does arithmetic on K/V; torch.distributed is plumbing.
"""
from __future__ import annotations

from dataclasses import dataclass

__all__ = ["shard_sequences", "HeadShard", "head_shard", "gather_head_outputs", "merge_gathered"]


def shard_sequences(ctx_lens, world: int) -> list[list[int]]:
    """Partition A.  Longest-processing-time greedy: sequences sorted by context length (descending, index as
    tie-break) go to the currently lightest rank.  Deterministic, so every rank computes the same assignment
    without communicating.  Returns, per rank, the ascending list of sequence indices it owns."""
    if world <= 0:
        raise ValueError("world must be positive")
    order = sorted(range(len(ctx_lens)), key=lambda i: (-int(ctx_lens[i]), i))
    load = [0] * world
    owned: list[list[int]] = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda j: (load[j], len(owned[j]), j))
        owned[r].append(i)
        load[r] += int(ctx_lens[i])
    return [sorted(o) for o in owned]


@dataclass(frozen=True)
class HeadShard:
    """Partition B for one rank: kv heads [kv0, kv0 + n_head_kv) and query heads [q0, q0 + n_head)."""
    rank: int
    world: int
    kv0: int
    n_head_kv: int
    q0: int
    n_head: int


def head_shard(rank: int, world: int, n_head: int, n_head_kv: int) -> HeadShard:
    """GQA maps query head h to kv head h // (n_head / n_head_kv) (ggml-cpu.c:12264), so a contiguous range of
    kv heads owns a contiguous range of query heads."""
    if n_head % n_head_kv:
        raise ValueError("n_head must be a multiple of n_head_kv")
    if n_head_kv % world:
        raise ValueError(f"{n_head_kv} kv heads do not split over {world} ranks")
    if not 0 <= rank < world:
        raise ValueError("rank out of range")
    per = n_head_kv // world
    gqa = n_head // n_head_kv
    return HeadShard(rank, world, rank * per, per, rank * per * gqa, per * gqa)


def gather_head_outputs(local_out, gathered=None, group=None, async_op: bool = False):
    """The single collective of partition B: local_out [n_q, n_head/G, d] F32 (contiguous) from every rank ->
    gathered [G, n_q, n_head/G, d].  One all_gather_into_tensor (NCCL over NVLink/NVSwitch on GPUs)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    if gathered is None:
        gathered = torch.empty((world,) + tuple(local_out.shape), dtype=local_out.dtype, device=local_out.device)
    # output passed as [G*n_q, ...]: the concatenation form both NCCL and gloo accept
    work = dist.all_gather_into_tensor(gathered.view((-1,) + tuple(local_out.shape[1:])), local_out.contiguous(), group=group,
                                       async_op=async_op)
    return (gathered, work) if async_op else gathered


def merge_gathered(gathered):
    """[G, n_q, n_head/G, d] -> the op's dst layout [n_q, n_head, d] (a view + one copy; the consumer of the
    attention output — the wo matmul — can equally read the permuted view)."""
    g, n_q, hpg, d = gathered.shape
    return gathered.permute(1, 0, 2, 3).reshape(n_q, g * hpg, d)
