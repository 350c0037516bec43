"""sharding.py — how the paged-attention path spreads over the GPUs of one NVSwitch box.

One process per GPU (torch.distributed: NCCL on GPUs, gloo in the CPU tests).  The units of the path are
independent — attention for (sequence s, kv head h) touches only that pair's K/V blocks — so there are exactly
two partitionings and no collective inside attention:

  A. sequence-batch sharding   sequences -> ranks, balanced by context length; every rank keeps its own block
                               manager and caches; no data-path collective at all.
  B. KV-head sharding          (70B-class GQA)  rank g holds kv heads [g*Hkv/G, (g+1)*Hkv/G) of EVERY block and
                               computes the query heads that map to them; block table / seq_lens are replicated;
                               one exchange of the F32 outputs afterwards (the reference has no counterpart:
                               its multi-GPU path is cudaMemcpyPeerAsync row splits, ggml-cuda.cu:598,1592):
                               PeerGather pushes the slices over NVLink peer memory with the library's own
                               kernels (csrc/peer_gather.cu); gather_head_outputs is the NCCL/gloo form.

Nothing here does arithmetic on K/V; torch.distributed is plumbing.
"""
from __future__ import annotations

from dataclasses import dataclass

__all__ = ["shard_sequences", "HeadShard", "head_shard", "gather_head_outputs", "merge_gathered", "PeerGather"]


def shard_sequences(ctx_lens, world: int) -> list[list[int]]:
    """Partition A.  Longest-processing-time greedy: sequences sorted by context length (descending, index as
    tie-break) go to the currently lightest rank.  Deterministic, so every rank computes the same assignment
    without communicating.  Returns, per rank, the ascending list of sequence indices it owns.
    The policy itself is the library's (b200_shard_sequences, host/sharding.cpp): a C host gets the same answer."""
    import ctypes as C
    from . import lib
    if world <= 0:
        raise ValueError("world must be positive")
    n = len(ctx_lens)
    lens = (C.c_int32 * max(n, 1))(*[int(x) for x in ctx_lens])
    rank_of = (C.c_int32 * max(n, 1))()
    if lib.b200_shard_sequences(lens, n, world, rank_of) != 0:
        raise ValueError("b200_shard_sequences: bad arguments")
    owned: list[list[int]] = [[] for _ in range(world)]
    for i in range(n):
        owned[rank_of[i]].append(i)
    return owned


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
    import ctypes as C
    from . import lib
    out = (C.c_int32 * 4)()
    rc = lib.b200_head_shard(rank, world, n_head, n_head_kv, out)     # the library's policy (host/sharding.cpp)
    if rc == -2:
        raise ValueError("n_head must be a multiple of n_head_kv")
    if rc == -3:
        raise ValueError(f"{n_head_kv} kv heads do not split over {world} ranks")
    if rc != 0:
        raise ValueError("rank out of range")
    return HeadShard(rank, world, out[0], out[1], out[2], out[3])


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


class _DevArray:
    """raw device memory presented through __cuda_array_interface__ so torch can alias it without a copy"""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (ptr, False), "version": 2}


class PeerGather:
    """The exchange step of partition B over NVLink peer memory (csrc/peer_gather.cu): every rank pushes its output
    slice into every peer's gathered buffer with plain stores through CUDA-IPC-mapped pointers and raises an epoch
    flag; `wait` spins on the local flags.  torch.distributed is used once, to exchange the 64-byte IPC handles.

        pg = PeerGather(local_shape, n_slots)        # e.g. (n_q, n_head/G, d), F32; one slot per layer
        pg.publish(local_out, slot); pg.wait()       # on the current stream; CUDA-graph capturable
        pg.gathered(slot)                            # [G, n_q, n_head/G, d] view, valid after that wait
    Two consecutive exchanges must use different slots (a fast rank may already be pushing exchange e+1 while a slow
    one still reads the slot of exchange e).
    """

    def __init__(self, local_shape, n_slots: int = 2, group=None):
        import ctypes as C
        import torch
        import torch.distributed as dist
        from . import lib
        self._lib, self._C, self._torch = lib, C, torch
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.shape = tuple(int(x) for x in local_shape)
        n = 1
        for x in self.shape:
            n *= x
        self.bytes = n*4
        if self.bytes % 16:
            raise ValueError("slice size must be a multiple of 16 bytes")
        if n_slots < 2:
            raise ValueError("n_slots must be at least 2")
        self.n_slots = n_slots
        self._buf = lib.b200_peer_alloc(n_slots*self.world*self.bytes)
        self._flags = lib.b200_peer_alloc(64*4)
        self._state = lib.b200_peer_alloc(16)
        self._ll = lib.b200_peer_alloc(n_slots*self.world*2*self.bytes)      # (payload, epoch) pairs, see exchange_ll
        if not (self._buf and self._flags and self._state and self._ll):
            raise RuntimeError("b200_peer_alloc failed: " + lib.b200_last_error().decode(errors="replace"))
        hb, hf, hl = C.create_string_buffer(64), C.create_string_buffer(64), C.create_string_buffer(64)
        if lib.b200_peer_export(self._buf, hb) or lib.b200_peer_export(self._flags, hf) or lib.b200_peer_export(self._ll, hl):
            raise RuntimeError("b200_peer_export failed: " + lib.b200_last_error().decode(errors="replace"))
        handles = [None]*self.world
        dist.all_gather_object(handles, (hb.raw, hf.raw, hl.raw), group=group)  # plumbing: 192 bytes per rank, once
        self._peer_buf, self._peer_flag, self._peer_ll = (C.c_void_p*self.world)(), (C.c_void_p*self.world)(), (C.c_void_p*self.world)()
        self._opened = []
        for r, (b, f, l) in enumerate(handles):
            if r == self.rank:
                self._peer_buf[r], self._peer_flag[r], self._peer_ll[r] = self._buf, self._flags, self._ll
            else:
                pb, pf, pl = lib.b200_peer_open(b), lib.b200_peer_open(f), lib.b200_peer_open(l)
                if not (pb and pf and pl):
                    raise RuntimeError(f"b200_peer_open(rank {r}) failed: " + lib.b200_last_error().decode(errors="replace"))
                self._peer_buf[r], self._peer_flag[r], self._peer_ll[r] = pb, pf, pl
                self._opened += [pb, pf, pl]
        self._all = torch.as_tensor(_DevArray(self._buf, (n_slots, self.world) + self.shape, "<f4"), device="cuda")
        self.epoch = 0                       # host mirror of the device epoch (publishes issued so far), for check()
        self._last_slot = -1
        dist.barrier(group=group)                                               # nobody publishes into an unmapped peer

    def exchange(self, local_out, slot: int, stream=None):
        """publish + wait in one launch"""
        self._push(self._lib.b200_peer_exchange, "b200_peer_exchange", local_out, slot, stream)

    def exchange_ll(self, local_out, slot: int, stream=None):
        """lowest latency: payload and epoch travel in the same 8-byte stores (no fence, no flag round trip); one
        launch pushes to every peer and unpacks what the peers pushed into gathered(slot)"""
        st = self._torch.cuda.current_stream().cuda_stream if stream is None else stream
        if tuple(local_out.shape) != self.shape or not local_out.is_contiguous():
            raise ValueError("local_out must be contiguous with the shape given at construction")
        if not 0 <= slot < self.n_slots or slot == self._last_slot:
            raise ValueError("slot out of range, or the same slot as the previous exchange")
        self._last_slot = slot
        rc = self._lib.b200_peer_exchange_ll(local_out.data_ptr(), self.bytes, self._peer_ll, self._buf, self._state,
                                             slot, self.rank, self.world, self._C.c_void_p(st))
        if rc:
            raise RuntimeError("b200_peer_exchange_ll: " + self._lib.b200_last_error().decode(errors="replace"))
        self.epoch += 1

    def publish(self, local_out, slot: int, stream=None):
        self._push(self._lib.b200_peer_publish, "b200_peer_publish", local_out, slot, stream)

    def _push(self, fn, what, local_out, slot, stream):
        st = self._torch.cuda.current_stream().cuda_stream if stream is None else stream
        if tuple(local_out.shape) != self.shape or not local_out.is_contiguous():
            raise ValueError("local_out must be contiguous with the shape given at construction")
        if not 0 <= slot < self.n_slots or slot == self._last_slot:
            raise ValueError("slot out of range, or the same slot as the previous exchange")
        self._last_slot = slot
        rc = fn(local_out.data_ptr(), self.bytes, self._peer_buf, self._peer_flag, self._state,
                slot, self.rank, self.world, self._C.c_void_p(st))
        if rc:
            raise RuntimeError(what + ": " + self._lib.b200_last_error().decode(errors="replace"))
        self.epoch += 1

    def note_replay(self, n_publishes: int):
        """a captured CUDA graph holding n_publishes publish launches was replayed: keep the host epoch in step"""
        self.epoch += n_publishes

    def wait(self, stream=None):
        st = self._torch.cuda.current_stream().cuda_stream if stream is None else stream
        rc = self._lib.b200_peer_wait(self._flags, self._state, self.world, self._C.c_void_p(st))
        if rc:
            raise RuntimeError("b200_peer_wait: " + self._lib.b200_last_error().decode(errors="replace"))

    def state(self):
        """(epoch, ctas_done, wait_error, _) after a device synchronise"""
        out = (self._C.c_uint32*4)()
        self._lib.b200_peer_read_state(self._state, out)
        return tuple(out)

    def gathered(self, slot: int):
        """[G, *local_shape] view of a slot (valid in stream order after the wait of the exchange that filled it)"""
        return self._all[slot]

    def all_slots(self):
        """[n_slots, G, *local_shape] view of every slot"""
        return self._all

    def check(self):
        """synchronise and verify the device epoch matches the host's and no wait timed out"""
        self._torch.cuda.synchronize()
        epoch, _, err, _ = self.state()
        if err:
            raise RuntimeError(f"peer wait timed out on rank {err - 1}'s flag")
        if epoch != self.epoch:
            raise RuntimeError(f"device epoch {epoch} != host epoch {self.epoch} (graph replays not reported with note_replay?)")

    def close(self):
        for p in self._opened:
            self._lib.b200_peer_close(p)
        self._opened = []
        for p in (self._buf, self._flags, self._state, self._ll):
            if p:
                self._lib.b200_peer_free(p)
        self._buf = self._flags = self._state = self._ll = None
        self._all = None
