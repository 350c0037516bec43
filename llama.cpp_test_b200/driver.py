"""driver.py — thin helpers that hand torch device buffers to the C ABI, plus ctypes wrappers of the
host-side block managers.  No arithmetic happens here."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

from . import (AttnDesc, B200Error, KShiftDesc, KVWriteDesc, KERNEL_AUTO, STATE_ERRORS, TYPE_F16, TYPE_F32, TYPE_Q4_0, TYPE_Q8_0,  # noqa: F401
               check, lib, row_bytes)

__all__ = ["KVLayout", "AttnCall", "StepCall", "kv_write_desc", "paged_attention", "dense_attention", "kv_write", "block_copy",
           "kv_rows_gather", "kv_rows_scatter", "k_shift", "context_shift", "state_seq_pack", "state_seq_unpack", "state_seq_save", "state_seq_load", "state_all_pack", "state_all_restore",
           "BlockManager", "PagedKVCache", "LlamaKVCachePaged", "current_stream"]


def current_stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


@dataclass(frozen=True)
class KVLayout:
    """The flat llama.cpp cache (src/llama-kv-cache.cpp:100-101) seen as blocks of `block_size` cells:
    cell c, kv-head g at  c*token_bytes + g*head_bytes ; block b = cells [b*block_size, (b+1)*block_size)."""
    type_id: int
    n_head_kv: int
    head_dim: int
    block_size: int
    n_blocks: int

    @property
    def head_bytes(self) -> int:
        return row_bytes(self.type_id, self.head_dim)

    @property
    def token_bytes(self) -> int:
        return self.head_bytes * self.n_head_kv

    @property
    def block_bytes(self) -> int:
        return self.token_bytes * self.block_size

    @property
    def n_cells(self) -> int:
        return self.n_blocks * self.block_size

    @property
    def total_bytes(self) -> int:
        return self.n_cells * self.token_bytes

    def alloc(self, device="cuda"):
        import torch
        return torch.zeros(self.total_bytes, dtype=torch.uint8, device=device)


def _ptr(t) -> int:
    return 0 if t is None else t.data_ptr()


class AttnCall:
    """A prepared b200_paged_attention call (descriptor + workspace), reusable across launches."""

    def __init__(self, q, k_cache, v_cache, layout_k: KVLayout, out, scale: float, *, layout_v: KVLayout | None = None,
                 block_table=None, seq_lens=None, q_seq=None, q_pos=None, mask=None, n_kv: int | None = None,
                 max_seq_len: int = 0, total_tokens: int = 0, softcap: float = 0.0, max_bias: float = 0.0,
                 kernel: int = KERNEL_AUTO, workspace=None, reuse_plan: bool = False):
        """workspace: a uint8 device tensor shared with other calls on the same stream (the layers of a step; big
        enough for each) instead of a private one.  reuse_plan: B200_ATTN_REUSE_PLAN — the caller guarantees the
        previous call on this stream with this workspace had the same n_q / heads / kernel and unchanged seq_lens /
        q_seq / q_pos (b200_decode_step / StepCall works that out by itself)."""
        import torch
        layout_v = layout_v or layout_k
        assert q.dtype == torch.float32 and q.dim() == 3 and q.is_contiguous()
        assert out.dtype == torch.float32 and out.is_contiguous() and out.shape == q.shape
        n_q, n_head, d = q.shape
        desc = AttnDesc()
        desc.n_q, desc.n_head, desc.n_head_kv, desc.head_dim = n_q, n_head, layout_k.n_head_kv, d
        desc.type_k, desc.type_v = layout_k.type_id, layout_v.type_id
        desc.q, desc.nbq1, desc.nbq2 = q.data_ptr(), n_head * d * 4, d * 4       # [d, n_q, n_head] permuted view
        desc.k, desc.nbk1, desc.nbk2 = k_cache.data_ptr(), layout_k.token_bytes, layout_k.head_bytes
        desc.v, desc.nbv1, desc.nbv2 = v_cache.data_ptr(), layout_v.token_bytes, layout_v.head_bytes
        desc.dst = out.data_ptr()
        desc.scale, desc.max_bias, desc.logit_softcap = scale, max_bias, softcap
        desc.n_kv = layout_k.n_cells if n_kv is None else n_kv
        if block_table is not None:
            assert block_table.dtype == torch.int32 and block_table.dim() == 2 and block_table.stride(1) == 1
            assert seq_lens is not None and seq_lens.dtype == torch.int32
            desc.block_table, desc.bt_stride = block_table.data_ptr(), block_table.stride(0)
            desc.seq_lens, desc.n_seqs, desc.block_size = seq_lens.data_ptr(), block_table.shape[0], layout_k.block_size
            desc.q_seq, desc.q_pos = _ptr(q_seq), _ptr(q_pos)
            desc.max_seq_len, desc.total_tokens = max_seq_len, total_tokens
        elif mask is not None:
            assert mask.dtype == torch.float16 and mask.dim() == 2 and mask.stride(1) == 1
            desc.mask, desc.nb_mask1 = mask.data_ptr(), mask.stride(0) * 2
        desc.kernel = kernel
        desc.flags = 1 if reuse_plan else 0
        self.desc = desc
        ws = lib.b200_paged_attention_workspace(C.byref(desc))
        if workspace is not None:
            if workspace.numel() < ws:
                raise ValueError(f"shared workspace of {workspace.numel()} bytes is smaller than the {ws} this call needs")
            self.workspace = workspace
        else:
            self.workspace = torch.empty(max(ws, 16), dtype=torch.uint8, device=q.device)
        desc.workspace, desc.workspace_bytes = self.workspace.data_ptr(), self.workspace.numel()
        self._keep = (q, k_cache, v_cache, out, block_table, seq_lens, q_seq, q_pos, mask)

    def supported(self) -> bool:
        return bool(lib.b200_supports_attention(C.byref(self.desc)))

    def run(self, stream: int | None = None) -> None:
        check(lib.b200_paged_attention(C.byref(self.desc), C.c_void_p(current_stream() if stream is None else stream)),
              "b200_paged_attention")


def paged_attention(q, k_cache, v_cache, layout: KVLayout, block_table, seq_lens, scale: float, **kw):
    """q F32 [n_q, n_head, d] -> out F32 [n_q, n_head, d] through the block table."""
    import torch
    out = kw.pop("out", None)
    if out is None:
        out = torch.empty_like(q)
    AttnCall(q, k_cache, v_cache, layout, out, scale, block_table=block_table, seq_lens=seq_lens, **kw).run()
    return out


def dense_attention(q, k_cache, v_cache, layout: KVLayout, n_kv: int, mask, scale: float, **kw):
    """reference-semantics mode: flat cache cells [0, n_kv) + F16 mask [>=n_q, n_kv]."""
    import torch
    out = kw.pop("out", None)
    if out is None:
        out = torch.empty_like(q)
    AttnCall(q, k_cache, v_cache, layout, out, scale, mask=mask, n_kv=n_kv, **kw).run()
    return out


class StepCall:
    """b200_decode_step over a fixed list of layers: descriptor arrays are built once, one FFI call per step."""

    def __init__(self, attn_calls, write_descs=None):
        n = len(attn_calls)
        self.n = n
        self.attns = (AttnDesc * n)(*[c.desc for c in attn_calls])
        self.writes = None if write_descs is None else (KVWriteDesc * n)(*write_descs)
        self._keep = (attn_calls, write_descs)

    def run(self, stream: int | None = None) -> None:
        check(lib.b200_decode_step(None if self.writes is None else C.cast(self.writes, C.c_void_p), C.cast(self.attns, C.c_void_p),
                                   self.n, C.c_void_p(current_stream() if stream is None else stream)), "b200_decode_step")


def kv_write_desc(k_src, v_src, slot_mapping, k_cache, v_cache, layout_k: KVLayout, layout_v: KVLayout | None = None) -> KVWriteDesc:
    layout_v = layout_v or layout_k
    src = k_src if k_src is not None else v_src
    d = KVWriteDesc()
    d.k_src, d.v_src, d.src_nb1 = _ptr(k_src), _ptr(v_src), src.stride(0) * 4
    d.n_tokens = src.shape[0]
    d.slot_mapping = _ptr(slot_mapping)
    d.k_cache, d.v_cache = _ptr(k_cache), _ptr(v_cache)
    d.type_k, d.type_v, d.n_head_kv, d.head_dim = layout_k.type_id, layout_v.type_id, layout_k.n_head_kv, layout_k.head_dim
    d.nbk_token, d.nbk_head = layout_k.token_bytes, layout_k.head_bytes
    d.nbv_token, d.nbv_head = layout_v.token_bytes, layout_v.head_bytes
    d.n_cells = layout_k.n_cells
    return d


def kv_write(k_src, v_src, slot_mapping, k_cache, v_cache, layout_k: KVLayout, layout_v: KVLayout | None = None,
             stream: int | None = None) -> None:
    """k_src/v_src F32 [n_tokens, n_head_kv*head_dim] -> typed cache cells slot_mapping[t]."""
    layout_v = layout_v or layout_k
    src = k_src if k_src is not None else v_src
    d = KVWriteDesc()
    d.k_src, d.v_src, d.src_nb1 = _ptr(k_src), _ptr(v_src), src.stride(0) * 4
    d.n_tokens = src.shape[0]
    d.slot_mapping = _ptr(slot_mapping)
    d.k_cache, d.v_cache = _ptr(k_cache), _ptr(v_cache)
    d.type_k, d.type_v, d.n_head_kv, d.head_dim = layout_k.type_id, layout_v.type_id, layout_k.n_head_kv, layout_k.head_dim
    d.nbk_token, d.nbk_head = layout_k.token_bytes, layout_k.head_bytes
    d.nbv_token, d.nbv_head = layout_v.token_bytes, layout_v.head_bytes
    d.n_cells = layout_k.n_cells
    check(lib.b200_kv_write(C.byref(d), C.c_void_p(current_stream() if stream is None else stream)), "b200_kv_write")


def block_copy(caches, block_bytes: int, n_blocks: int, pairs, stream: int | None = None) -> None:
    """caches: list of device byte tensors sharing one block geometry; pairs: int32 device tensor [n, 2]."""
    arr = (C.c_void_p * len(caches))(*[c.data_ptr() for c in caches])
    check(lib.b200_block_copy(arr, len(caches), block_bytes, n_blocks, C.c_void_p(pairs.data_ptr()), pairs.shape[0],
                              C.c_void_p(current_stream() if stream is None else stream)), "b200_block_copy")


def _kv_rows(fn, what, caches, layout: KVLayout, slots, staging, stream):
    arr = (C.c_void_p * len(caches))(*[c.data_ptr() for c in caches])
    check(fn(arr, len(caches), layout.token_bytes, layout.token_bytes, layout.n_cells, C.c_void_p(slots.data_ptr()), slots.shape[0],
             C.c_void_p(staging.data_ptr()), C.c_void_p(current_stream() if stream is None else stream)), what)


def kv_rows_gather(caches, layout: KVLayout, slots, staging, stream: int | None = None) -> None:
    """staging[c][i] = token row of cell slots[i] in caches[c]; caches share one layout, slots int32 device [n],
    staging uint8 device [len(caches), n, token_bytes]"""
    _kv_rows(lib.b200_kv_rows_gather, "b200_kv_rows_gather", caches, layout, slots, staging, stream)


def kv_rows_scatter(caches, layout: KVLayout, slots, staging, stream: int | None = None) -> None:
    _kv_rows(lib.b200_kv_rows_scatter, "b200_kv_rows_scatter", caches, layout, slots, staging, stream)


def k_shift(k_caches, layout: KVLayout, shift, cells=None, *, n_rot: int | None = None, mode: int = 0, n_ctx_orig: int = 8192,
            freq_base: float = 500000.0, freq_scale: float = 1.0, ext_factor: float = 0.0, attn_factor: float = 1.0,
            beta_fast: float = 32.0, beta_slow: float = 1.0, freq_factors=None, stream: int | None = None) -> None:
    """re-rotate K rows in place (the reference's K-shift graph, src/llama-context.cpp:524-571) for every layer in one
    launch.  shift int32 device [n_cells] indexed by cell; cells int32 device list of cells to visit (None: all);
    the rope parameters are ggml_rope_ext's (defaults: Llama-3)."""
    d = KShiftDesc()
    arr = (C.c_void_p * len(k_caches))(*[c.data_ptr() for c in k_caches])
    d.caches, d.n_caches = arr, len(k_caches)
    d.type_k, d.n_head_kv, d.head_dim = layout.type_id, layout.n_head_kv, layout.head_dim
    d.nb_token, d.nb_head, d.n_cells = layout.token_bytes, layout.head_bytes, layout.n_cells
    d.shift = _ptr(shift)
    d.cells, d.n_rows = _ptr(cells), (0 if cells is None else cells.shape[0])
    d.n_rot, d.mode, d.n_ctx_orig = (n_rot or layout.head_dim), mode, n_ctx_orig
    d.freq_base, d.freq_scale, d.ext_factor, d.attn_factor = freq_base, freq_scale, ext_factor, attn_factor
    d.beta_fast, d.beta_slow = beta_fast, beta_slow
    d.freq_factors = _ptr(freq_factors)
    check(lib.b200_k_shift(C.byref(d), C.c_void_p(current_stream() if stream is None else stream)), "b200_k_shift")


def context_shift(cache: "PagedKVCache", seq: int, p0: int, p1: int, k_caches, v_caches, layout_k: KVLayout,
                  layout_v: KVLayout | None = None, **rope) -> bool:
    """llama_kv_self_seq_rm(seq, p0, p1) + llama_kv_self_seq_add(seq, p1, -1, p0 - p1) + the K-shift of the next
    kv_self_update (src/llama-context.cpp:719-790) on the paged layout: block-table edit on the host, then at most four
    launches (copy-on-write copies, row gather, row scatter, K rotation) over all layers.  False: refused, nothing changed."""
    import torch
    layout_v = layout_v or layout_k
    plan = cache.context_shift(seq, p0, p1)
    if plan is None:
        return False
    cow, moves, cells = plan
    dev = k_caches[0].device
    if cow:
        pairs = torch.tensor(cow, dtype=torch.int32, device=dev)
        block_copy(list(k_caches), layout_k.block_bytes, layout_k.n_blocks, pairs)
        block_copy(list(v_caches), layout_v.block_bytes, layout_v.n_blocks, pairs)
    if moves:
        src = torch.tensor([m[0] for m in moves], dtype=torch.int32, device=dev)
        dst = torch.tensor([m[1] for m in moves], dtype=torch.int32, device=dev)
        for caches, lay in ((k_caches, layout_k), (v_caches, layout_v)):
            st = torch.empty((len(caches), len(moves), lay.token_bytes), dtype=torch.uint8, device=dev)
            kv_rows_gather(list(caches), lay, src, st)
            kv_rows_scatter(list(caches), lay, dst, st)
    if cells:
        cl = torch.tensor(cells, dtype=torch.int32, device=dev)
        shift = torch.zeros(layout_k.n_cells, dtype=torch.int32, device=dev)
        shift[cl.long()] = p0 - p1
        k_shift(list(k_caches), layout_k, shift, cl, **rope)
    return True


def state_seq_pack(k_rows, v_rows, type_k: int, type_v: int) -> bytes:
    """frame token-ordered rows (uint8 numpy [n_layer, n_cells, size_row]) as the reference's single-sequence KV stream
    (llama_kv_cache_unified::state_write, src/llama-kv-cache.cpp:848-971)"""
    import numpy as np
    k_rows, v_rows = np.ascontiguousarray(k_rows, dtype=np.uint8), np.ascontiguousarray(v_rows, dtype=np.uint8)
    n_layer, n_cells, k_row = k_rows.shape
    v_row = v_rows.shape[2]
    size = lib.b200_kv_state_seq_size(n_cells, n_layer, k_row, v_row)
    out = np.empty(size, dtype=np.uint8)
    n = lib.b200_kv_state_seq_write(out.ctypes.data, size, n_cells, n_layer, type_k, k_row, type_v, v_row,
                                    k_rows.ctypes.data, v_rows.ctypes.data)
    if n != size:
        raise B200Error(f"b200_kv_state_seq_write: {n}")
    return out.tobytes()


def state_seq_unpack(data: bytes, n_layer: int, type_k: int, k_row: int, type_v: int, v_row: int):
    """validate a stream against the cache's shape like state_read_data (:1105-1240) -> (k_rows, v_rows) uint8 numpy
    [n_layer, n_cells, size_row]; raises B200Error with the reference's reason on a mismatch"""
    import numpy as np
    buf = np.frombuffer(data, dtype=np.uint8)
    n_cells = C.c_int64()
    k_off, v_off = (C.c_int64 * max(n_layer, 1))(), (C.c_int64 * max(n_layer, 1))()
    rc = lib.b200_kv_state_seq_parse(buf.ctypes.data, buf.size, n_layer, type_k, k_row, type_v, v_row, C.byref(n_cells), k_off, v_off)
    if rc != 0:
        raise B200Error(f"failed to restore kv cache: {STATE_ERRORS.get(rc, rc)}")
    n = n_cells.value
    k = np.stack([buf[k_off[l]:k_off[l] + n * k_row].reshape(n, k_row) for l in range(n_layer)]) if n_layer else np.zeros((0, n, k_row), np.uint8)
    v = np.stack([buf[v_off[l]:v_off[l] + n * v_row].reshape(n, v_row) for l in range(n_layer)]) if n_layer else np.zeros((0, n, v_row), np.uint8)
    # the stream's rows come in the writer's cell order, which need not be position order: hand them out by position
    pos = (C.c_int32 * max(n, 1))()
    if lib.b200_kv_state_seq_positions(buf.ctypes.data, buf.size, pos, n) != n:
        raise B200Error("failed to restore kv cache: truncated stream")
    order = np.argsort(np.frombuffer(pos, dtype=np.int32, count=n), kind="stable")
    if not np.array_equal(order, np.arange(n)):
        k, v = k[:, order], v[:, order]
    return k, v


def state_all_pack(cache: "PagedKVCache", k_rows, v_rows, type_k: int, type_v: int) -> bytes:
    """the whole-cache stream (llama_state_get_data's KV section) from rows gathered in cache.used_cells() order:
    k_rows / v_rows uint8 [n_layer, n_cells, size_row]"""
    import numpy as np
    k_rows = np.ascontiguousarray(k_rows, dtype=np.uint8); v_rows = np.ascontiguousarray(v_rows, dtype=np.uint8)
    n_layer, _, k_row = k_rows.shape
    v_row = v_rows.shape[2]
    need = -lib.b200_kv_state_all_write(None, 0, cache._h, n_layer, type_k, k_row, type_v, v_row, None, None)
    out = np.empty(need, dtype=np.uint8)
    n = lib.b200_kv_state_all_write(out.ctypes.data, need, cache._h, n_layer, type_k, k_row, type_v, v_row,
                                    k_rows.ctypes.data, v_rows.ctypes.data)
    if n != need:
        raise B200Error(f"b200_kv_state_all_write: {n}")
    return out.tobytes()


def state_all_restore(cache: "PagedKVCache", data: bytes, n_layer: int, type_k: int, k_row: int, type_v: int, v_row: int):
    """llama_state_set_data's KV section: validates, clears the cache, rebuilds every sequence.
    -> (src_rows, dst_cells, k_rows, v_rows): copy row src_rows[i] of each layer's rows to cell dst_cells[i]"""
    import numpy as np
    buf = np.frombuffer(data, dtype=np.uint8)
    n_cells = C.c_int64(0)
    k_off, v_off = (C.c_int64 * max(n_layer, 1))(), (C.c_int64 * max(n_layer, 1))()
    cap = 16
    while True:
        src, dst = (C.c_int32 * cap)(), (C.c_int32 * cap)()
        rc = lib.b200_kv_state_all_restore(cache._h, buf.ctypes.data, buf.size, n_layer, type_k, k_row, type_v, v_row,
                                           src, dst, cap, C.byref(n_cells), k_off, v_off)
        if rc <= -100:
            cap = -rc - 100
            continue
        break
    if rc < 0:
        raise B200Error(f"kv state: {STATE_ERRORS.get(int(rc), rc)}")
    n = n_cells.value
    k_rows = np.stack([buf[k_off[l]:k_off[l] + n * k_row].reshape(n, k_row) for l in range(n_layer)]) if n_layer else np.zeros((0, n, k_row), np.uint8)
    v_rows = np.stack([buf[v_off[l]:v_off[l] + n * v_row].reshape(n, v_row) for l in range(n_layer)]) if n_layer else np.zeros((0, n, v_row), np.uint8)
    return list(src[:rc]), list(dst[:rc]), k_rows, v_rows


def state_seq_save(cache: "PagedKVCache", seq: int, k_caches, v_caches, layout_k: KVLayout, layout_v: KVLayout | None = None) -> bytes:
    """llama_state_seq_get_data for the paged layout: gather the sequence's rows of every layer on the device (two
    launches), one D2H copy each for K and V, frame them in the reference's stream format"""
    import torch
    layout_v = layout_v or layout_k
    cells = cache.slots_of(seq)
    if cells is None:
        cells = []
    dev = k_caches[0].device
    slots = torch.tensor(cells, dtype=torch.int32, device=dev)
    n, nl = len(cells), len(k_caches)
    ks = torch.empty((nl, n, layout_k.token_bytes), dtype=torch.uint8, device=dev)
    vs = torch.empty((nl, n, layout_v.token_bytes), dtype=torch.uint8, device=dev)
    if n:
        kv_rows_gather(k_caches, layout_k, slots, ks)
        kv_rows_gather(v_caches, layout_v, slots, vs)
    return state_seq_pack(ks.cpu().numpy(), vs.cpu().numpy(), layout_k.type_id, layout_v.type_id)


def state_seq_load(cache: "PagedKVCache", seq: int, data: bytes, k_caches, v_caches, layout_k: KVLayout,
                   layout_v: KVLayout | None = None) -> int:
    """llama_state_seq_set_data for the paged layout: validate, drop `seq`, reserve blocks, scatter the rows into
    them.  Returns the number of restored tokens; raises B200Error("failed to restore kv cache ...") like the reference"""
    import torch
    layout_v = layout_v or layout_k
    k, v = state_seq_unpack(data, len(k_caches), layout_k.type_id, layout_k.token_bytes, layout_v.type_id, layout_v.token_bytes)
    n = k.shape[1]
    cells = cache.restore_seq(seq, n)
    if cells is None:
        raise B200Error("failed to restore kv cache: failed to find available cells in kv cache")
    if n:
        dev = k_caches[0].device
        slots = torch.tensor(cells, dtype=torch.int32, device=dev)
        kv_rows_scatter(k_caches, layout_k, slots, torch.from_numpy(k.copy()).to(dev))
        kv_rows_scatter(v_caches, layout_v, slots, torch.from_numpy(v.copy()).to(dev))
    return n


class BlockManager:
    """the fork's BlockManager through the C ABI (include/b200_host.h)"""

    def __init__(self, num_blocks: int):
        self._h = lib.b200_bm_create(num_blocks)
        self.num_blocks = num_blocks

    def __del__(self):
        if getattr(self, "_h", None):
            lib.b200_bm_destroy(self._h)
            self._h = None

    def allocateBlock(self) -> int:
        return lib.b200_bm_allocate(self._h)

    def freeBlock(self, i: int) -> None:
        lib.b200_bm_free(self._h, i)

    def evictBlock(self) -> int:
        return lib.b200_bm_evict(self._h)

    def getBlockStatus(self) -> list[bool]:
        buf = (C.c_uint8 * max(self.num_blocks, 1))()
        n = lib.b200_bm_status(self._h, buf, self.num_blocks)
        return [bool(buf[i]) for i in range(n)]


class PagedKVCache:
    """PagedKVCache through the C ABI"""

    def __init__(self, n_blocks: int, block_size: int):
        self._h = lib.b200_pc_create(n_blocks, block_size)
        self.n_blocks, self.block_size = n_blocks, block_size

    def __del__(self):
        if getattr(self, "_h", None):
            lib.b200_pc_destroy(self._h)
            self._h = None

    def append(self, seq: int, n_new: int):
        """-> (slot_mapping list, cow_pairs list of (src,dst)) or None when out of blocks"""
        slots = (C.c_int32 * max(n_new, 1))()
        cow = (C.c_int32 * 2)()
        n = lib.b200_pc_append(self._h, seq, n_new, slots, cow, 1)
        if n < 0:
            return None
        return [slots[i] for i in range(n_new)], [(cow[2 * i], cow[2 * i + 1]) for i in range(n)]

    def prepare_batch(self, seq, pos=None):
        """one ubatch (llama_context::decode's find_slot + mask fill, paged form): -> dict(slot_mapping, q_seq, q_pos, seqs,
        cow_pairs), or the status 1 (no room) / -1 (positions do not continue their sequence); see include/b200_host.h"""
        n = len(seq)
        arr = lambda v: (C.c_int32 * max(n, 1))(*v)
        s_in, p_in = arr(seq), (arr(pos) if pos is not None else None)
        slots, q_seq, q_pos, seqs, cow = (C.c_int32 * max(n, 1))(), (C.c_int32 * max(n, 1))(), (C.c_int32 * max(n, 1))(), \
            (C.c_int32 * max(n, 1))(), (C.c_int32 * (2 * max(n, 1)))()
        n_seqs, n_cow = C.c_int64(0), C.c_int64(0)
        rc = lib.b200_pc_prepare_batch(self._h, n, s_in, p_in, slots, q_seq, q_pos, seqs, C.byref(n_seqs), cow, max(n, 1),
                                       C.byref(n_cow))
        if rc != 0:
            return rc
        return dict(slot_mapping=list(slots[:n]), q_seq=list(q_seq[:n]), q_pos=list(q_pos[:n]), seqs=list(seqs[:n_seqs.value]),
                    cow_pairs=[(cow[2 * i], cow[2 * i + 1]) for i in range(n_cow.value)])

    def undo_batch(self, seq, q_pos, cow_pairs=()) -> None:
        """llama_kv_slot_restorer::restore for a batch prepare_batch placed (its seq list, the q_pos and cow_pairs it returned)"""
        n = len(seq)
        flat = [x for pr in cow_pairs for x in pr]
        lib.b200_pc_undo_batch(self._h, n, (C.c_int32 * max(n, 1))(*seq), (C.c_int32 * max(n, 1))(*q_pos),
                               (C.c_int32 * max(len(flat), 1))(*flat), len(flat) // 2)

    def seq_cp(self, src: int, dst: int, p0: int = 0, p1: int = -1) -> bool:
        return bool(lib.b200_pc_seq_cp(self._h, src, dst, p0, p1))

    def seq_rm(self, seq: int, p0: int = 0, p1: int = -1) -> bool:
        return bool(lib.b200_pc_seq_rm(self._h, seq, p0, p1))

    def seq_keep(self, seq: int) -> None:
        lib.b200_pc_seq_keep(self._h, seq)

    def clear(self) -> None:
        lib.b200_pc_clear(self._h)

    def seq_len(self, seq: int) -> int:
        return lib.b200_pc_seq_len(self._h, seq)

    def n_free(self) -> int:
        return lib.b200_pc_n_free(self._h)

    def ref_count(self, block: int) -> int:
        return lib.b200_pc_ref_count(self._h, block)

    def used_cells(self):
        """every cell some sequence holds, ascending: the row order of the whole-cache stream"""
        n = -lib.b200_pc_used_cells(self._h, None, 0) - 1
        buf = (C.c_int32 * max(n, 1))()
        n = lib.b200_pc_used_cells(self._h, buf, max(n, 1))
        return list(buf[:n])

    def slots_of(self, seq: int):
        """cell of every cached position, in order, or None for an unknown sequence"""
        n = max(self.seq_len(seq), 1)
        buf = (C.c_int32 * n)()
        r = lib.b200_pc_slots_of(self._h, seq, buf, n)
        return None if r < 0 else [buf[i] for i in range(r)]

    def defrag(self, max_moves: int = -1):
        """compaction plan: list of (src_block, dst_block) for block_copy; block tables are already rewritten"""
        cap = max(self.n_blocks, 1)
        buf = (C.c_int32 * (2 * cap))()
        n = lib.b200_pc_defrag(self._h, buf, cap, max_moves)
        return [(buf[2 * i], buf[2 * i + 1]) for i in range(max(n, 0))]

    def context_shift(self, seq: int, p0: int, p1: int):
        """drop positions [p0, p1) of `seq`, later tokens move down -> (cow_pairs, row_moves, shift_cells) or None when refused"""
        counts = (C.c_int64 * 3)()
        rc = lib.b200_pc_context_shift(self._h, seq, p0, p1, None, 0, None, 0, None, 0, counts)
        if rc == 0:
            return [], [], []
        if rc != -3:
            return None
        nc, nm, ns = counts[0], counts[1], counts[2]
        cow, mv, cells = (C.c_int32 * max(2 * nc, 1))(), (C.c_int32 * max(2 * nm, 1))(), (C.c_int32 * max(ns, 1))()
        rc = lib.b200_pc_context_shift(self._h, seq, p0, p1, cow, nc, mv, nm, cells, ns, counts)
        if rc != 0:
            return None
        return ([(cow[2 * i], cow[2 * i + 1]) for i in range(counts[0])], [(mv[2 * i], mv[2 * i + 1]) for i in range(counts[1])],
                [cells[i] for i in range(counts[2])])

    def restore_seq(self, seq: int, n_cells: int):
        """drop `seq` and reserve n_cells tokens for it -> their cells, or None when the pool has no room"""
        buf = (C.c_int32 * max(n_cells, 1))()
        r = lib.b200_pc_restore_seq(self._h, seq, n_cells, buf)
        return None if r < 0 else [buf[i] for i in range(r)]

    def export(self, seqs: list[int], stride: int):
        """-> (block_table rows as list of lists, seq_lens list) or None if stride too small"""
        n = len(seqs)
        ids = (C.c_int32 * max(n, 1))(*seqs)
        bt = (C.c_int32 * max(n * stride, 1))()
        lens = (C.c_int32 * max(n, 1))()
        if not lib.b200_pc_export(self._h, ids, n, bt, stride, lens):
            return None
        return [[bt[i * stride + j] for j in range(stride)] for i in range(n)], [lens[i] for i in range(n)]


class LlamaKVCachePaged:
    """llama_kv_cache_paged through the C ABI (b200_kvc_*): the reference's llama_kv_cache method surface on the block manager"""

    def __init__(self, n_blocks: int, block_size: int):
        self._h = lib.b200_kvc_create(n_blocks, block_size)
        self.n_blocks, self.block_size = n_blocks, block_size

    def __del__(self):
        if getattr(self, "_h", None):
            lib.b200_kvc_destroy(self._h)
            self._h = None

    def clear(self): lib.b200_kvc_clear(self._h)
    def defrag(self): lib.b200_kvc_defrag(self._h)
    def seq_rm(self, seq, p0, p1): return bool(lib.b200_kvc_seq_rm(self._h, seq, p0, p1))
    def seq_cp(self, src, dst, p0, p1): lib.b200_kvc_seq_cp(self._h, src, dst, p0, p1)
    def seq_keep(self, seq): lib.b200_kvc_seq_keep(self._h, seq)
    def seq_add(self, seq, p0, p1, delta): lib.b200_kvc_seq_add(self._h, seq, p0, p1, delta)
    def seq_div(self, seq, p0, p1, d): lib.b200_kvc_seq_div(self._h, seq, p0, p1, d)
    def seq_pos_max(self, seq): return lib.b200_kvc_seq_pos_max(self._h, seq)
    def get_n_tokens(self): return lib.b200_kvc_n_tokens(self._h)
    def get_used_cells(self): return lib.b200_kvc_used_cells(self._h)
    def unsupported(self): return bool(lib.b200_kvc_unsupported(self._h))

    def slots_of(self, seq):
        n = lib.b200_kvc_slots_of(self._h, seq, None, 0)
        if n == -1:
            return None
        n = max(-n - 1, 0) if n < 0 else n
        buf = (C.c_int32 * max(n, 1))()
        n = lib.b200_kvc_slots_of(self._h, seq, buf, max(n, 1))
        return list(buf[:n])

    def find_slot(self, seq, pos=None):
        """find_slot(ubatch) -> dict like PagedKVCache.prepare_batch, or the status"""
        n = len(seq)
        arr = lambda v: (C.c_int32 * max(n, 1))(*v)
        out = [(C.c_int32 * max(n, 1))() for _ in range(4)]
        cow, n_seqs, n_cow = (C.c_int32 * (2 * max(n, 1)))(), C.c_int64(0), C.c_int64(0)
        rc = lib.b200_kvc_find_slot(self._h, n, arr(seq), arr(pos) if pos is not None else None, out[0], out[1], out[2], out[3],
                                    C.byref(n_seqs), cow, max(n, 1), C.byref(n_cow))
        if rc != 0:
            return rc
        return dict(slot_mapping=list(out[0][:n]), q_seq=list(out[1][:n]), q_pos=list(out[2][:n]), seqs=list(out[3][:n_seqs.value]),
                    cow_pairs=[(cow[2 * i], cow[2 * i + 1]) for i in range(n_cow.value)])

    def update(self):
        """kv_self_update -> dict(cow_pairs, row_moves, shift_cells, shift_delta, groups, defrag_moves), or None when
        unsupported.  groups = [(cow_end, move_end, shift_end)] per context shift, to be executed one after another."""
        counts = (C.c_int64 * 5)()
        rc = lib.b200_kvc_update(self._h, None, None, None, None, None, None, counts)
        if rc == 0:
            return None
        c = [int(x) for x in counts[:4]]
        ends = (C.c_int64 * max(3 * int(counts[4]), 1))()
        n_groups = lib.b200_kvc_update_groups(self._h, ends, int(counts[4])) if rc == -3 else 0
        groups = [(int(ends[3 * g]), int(ends[3 * g + 1]), int(ends[3 * g + 2])) for g in range(max(n_groups, 0))]
        bufs = [(C.c_int32 * max(2 * c[0], 1))(), (C.c_int32 * max(2 * c[1], 1))(), (C.c_int32 * max(c[2], 1))(),
                (C.c_int32 * max(c[2], 1))(), (C.c_int32 * max(2 * c[3], 1))()]
        if rc == -3:
            caps = (C.c_int64 * 4)(*c)
            rc = lib.b200_kvc_update(self._h, bufs[0], bufs[1], bufs[2], bufs[3], bufs[4], caps, counts)
        if rc != 1:
            return None
        pairs = lambda b, k: [(b[2 * i], b[2 * i + 1]) for i in range(k)]
        return dict(cow_pairs=pairs(bufs[0], c[0]), row_moves=pairs(bufs[1], c[1]), shift_cells=list(bufs[2][:c[2]]),
                    shift_delta=list(bufs[3][:c[2]]), groups=groups, defrag_moves=pairs(bufs[4], c[3]))
