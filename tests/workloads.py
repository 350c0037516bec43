"""tests/workloads.py — seeded synthetic inputs for the paged-attention path (numpy, CPU).

Inputs follow SURVEY.md §8(d): Q, K, V ~ U(-1, 1) like init_tensor_uniform of the reference's
tests/test-backend-ops.cpp:39-110 with the fixed seed the reference left commented out (:51), block
table = a seeded random permutation of physical blocks so gathers are non-contiguous, scale = 1/sqrt(d).
K/V are quantised with the ORACLE's from_float (bit-identical to the reference's *_ref quantisers,
tests/test_oracle_vs_ref.py) so checker and kernel see the same cache bytes.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

F32, F16, Q4_0, Q8_0 = 0, 1, 2, 8


def row_bytes(t: int, n: int) -> int:
    return {F32: n * 4, F16: n * 2, Q8_0: n // 32 * 34, Q4_0: n // 32 * 18}[t]


@dataclass
class Workload:
    n_head: int
    n_head_kv: int
    d: int
    block_size: int
    n_blocks: int
    type_k: int
    type_v: int
    scale: float
    q: np.ndarray                 # F32 [n_q, n_head, d]
    k_cache: np.ndarray           # uint8 [n_cells, token_bytes_k]
    v_cache: np.ndarray           # uint8 [n_cells, token_bytes_v]
    block_table: np.ndarray       # int32 [n_seqs, max_blocks]
    seq_lens: np.ndarray          # int32 [n_seqs]
    q_seq: np.ndarray | None = None
    q_pos: np.ndarray | None = None
    k_f32: np.ndarray | None = None     # [total_tokens, n_head_kv*d] source rows (for kv_write tests)
    v_f32: np.ndarray | None = None
    slot_mapping: np.ndarray | None = None   # int32 [total_tokens]
    extra: dict = field(default_factory=dict)

    @property
    def n_cells(self) -> int:
        return self.n_blocks * self.block_size

    @property
    def n_q(self) -> int:
        return self.q.shape[0]

    def kv_bytes(self) -> int:
        """algorithmic KV bytes of one op call: every cached element once, tails excluded"""
        tok = int(self.seq_lens.astype(np.int64).sum())
        return tok * self.n_head_kv * (row_bytes(self.type_k, self.d) + row_bytes(self.type_v, self.d))

    def dense_mask(self) -> np.ndarray:
        """the reference's dense F16 mask [pad64(n_q), n_cells] hiding every cell a row must not see
        (llm_graph_input_attn_kv_unified::set_input, src/llama-graph.cpp:403-477)"""
        n_q = self.n_q
        pad = (n_q + 63) // 64 * 64
        m = np.full((pad, self.n_cells), -np.inf, dtype=np.float16)
        for i in range(n_q):
            s = i if self.q_seq is None else int(self.q_seq[i])
            n = int(self.seq_lens[s])
            if self.q_pos is not None:
                n = min(n, int(self.q_pos[i]) + 1)
            pos = np.arange(n)
            cells = self.block_table[s, pos // self.block_size].astype(np.int64) * self.block_size + pos % self.block_size
            m[i, cells] = 0.0
        return m


def make_workload(oracle, seq_lens, n_head=32, n_head_kv=8, d=128, type_k=F16, type_v=None, block_size=16,
                  seed=1234, dist="uniform", spare_blocks=3, shuffle=True, q_rows=None) -> Workload:
    """q_rows: None => decode (one query per sequence, sees everything); else list of (seq, pos) query rows."""
    type_v = type_k if type_v is None else type_v
    rng = np.random.default_rng(seed)
    seq_lens = np.asarray(seq_lens, dtype=np.int32)
    n_seqs = len(seq_lens)
    blocks_per_seq = (seq_lens.astype(np.int64) + block_size - 1) // block_size
    max_blocks = max(int(blocks_per_seq.max()) if n_seqs else 1, 1)
    n_blocks = int(blocks_per_seq.sum()) + spare_blocks
    ids = np.arange(n_blocks, dtype=np.int32)
    if shuffle:
        np.random.default_rng(42).shuffle(ids)
    block_table = np.zeros((n_seqs, max_blocks), dtype=np.int32)
    o = 0
    for s in range(n_seqs):
        nb = int(blocks_per_seq[s])
        block_table[s, :nb] = ids[o:o + nb]
        o += nb
    total = int(seq_lens.astype(np.int64).sum())

    def draw(shape):
        if dist == "normal":
            return rng.standard_normal(shape, dtype=np.float32)
        return rng.uniform(-1.0, 1.0, shape).astype(np.float32)

    n_q = n_seqs if q_rows is None else len(q_rows)
    q = draw((n_q, n_head, d))
    k_f32 = draw((total, n_head_kv * d))
    v_f32 = draw((total, n_head_kv * d))
    slot = np.empty(total, dtype=np.int32)
    o = 0
    for s in range(n_seqs):
        n = int(seq_lens[s])
        pos = np.arange(n)
        slot[o:o + n] = block_table[s, pos // block_size] * block_size + pos % block_size
        o += n
    tk, tv = row_bytes(type_k, d) * n_head_kv, row_bytes(type_v, d) * n_head_kv
    n_cells = n_blocks * block_size
    # unused cells hold a poison pattern, not zeros: a kernel that reads past seq_len shows up
    k_cache = np.full((n_cells, tk), 0x7b, dtype=np.uint8)
    v_cache = np.full((n_cells, tv), 0x7b, dtype=np.uint8)
    if total:
        k_cache[slot] = oracle.from_float(type_k, k_f32.reshape(total * n_head_kv, d)).reshape(total, tk)
        v_cache[slot] = oracle.from_float(type_v, v_f32.reshape(total * n_head_kv, d)).reshape(total, tv)
    q_seq = q_pos = None
    if q_rows is not None:
        q_seq = np.asarray([r[0] for r in q_rows], dtype=np.int32)
        q_pos = np.asarray([r[1] for r in q_rows], dtype=np.int32)
    return Workload(n_head, n_head_kv, d, block_size, n_blocks, type_k, type_v, float(1.0 / np.sqrt(d)), q, k_cache,
                    v_cache, block_table, seq_lens, q_seq, q_pos, k_f32, v_f32, slot)


def mixed_ctx(n_seqs=64, lo=512, hi=8192, seed=7):
    """config 2's context lengths: ctx_i ~ U{lo..hi}, seed 7"""
    return np.random.default_rng(seed).integers(lo, hi + 1, size=n_seqs).astype(np.int32)


def oracle_attention(oracle, w: Workload) -> np.ndarray:
    return oracle.attention(w.q, w.k_cache, w.v_cache, w.type_k, w.type_v, w.n_head_kv, w.scale,
                            block_table=w.block_table, seq_lens=w.seq_lens, block_size=w.block_size,
                            q_seq=w.q_seq, q_pos=w.q_pos)
