"""PagedKVCache (llama.cpp_test_b200/host/paged_kv_cache.cpp) through the C ABI: block tables, slot mapping,
ref-counts and copy-on-write are integer work and must be exact.

The semantic bridge to the reference: llama_kv_cache_unified keeps one cell per token with
a set of sequence ids; seq_cp inserts the destination id into the cells of the source (src/llama-kv-cache.cpp:
244-293), seq_rm erases ids (:173-242), find_slot hands out fresh cells (:453-672), and a query of sequence s at
position p sees every cell whose set contains s and whose pos <= p (mask fill, src/llama-graph.cpp:428-460).
`UnifiedModel` below restates exactly that in a few lines; the property test drives both with the same random
operation stream and demands that every sequence sees the same token ids in the same positions."""
import numpy as np
import pytest


class UnifiedModel:
    """cell-per-token model of llama_kv_cache_unified: cells[i] = (pos, {seq ids}, token id)"""

    def __init__(self):
        self.cells = []

    def append(self, seq, token_ids):
        pos = self.seq_len(seq)
        for t in token_ids:                       # find_slot + cells[].pos / seq_id.insert, llama-kv-cache.cpp:658-667
            self.cells.append([pos, {seq}, t])
            pos += 1

    def seq_cp(self, src, dst, p1):               # llama-kv-cache.cpp:262-292 with p0 = 0
        if src == dst:
            return
        self.seq_rm(dst)                          # the paged cache replaces dst (documented restriction)
        for c in self.cells:
            if src in c[1] and c[0] < p1:
                c[1].add(dst)

    def seq_rm(self, seq, p0=0):                  # llama-kv-cache.cpp:199-240
        for c in self.cells:
            if seq in c[1] and c[0] >= p0:
                c[1].discard(seq)
        self.cells = [c for c in self.cells if c[1]]

    def seq_len(self, seq):                       # seq_pos_max + 1
        return max([c[0] + 1 for c in self.cells if seq in c[1]], default=0)

    def visible(self, seq):
        return [c[2] for c in sorted((c for c in self.cells if seq in c[1]), key=lambda c: c[0])]


def test_slot_mapping_and_first_fit(pkg):
    pc = pkg.PagedKVCache(8, 16)
    slots, cow = pc.append(0, 20)
    assert slots == list(range(20)) and cow == []            # blocks 0, 1 (lowest free index first)
    slots, cow = pc.append(1, 5)
    assert slots == [32 + i for i in range(5)] and cow == [] # block 2
    slots, _ = pc.append(0, 13)                              # fills block 1 (12 more) then opens block 3
    assert slots == list(range(20, 32)) + [48]
    assert pc.seq_len(0) == 33 and pc.seq_len(1) == 5 and pc.n_free() == 4
    bt, lens = pc.export([0, 1, 7], 4)
    assert bt == [[0, 1, 3, 0], [2, 0, 0, 0], [0, 0, 0, 0]] and lens == [33, 5, 0]
    assert pc.export([0], 2) is None                         # row does not fit


def test_fork_shares_blocks_and_copy_on_write(pkg):
    """fork at a non-block-aligned point: full blocks are shared by ref-count, the partially filled tail block
    is copied on the first append of EITHER owner (BASELINE config 3)"""
    pc = pkg.PagedKVCache(16, 16)
    pc.append(0, 40)                                         # blocks 0,1,2 (2 is partial: 8 tokens)
    assert pc.seq_cp(0, 1)
    assert [pc.ref_count(b) for b in range(4)] == [2, 2, 2, 0]
    slots, cow = pc.append(1, 1)                             # seq 1 writes position 40 -> copy block 2 to block 3
    assert cow == [(2, 3)] and slots == [3 * 16 + 8]
    assert [pc.ref_count(b) for b in range(4)] == [2, 2, 1, 1]
    slots, cow = pc.append(0, 1)                             # seq 0 now owns block 2 alone: no copy
    assert cow == [] and slots == [2 * 16 + 8]
    # block-aligned fork: nothing to copy, a fresh block is opened
    pc.clear()
    pc.append(0, 32)
    pc.seq_cp(0, 1)
    slots, cow = pc.append(1, 1)
    assert cow == [] and slots == [2 * 16]
    # prefix fork in the middle of a block
    pc.clear()
    pc.append(0, 40)
    assert pc.seq_cp(0, 2, 0, 20)                            # positions [0, 20): blocks 0 and 1 (1 is partial for seq 2)
    assert pc.seq_len(2) == 20 and [pc.ref_count(b) for b in range(3)] == [2, 2, 1]
    slots, cow = pc.append(2, 2)
    assert cow == [(1, 3)] and slots == [3 * 16 + 4, 3 * 16 + 5]


def test_all_or_nothing_on_exhaustion(pkg):
    pc = pkg.PagedKVCache(2, 16)
    assert pc.append(0, 30) is not None
    assert pc.append(1, 1) is None                           # no free block: llama_decode's "no KV slot" (-3 path... =1)
    assert pc.seq_len(1) == 0 and pc.n_free() == 0
    assert pc.append(0, 2) is not None                       # still fits in the tail block
    assert pc.append(0, 1) is None and pc.seq_len(0) == 32
    pc.seq_cp(0, 1)
    pc.seq_rm(0, 31, -1)                                     # truncate seq 0 to 31 tokens; block 1 still shared
    assert pc.append(0, 1) is None                           # would need a CoW copy and there is no free block
    assert pc.seq_len(0) == 31


def test_seq_rm_and_keep(pkg):
    pc = pkg.PagedKVCache(8, 4)
    pc.append(0, 10); pc.append(1, 3); pc.append(2, 4)
    assert not pc.seq_rm(0, 2, 5)                            # middle range refused
    assert pc.seq_rm(0, 5, -1) and pc.seq_len(0) == 5 and pc.n_free() == 8 - 2 - 1 - 1
    assert pc.seq_rm(1) and pc.seq_len(1) == 0
    pc.seq_keep(2)
    assert pc.seq_len(0) == 0 and pc.seq_len(2) == 4 and pc.n_free() == 7
    assert pc.seq_rm(-1) and pc.n_free() == 8


@pytest.mark.parametrize("seed", range(6))
def test_equivalent_to_unified_cache_visibility(pkg, seed):
    """random append / fork / truncate / remove streams: every sequence must see the same token ids in the same
    order as in the reference's cell-per-token model; CoW pairs are applied to a byte-free 'content' array"""
    rng = np.random.default_rng(seed)
    bs, nb = int(rng.choice([1, 4, 16])), 64
    pc, um = pkg.PagedKVCache(nb, bs), UnifiedModel()
    content = np.full(nb * bs, -1, dtype=np.int64)           # what each physical cell holds (token id)
    next_tok = 0
    for _ in range(300):
        op = rng.choice(["append", "append", "append", "cp", "trunc", "rm"])
        seqs = [s for s in range(6) if pc.seq_len(s) > 0]
        if op == "append":
            s, n = int(rng.integers(0, 6)), int(rng.integers(1, 2 * bs + 2))
            r = pc.append(s, n)
            if r is None:
                continue                                      # pool exhausted: nothing may have changed (checked below)
            slots, cow = r
            for a, b in cow:
                content[b * bs:(b + 1) * bs] = content[a * bs:(a + 1) * bs]
            toks = list(range(next_tok, next_tok + n)); next_tok += n
            content[slots] = toks
            um.append(s, toks)
        elif op == "cp" and seqs:
            src, dst = int(rng.choice(seqs)), int(rng.integers(0, 6))
            p1 = int(rng.integers(1, pc.seq_len(src) + 1))
            assert pc.seq_cp(src, dst, 0, p1)
            um.seq_cp(src, dst, p1)
        elif op == "trunc" and seqs:
            s = int(rng.choice(seqs)); p0 = int(rng.integers(0, pc.seq_len(s) + 1))
            assert pc.seq_rm(s, p0, -1)
            um.seq_rm(s, p0)
        elif op == "rm" and seqs:
            s = int(rng.choice(seqs))
            assert pc.seq_rm(s)
            um.seq_rm(s)
        for s in range(6):
            n = pc.seq_len(s)
            assert n == um.seq_len(s)
            if n:
                bt, lens = pc.export([s], (n + bs - 1) // bs)
                pos = np.arange(n)
                cells = np.asarray(bt[0])[pos // bs] * bs + pos % bs
                assert content[cells].tolist() == um.visible(s)
    # ref-counts are consistent with the tables and nothing leaked
    refs = np.zeros(nb, dtype=np.int64)
    for s in range(6):
        n = pc.seq_len(s)
        if n:
            bt, _ = pc.export([s], (n + bs - 1) // bs)
            refs[bt[0]] += 1
    assert refs.tolist() == [pc.ref_count(b) for b in range(nb)]
    assert pc.n_free() == int((refs == 0).sum())


def test_prepare_batch_interleaved_sequences(pkg):
    """one ubatch with interleaved sequences: cells, query rows and positions, and the row order of the exported table"""
    pc = pkg.PagedKVCache(16, 4)
    pc.append(7, 6)                                           # seq 7 holds 6 tokens: blocks 0, 1
    r = pc.prepare_batch([7, 3, 7, 3, 9], pos=[6, 0, 7, 1, 0])
    assert r["seqs"] == [7, 3, 9] and r["q_seq"] == [0, 1, 0, 1, 2] and r["q_pos"] == [6, 0, 7, 1, 0]
    assert r["slot_mapping"] == [6, 8, 7, 9, 12] and r["cow_pairs"] == []     # seq 7 fills block 1; seq 3 -> block 2; seq 9 -> block 3
    bt, lens = pc.export(r["seqs"], 3)
    assert lens == [8, 2, 1] and bt == [[0, 1, 0], [2, 0, 0], [3, 0, 0]]
    assert pc.prepare_batch([3, 3], pos=[2, 4]) == -1 and pc.seq_len(3) == 2  # a hole: refused, nothing changed
    assert pc.prepare_batch([3], pos=[1]) == -1                               # position already cached
    r = pc.prepare_batch([9, 9, 3])                                           # pos None: positions continue
    assert r["q_pos"] == [1, 2, 2] and r["seqs"] == [9, 3]


def test_prepare_batch_is_all_or_nothing(pkg):
    pc = pkg.PagedKVCache(4, 4)
    pc.append(0, 6)                                           # blocks 0, 1 (1 is half full)
    pc.seq_cp(0, 1)
    before = (pc.n_free(), pc.seq_len(0), pc.seq_len(1), [pc.ref_count(b) for b in range(4)])
    # seq 1 needs a copy-on-write block, seq 2 needs 2 blocks: 3 > 2 free -> "no KV slot", nothing may change
    assert pc.prepare_batch([1, 2, 2, 2, 2, 2]) == 1
    assert before == (pc.n_free(), pc.seq_len(0), pc.seq_len(1), [pc.ref_count(b) for b in range(4)]) and pc.seq_len(2) == 0
    r = pc.prepare_batch([1, 2, 2])                           # CoW (block 1 -> 2) + one block for seq 2
    assert r["cow_pairs"] == [(1, 2)] and r["slot_mapping"] == [2 * 4 + 2, 12, 13] and pc.n_free() == 0


def test_undo_batch_is_the_slot_restorer(pkg):
    """llama_context::decode restores the cache when a ubatch's graph fails (llama_kv_slot_restorer, src/llama-kv-cache.h:
    192-237): the sequences are back to their lengths, new sequences are gone, the blocks are free again"""
    pc = pkg.PagedKVCache(16, 4)
    pc.append(7, 6)
    free0 = pc.n_free()
    seq = [7, 3, 7, 3, 9, 7, 7]
    r = pc.prepare_batch(seq)
    assert pc.seq_len(7) == 10 and pc.seq_len(3) == 2 and pc.seq_len(9) == 1
    pc.undo_batch(seq, r["q_pos"])
    assert pc.seq_len(7) == 6 and pc.seq_len(3) == 0 and pc.seq_len(9) == 0 and pc.n_free() == free0
    assert pc.prepare_batch(seq)["slot_mapping"] == r["slot_mapping"]          # and the same cells come out again
    # a batch that forced a copy-on-write: the device copy may never have run, so the undo hands the sequence its shared
    # block back and frees the copy — the cache is exactly what it was before the batch
    pc = pkg.PagedKVCache(8, 4)
    pc.append(0, 6)
    pc.seq_cp(0, 1)
    state = lambda: (pc.n_free(), pc.seq_len(0), pc.seq_len(1), [pc.ref_count(b) for b in range(8)], pc.export([0, 1], 4))
    before = state()
    r = pc.prepare_batch([1, 1, 1])
    assert r["cow_pairs"] == [(1, 2)] and pc.ref_count(1) == 1 and pc.ref_count(2) == 1
    pc.undo_batch([1, 1, 1], r["q_pos"], r["cow_pairs"])
    assert state() == before


@pytest.mark.parametrize("seed", range(5))
def test_prepare_batch_equals_the_reference_mask(pkg, seed):
    """llm_graph_input_attn_kv_unified::set_input (src/llama-graph.cpp:428-460) lets token (seq s, pos p) see the cells whose
    seq set holds s and whose pos <= p.  The paged inputs must name exactly those cells, in position order."""
    rng = np.random.default_rng(100 + seed)
    bs, nb = int(rng.choice([1, 4, 16])), 96
    pc, um = pkg.PagedKVCache(nb, bs), UnifiedModel()
    content = np.full(nb * bs, -1, dtype=np.int64)
    next_tok = 0
    for _ in range(120):
        op = rng.choice(["batch", "batch", "batch", "cp", "trunc", "rm"])
        seqs = [s for s in range(5) if pc.seq_len(s) > 0]
        if op == "batch":
            n = int(rng.integers(1, 3 * bs + 3))
            seq = [int(s) for s in rng.integers(0, 5, n)]                    # interleaved
            r = pc.prepare_batch(seq)
            if r == 1:
                continue
            for a, b in r["cow_pairs"]:
                content[b * bs:(b + 1) * bs] = content[a * bs:(a + 1) * bs]
            toks = list(range(next_tok, next_tok + n)); next_tok += n
            content[r["slot_mapping"]] = toks
            for s, t in zip(seq, toks):
                um.append(s, [t])
            stride = max((pc.seq_len(s) + bs - 1) // bs for s in r["seqs"])
            bt, lens = pc.export(r["seqs"], stride)
            for i in range(n):                                               # every query row of the batch
                s, p = seq[i], r["q_pos"][i]
                want = [c[2] for c in sorted((c for c in um.cells if s in c[1] and c[0] <= p), key=lambda c: c[0])]
                row = np.asarray(bt[r["q_seq"][i]])
                n_vis = min(lens[r["q_seq"][i]], p + 1)
                pos = np.arange(n_vis)
                assert content[row[pos // bs] * bs + pos % bs].tolist() == want
                assert want[-1] == toks[i]                                   # it sees itself last
        elif op == "cp" and seqs:
            src, dst = int(rng.choice(seqs)), int(rng.integers(0, 5))
            p1 = int(rng.integers(1, pc.seq_len(src) + 1))
            assert pc.seq_cp(src, dst, 0, p1)
            um.seq_cp(src, dst, p1)
        elif op == "trunc" and seqs:
            s = int(rng.choice(seqs)); p0 = int(rng.integers(0, pc.seq_len(s) + 1))
            assert pc.seq_rm(s, p0, -1)
            um.seq_rm(s, p0)
        elif op == "rm" and seqs:
            s = int(rng.choice(seqs))
            assert pc.seq_rm(s)
            um.seq_rm(s)
