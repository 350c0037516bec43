"""KV session state on the paged layout and block defrag — host side, no GPU.

* the stream the product writes for one sequence equals, byte for byte, what llama_kv_cache_unified::state_write
  produces for the same tokens held in a flat unified cache (oracle/kv_state_oracle.py restates
  src/llama-kv-cache.cpp:848-971 over the reference's own cell layout);
* the parser rejects exactly the streams state_read_meta / state_read_data reject (:1008-1240), for the same reason;
* restore bookkeeping = seq_rm + find_slot; the defrag plan compacts blocks without changing what any sequence sees.
"""
import struct

import numpy as np
import pytest

from oracle.kv_state_oracle import unified_state_read_errors, unified_state_write
from workloads import F16, Q4_0, Q8_0, row_bytes

N_EMBD_GQA = 1024          # 8 kv heads x 128 (Llama-3-8B)


def _rows(rng, n_layer, n, size_row):
    return rng.integers(0, 256, size=(n_layer, n, size_row), dtype=np.uint8)


@pytest.mark.parametrize("tk,tv", [(F16, F16), (Q8_0, Q8_0), (Q4_0, Q4_0), (Q8_0, F16)])
@pytest.mark.parametrize("placement", ["contiguous", "two_ranges", "empty"])
def test_stream_equals_reference_writer(pkg, tk, tv, placement):
    rng = np.random.default_rng(3)
    n_layer, size = 3, 96
    k_row, v_row = row_bytes(tk, N_EMBD_GQA), row_bytes(tv, N_EMBD_GQA)
    # the reference side: a flat unified cache; sequence 5 holds n tokens, another sequence sits in between
    if placement == "contiguous":
        cell_of = list(range(17, 17 + 40))
    elif placement == "two_ranges":
        cell_of = list(range(4, 20)) + list(range(33, 60))
    else:
        cell_of = []
    n = len(cell_of)
    cells = [(-1, set()) for _ in range(size)]
    for c in range(20, 33):
        cells[c] = (c - 20, {9})
    for p, c in enumerate(cell_of):
        cells[c] = (p, {5} if p >= 8 else {5, 9})                 # a shared prefix: cells may carry several ids
    k_flat = [rng.integers(0, 256, size=(size, k_row), dtype=np.uint8) for _ in range(n_layer)]
    v_flat = [rng.integers(0, 256, size=(size, v_row), dtype=np.uint8) for _ in range(n_layer)]
    want = unified_state_write(cells, 5, k_flat, v_flat, tk, tv)
    # the product side: the same rows in token order (what b200_kv_rows_gather delivers)
    k_tok = np.stack([k[cell_of] for k in k_flat]) if n else np.zeros((n_layer, 0, k_row), np.uint8)
    v_tok = np.stack([v[cell_of] for v in v_flat]) if n else np.zeros((n_layer, 0, v_row), np.uint8)
    got = pkg.state_seq_pack(k_tok, v_tok, tk, tv)
    assert got == want
    assert len(got) == pkg.lib.b200_kv_state_seq_size(n, n_layer, k_row, v_row)
    k2, v2 = pkg.state_seq_unpack(got, n_layer, tk, k_row, tv, v_row)
    assert np.array_equal(k2, k_tok) and np.array_equal(v2, v_tok)
    assert unified_state_read_errors(got, n_layer, tk, k_row, tv, v_row) is None


def test_parser_rejects_what_the_reference_rejects(pkg):
    rng = np.random.default_rng(4)
    n_layer, n = 2, 5
    k_row, v_row = row_bytes(Q8_0, N_EMBD_GQA), row_bytes(F16, N_EMBD_GQA)
    good = pkg.state_seq_pack(_rows(rng, n_layer, n, k_row), _rows(rng, n_layer, n, v_row), Q8_0, F16)
    meta_end = 4 + 8*n
    k0 = meta_end + 8                                             # first key header
    v0 = k0 + n_layer*(12 + n*k_row)                              # first value header

    def patched(off, fmt, val):
        b = bytearray(good)
        struct.pack_into(fmt, b, off, val)
        return bytes(b)

    cases = {
        "invalid seq_id-agnostic kv cell": patched(4 + 8*2 + 4, "<I", 1),
        "mismatched layer count": patched(meta_end + 4, "<I", 3),
        "incompatible V transposition": patched(meta_end, "<I", 1),
        "mismatched key type": patched(k0, "<i", F16),
        "mismatched key row size": patched(k0 + 4, "<Q", k_row + 34),
        "mismatched value type": patched(v0, "<i", Q8_0),
        "mismatched value row size": patched(v0 + 4, "<Q", v_row - 2),
        "truncated": good[:-7],
    }
    for reason, blob in cases.items():
        assert unified_state_read_errors(blob, n_layer, Q8_0, k_row, F16, v_row) == reason   # the oracle agrees it is bad
        with pytest.raises(pkg.B200Error) as e:
            pkg.state_seq_unpack(blob, n_layer, Q8_0, k_row, F16, v_row)
        msg = str(e.value)
        assert msg.startswith("failed to restore kv cache")       # the reference's exception text, llama-kv-cache.cpp:895
        ours = {"mismatched key type": "mismatched key/value type", "mismatched value type": "mismatched key/value type",
                "mismatched key row size": "mismatched key/value row size", "mismatched value row size": "mismatched key/value row size",
                "truncated": "truncated stream"}.get(reason, reason)
        assert ours in msg
    # shape of the cache differs from the stream's: rejected as the reference would
    with pytest.raises(pkg.B200Error):
        pkg.state_seq_unpack(good, n_layer, Q8_0, k_row, Q8_0, row_bytes(Q8_0, N_EMBD_GQA))
    # positions that are not 0..n-1 cannot live in a block table
    with pytest.raises(pkg.B200Error, match="positions"):
        pkg.state_seq_unpack(patched(4 + 8*3, "<i", 7), n_layer, Q8_0, k_row, F16, v_row)
    with pytest.raises(pkg.B200Error, match="positions"):
        pkg.state_seq_unpack(patched(4 + 8*3, "<i", 2), n_layer, Q8_0, k_row, F16, v_row)      # position 2 twice, 3 missing


def test_stream_rows_in_cell_order_not_position_order(pkg):
    """llama_kv_cache_unified::state_write(seq_id) walks the cells in cell-index order (src/llama-kv-cache.cpp:856-872), which
    is not position order once find_slot's head has wrapped, cells were reused after a seq_rm or a defrag moved them.  Such a
    stream — the positions 0..n-1 in any row order — restores to the same rows, by position."""
    rng = np.random.default_rng(12)
    n, n_layer = 9, 2
    k_row, v_row = row_bytes(Q4_0, N_EMBD_GQA), row_bytes(Q8_0, N_EMBD_GQA)
    k, v = _rows(rng, n_layer, n, k_row), _rows(rng, n_layer, n, v_row)
    perm = rng.permutation(n)                                   # stream row i holds position perm[i]
    blob = bytearray(pkg.state_seq_pack(k[:, perm], v[:, perm], Q4_0, Q8_0))
    for i in range(n):
        struct.pack_into("<i", blob, 4 + 8*i, int(perm[i]))
    assert unified_state_read_errors(bytes(blob), n_layer, Q4_0, k_row, Q8_0, v_row) is None    # a stream the reference reads
    k2, v2 = pkg.state_seq_unpack(bytes(blob), n_layer, Q4_0, k_row, Q8_0, v_row)
    assert np.array_equal(k2, k) and np.array_equal(v2, v)


def test_restore_seq_is_seq_rm_plus_find_slot(pkg):
    c = pkg.PagedKVCache(8, 4)
    c.append(0, 10)                      # blocks 0,1,2
    c.append(1, 6)                       # blocks 3,4
    assert c.seq_cp(1, 2, 0, -1)         # 2 shares blocks 3,4
    cells = c.restore_seq(0, 7)          # old blocks of 0 come back first: lowest free indices again
    assert cells == [0, 1, 2, 3, 4, 5, 6] and c.seq_len(0) == 7 and c.n_free() == 8 - 2 - 2
    assert c.restore_seq(1, 3) == [8, 9, 10]                      # blocks 3,4 stay with seq 2; new block 2
    assert c.ref_count(3) == 1 and c.ref_count(4) == 1 and c.seq_len(2) == 6
    assert c.slots_of(2) == [12, 13, 14, 15, 16, 17]
    # no room: like the reference, the destination sequence is gone afterwards (state_read: seq_rm, then throw)
    assert c.restore_seq(0, 4*8) is None
    assert c.seq_len(0) == 0 and c.slots_of(0) is None
    assert c.restore_seq(7, 0) == [] and c.seq_len(7) == 0


def test_defrag_compacts_and_preserves_every_sequence(pkg):
    rng = np.random.default_rng(11)
    bs, nb = 4, 64
    c = pkg.PagedKVCache(nb, bs)
    store = np.full(nb*bs, -1, dtype=np.int64)                    # what each cell holds: a unique token id
    next_tok, live = 0, {}
    for step in range(200):
        op = rng.integers(0, 4)
        seq = int(rng.integers(0, 10))
        if op <= 1:
            n = int(rng.integers(1, 9))
            r = c.append(seq, n)
            if r is None:
                continue
            slots, cow = r
            for s, d in cow:
                store[d*bs:(d + 1)*bs] = store[s*bs:(s + 1)*bs]
            for s in slots:
                store[s] = next_tok; next_tok += 1
            live[seq] = True
        elif op == 2 and live:
            victim = int(rng.choice(list(live)))
            c.seq_rm(victim); live.pop(victim)
        elif op == 3 and live:
            src = int(rng.choice(list(live)))
            if src != seq and c.seq_cp(src, seq, 0, int(rng.integers(1, c.seq_len(src) + 1))):
                live[seq] = True
    assert len(live) >= 3
    before = {s: store[c.slots_of(s)].copy() for s in live}
    refs_before = sorted(c.ref_count(b) for b in range(nb) if c.ref_count(b) > 0)
    n_used = nb - c.n_free()
    assert any(c.ref_count(b) == 0 for b in range(n_used)), "the op stream left no hole to compact"

    limited = c.defrag(max_moves=1)                               # the reference caps moves per graph (:724-729)
    assert len(limited) == 1
    moves = limited + c.defrag()
    srcs, dsts = [m[0] for m in moves], [m[1] for m in moves]
    assert not set(srcs) & set(dsts), "a block is both source and destination"
    assert all(s >= n_used > d for s, d in moves)
    for s, d in moves:                                            # what b200_block_copy does on the device
        store[d*bs:(d + 1)*bs] = store[s*bs:(s + 1)*bs]
    assert [c.ref_count(b) > 0 for b in range(nb)] == [b < n_used for b in range(nb)]
    assert sorted(c.ref_count(b) for b in range(n_used)) == refs_before
    assert c.n_free() == nb - n_used
    for s in live:
        assert np.array_equal(store[c.slots_of(s)], before[s]), f"sequence {s} changed"
    assert c.defrag() == []                                       # idempotent
    r = c.append(99, bs)                                          # the allocator still hands out the lowest free block
    assert r is not None and r[0][0] == n_used*bs


def test_parser_on_damaged_streams_under_address_sanitizer(tmp_path):
    """tests/cpp/kv_state_fuzz.cpp: every truncation and 4000 seeded corruptions of a valid stream, each in a heap buffer of
    exactly its own length (a read past the end faults under ASan); the parser answers 0 or a B200_STATE_ERR_* and the
    offsets it hands out stay inside the buffer — the reference throws on the same conditions (llama-kv-cache.cpp:1005-1240)"""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    host = os.path.join(root, "llama.cpp_test_b200", "host")
    exe = str(tmp_path / "kv_state_fuzz")
    base = ["g++", "-std=c++17", "-g", "-O1", "-Wall", "-Wextra", "-I", os.path.join(root, "include"), "-I", os.path.join(host, "src"),
            os.path.join(root, "tests", "cpp", "kv_state_fuzz.cpp"), os.path.join(host, "kv_state.cpp"),
            os.path.join(host, "paged_kv_cache.cpp"), os.path.join(host, "host_capi.cpp"), os.path.join(host, "fork_compat.cpp"),
            os.path.join(host, "llama_kv_cache_paged.cpp"), "-o", exe]
    r = subprocess.run(base + ["-fsanitize=address,undefined", "-fno-omit-frame-pointer"], capture_output=True, text=True)
    if r.returncode != 0:
        r = subprocess.run(base, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    run = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert run.returncode == 0 and run.stdout.strip().endswith("OK"), (run.stdout[-1500:], run.stderr[-3000:])
    assert "runtime error" not in run.stderr, run.stderr[-3000:]


# ---- the whole-cache stream: llama_state_get_data / llama_state_set_data (state_write / state_read with seq_id == -1)

def _unified_read_all(data, n_layer, k_row, v_row):
    """state_read_meta's whole-cache branch (src/llama-kv-cache.cpp:1038-1087) + state_read_data, restated: the cells
    (pos, ids) in stream order and each layer's rows"""
    off = 0
    (count,) = struct.unpack_from("<I", data, off); off += 4
    cells = []
    for _ in range(count):
        pos, n = struct.unpack_from("<iI", data, off); off += 8
        ids = set(struct.unpack_from(f"<{n}i", data, off)); off += 4 * n
        cells.append((pos, ids))
    off += 8
    k, v = [], []
    for rows, size_row in ((k, k_row), (v, v_row)):
        for _ in range(n_layer):
            off += 12
            rows.append(np.frombuffer(data, np.uint8, count * size_row, off).reshape(count, size_row)); off += count * size_row
    assert off == len(data)
    return cells, k, v


def _visible_rows(cells, rows, seq):
    """what sequence `seq` sees, in position order"""
    idx = sorted((pos, i) for i, (pos, ids) in enumerate(cells) if seq in ids)
    assert [p for p, _ in idx] == list(range(len(idx)))
    return rows[[i for _, i in idx]] if idx else rows[:0]


@pytest.mark.parametrize("bs", [1, 4, 16])
def test_whole_cache_stream_equals_reference_writer(pkg, bs):
    """a paged cache with forks (shared blocks), a truncated and a removed sequence: its stream is byte for byte what
    state_write(io, -1) writes for a unified cache whose cells are the paged cells"""
    rng = np.random.default_rng(bs)
    n_layer, nb = 2, 96
    k_row, v_row = row_bytes(Q8_0, N_EMBD_GQA), row_bytes(F16, N_EMBD_GQA)
    pc = pkg.PagedKVCache(nb, bs)
    pc.append(3, 37); pc.append(0, 5); pc.seq_cp(3, 7, 0, 22); pc.append(7, 9); pc.append(1, 12); pc.seq_rm(1)
    pc.append(3, 2); pc.seq_cp(0, 2); pc.seq_rm(3, 30, -1)
    size = nb * bs
    k_flat = [rng.integers(0, 256, size=(size, k_row), dtype=np.uint8) for _ in range(n_layer)]
    v_flat = [rng.integers(0, 256, size=(size, v_row), dtype=np.uint8) for _ in range(n_layer)]
    cells = [(-1, set()) for _ in range(size)]                 # the unified view of the same cache
    for s in (0, 2, 3, 7):
        for p, c in enumerate(pc.slots_of(s)):
            assert cells[c][0] in (-1, p)
            cells[c] = (p, cells[c][1] | {s})
    want = unified_state_write(cells, -1, k_flat, v_flat, Q8_0, F16)
    used = pc.used_cells()
    assert used == [c for c in range(size) if cells[c][1]]
    got = pkg.state_all_pack(pc, np.stack([k[used] for k in k_flat]), np.stack([v[used] for v in v_flat]), Q8_0, F16)
    assert got == want


@pytest.mark.parametrize("seed", range(4))
def test_whole_cache_restore_from_a_unified_layout(pkg, seed):
    """a stream the reference would write from its own layout — sequences interleaved cell by cell (batched decode),
    a prompt prefix shared by three sequences — restored into blocks: every sequence sees the same rows in the same
    order, whole blocks of the common prefix stay shared"""
    rng = np.random.default_rng(40 + seed)
    bs = int(rng.choice([1, 4, 16]))
    n_layer, size = 2, 400
    k_row, v_row = row_bytes(F16, N_EMBD_GQA), row_bytes(Q4_0, N_EMBD_GQA)
    cells = [(-1, set()) for _ in range(size)]
    prefix = int(rng.integers(20, 50))
    for p in range(prefix):
        cells[10 + p] = (p, {0, 4, 6})                         # shared prompt (llama-batched-bench -pps)
    c, lens = 10 + prefix + int(rng.integers(0, 5)), {0: prefix, 4: prefix, 6: prefix, 9: 0}
    for _ in range(int(rng.integers(30, 60))):                 # decode steps: one cell per sequence per step, interleaved
        for s in (6, 0, 9, 4):
            if rng.random() < 0.8:
                cells[c] = (lens[s], {s}); lens[s] += 1; c += 1
    k_flat = [rng.integers(0, 256, size=(size, k_row), dtype=np.uint8) for _ in range(n_layer)]
    v_flat = [rng.integers(0, 256, size=(size, v_row), dtype=np.uint8) for _ in range(n_layer)]
    data = unified_state_write(cells, -1, k_flat, v_flat, F16, Q4_0)
    st_cells, st_k, st_v = _unified_read_all(data, n_layer, k_row, v_row)
    pc = pkg.PagedKVCache(64 * 16 // bs, bs)
    pc.append(11, 3)                                           # state_read clears whatever was there
    src, dst, k_rows, v_rows = pkg.state_all_restore(pc, data, n_layer, F16, k_row, Q4_0, v_row)
    assert pc.seq_len(11) == 0 and len(set(dst)) == len(dst)
    k_new = [np.zeros((pc.n_blocks * bs, k_row), np.uint8) for _ in range(n_layer)]
    for l in range(n_layer):
        assert np.array_equal(k_rows[l], st_k[l]) and np.array_equal(v_rows[l], st_v[l])
        k_new[l][dst] = k_rows[l][src]                         # what b200_kv_rows_scatter does
    for s, n in lens.items():
        assert pc.seq_len(s) == n
        if n:
            for l in range(n_layer):
                assert np.array_equal(k_new[l][pc.slots_of(s)], _visible_rows(st_cells, st_k[l], s))
    shared = prefix // bs                                      # whole blocks of the prompt are held by all three
    assert sum(pc.ref_count(b) == 3 for b in range(pc.n_blocks)) == shared
    # and the stream written back from the blocks restores to the same thing (round trip through the product's own writer)
    used = pc.used_cells()
    again = pkg.state_all_pack(pc, np.stack([k[used] for k in k_new]), np.stack([np.zeros((len(used), v_row), np.uint8)] * n_layer), F16, Q4_0)
    c2, k2, _ = _unified_read_all(again, n_layer, k_row, v_row)
    for s, n in lens.items():
        if n:
            assert np.array_equal(_visible_rows(c2, k2[0], s), _visible_rows(st_cells, st_k[0], s))


def test_whole_cache_restore_rejects_like_the_reference(pkg):
    n_layer, k_row, v_row = 1, row_bytes(F16, N_EMBD_GQA), row_bytes(F16, N_EMBD_GQA)
    rows = [np.zeros((6, k_row), np.uint8)]

    def stream(cells):
        return unified_state_write(cells + [(-1, set())], -1, rows[:1], rows[:1], F16, F16) if len(cells) == 5 else None

    pc = pkg.PagedKVCache(4, 2)
    pc.append(1, 3)
    ok = [(0, {2}), (1, {2}), (2, {2}), (0, {5}), (1, {5})]
    for bad, msg in (([(0, {2}), (2, {2}), (3, {2}), (0, {5}), (1, {5})], "positions are not 0..n-1"),      # a hole
                     ([(0, {2}), (1, {2}), (1, {2}), (0, {5}), (1, {5})], "positions are not 0..n-1"),      # twice
                     ([(0, {2}), (1, {2}), (2, {2}), (0, {-3}), (1, {5})], "invalid seq_id")):
        with pytest.raises(pkg.B200Error, match=msg):
            pkg.state_all_restore(pc, stream(bad), n_layer, F16, k_row, F16, v_row)
        assert pc.seq_len(1) == 3                              # a parse error changes nothing
    good = stream(ok)
    with pytest.raises(pkg.B200Error, match="truncated"):
        pkg.state_all_restore(pc, good[:-5], n_layer, F16, k_row, F16, v_row)
    with pytest.raises(pkg.B200Error, match="mismatched key/value type"):
        pkg.state_all_restore(pc, good, n_layer, Q8_0, k_row, F16, v_row)
    src, dst, _, _ = pkg.state_all_restore(pc, good, n_layer, F16, k_row, F16, v_row)
    assert pc.seq_len(1) == 0 and pc.seq_len(2) == 3 and pc.seq_len(5) == 2 and src == [0, 1, 2, 3, 4]
    small = pkg.PagedKVCache(2, 2)                             # 3 + 2 tokens need 2 + 1 blocks
    small.append(0, 1)
    with pytest.raises(pkg.B200Error, match="not enough cells"):
        pkg.state_all_restore(small, good, n_layer, F16, k_row, F16, v_row)
    assert small.n_free() == 2                                 # cleared, as the reference leaves it
