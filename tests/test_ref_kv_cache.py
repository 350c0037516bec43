"""Pins the restatements of the reference's KV-cache bookkeeping to the reference itself.

oracle/_ref/refkv is the reference's OWN llama_kv_cache_unified (src/llama-kv-cache.cpp, compiled where it lies with the rest
of libllama's sources by `make -C oracle refkv`; oracle/ref_kv_harness.cpp only constructs it and relays calls).  The same
seeded call stream goes to it, to the cell-per-token model the host-class tests use (tests/test_paged_cache.py::UnifiedModel),
to the product's PagedKVCache, and its session streams are compared byte for byte with oracle/kv_state_oracle.py and with the
product's writers.  Skipped where the binary is absent (no reference tree when the checker was built)."""
import os
import struct
import subprocess

import numpy as np
import pytest

from oracle.kv_state_oracle import unified_state_write
from test_paged_cache import UnifiedModel
from workloads import F16, row_bytes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "oracle", "_ref", "refkv")


class RefKV:
    def __init__(self, n_layer, head_dim, n_head_kv, size, tk=F16, tv=F16):
        if not os.path.exists(EXE):
            pytest.skip("oracle/_ref/refkv not built (reference tree absent when the checker was built)")
        self.p = subprocess.Popen([EXE], stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True)
        assert self.cmd(f"I {n_layer} {head_dim} {n_head_kv} {size} {tk} {tv}") == "ok"

    def cmd(self, line):
        self.p.stdin.write(line + "\n"); self.p.stdin.flush()
        return self.p.stdout.readline().strip()

    def visible(self, seq):
        return [tuple(map(int, e.split(":"))) for e in self.cmd(f"V {seq}").split()[1:]]

    def close(self):
        self.p.stdin.close(); self.p.wait(timeout=10)


def test_seq_pos_max_of_an_empty_sequence_is_zero(pkg):
    """src/llama-kv-cache.cpp:431-441 starts its maximum at 0: the adapter answers the same (found by this pin)"""
    r = RefKV(1, 16, 2, 32)
    assert r.cmd("P 3") == "0"
    r.cmd("A 3 1"); assert r.cmd("P 3") == "0"
    r.cmd("A 3 2"); assert r.cmd("P 3") == "2"
    r.close()


@pytest.mark.parametrize("seed", range(6))
def test_cell_model_and_paged_cache_against_the_reference_cache(pkg, seed):
    """random append (single sequence and interleaved ubatches) / fork / truncate / remove / keep streams: what every sequence
    sees — token ids in position order — is the same in the reference's cache, in the cell model and in the block tables;
    the reference's get_n_tokens equals the sum of the sequence lengths"""
    rng = np.random.default_rng(700 + seed)
    bs = int(rng.choice([1, 4, 16]))
    size, n_seq = 512, 5
    r, um, pc = RefKV(1, 16, 2, size), UnifiedModel(), pkg.PagedKVCache(size // bs + 64, bs)
    content = np.full((size // bs + 64) * bs, -1, dtype=np.int64)
    next_tok = 0
    for step in range(150):
        op = rng.choice(["append", "append", "batch", "batch", "cp", "trunc", "rm", "keep"])
        live = [s for s in range(n_seq) if pc.seq_len(s) > 0]
        if op in ("append", "batch"):
            seq = [int(rng.integers(0, n_seq))] * int(rng.integers(1, 20)) if op == "append" else \
                  [int(s) for s in rng.integers(0, n_seq, int(rng.integers(1, 24)))]
            ans = r.cmd("B %d %s" % (len(seq), " ".join(map(str, seq))))
            if ans == "noslot":                                # the flat cache fragments; the paged one need not follow
                continue
            got = pc.prepare_batch(seq)
            assert isinstance(got, dict)
            for a, b in got["cow_pairs"]:
                content[b * bs:(b + 1) * bs] = content[a * bs:(a + 1) * bs]
            toks = list(range(next_tok, next_tok + len(seq))); next_tok += len(seq)
            content[got["slot_mapping"]] = toks
            for s, t in zip(seq, toks):
                um.append(s, [t])
        elif op == "cp" and live:
            src, dst = int(rng.choice(live)), int(rng.integers(0, n_seq))
            if dst == src:
                continue
            p1 = int(rng.integers(1, pc.seq_len(src) + 1))
            r.cmd(f"R {dst} -1 -1")                            # the paged seq_cp replaces dst: give the reference the same request
            r.cmd(f"C {src} {dst} 0 {p1}")
            assert pc.seq_cp(src, dst, 0, p1)
            um.seq_cp(src, dst, p1)
        elif op == "trunc" and live:
            s = int(rng.choice(live)); p0 = int(rng.integers(0, pc.seq_len(s) + 1))
            assert r.cmd(f"R {s} {p0} -1") == "1" and pc.seq_rm(s, p0, -1)
            um.seq_rm(s, p0)
        elif op == "rm" and live:
            s = int(rng.choice(live))
            assert r.cmd(f"R {s} -1 -1") == "1" and pc.seq_rm(s)
            um.seq_rm(s)
        elif op == "keep" and live and rng.random() < 0.3:
            s = int(rng.choice(live))
            r.cmd(f"K {s}"); pc.seq_keep(s)
            for o in range(n_seq):
                if o != s:
                    um.seq_rm(o)
        n_tokens = 0
        for s in range(n_seq):
            vis = r.visible(s)
            assert [p for p, _ in vis] == list(range(len(vis)))
            ids = [t for _, t in vis]
            assert ids == um.visible(s)
            n = pc.seq_len(s)
            assert n == len(ids)
            if n:
                assert content[pc.slots_of(s)].tolist() == ids
            n_tokens += n
        assert int(r.cmd("N").split()[0]) == n_tokens
    r.close()


@pytest.mark.parametrize("whole", [False, True])
def test_session_streams_against_the_reference_writer(pkg, whole):
    """state_write(io, seq) / state_write(io, -1) of the reference on a cache with a shared prefix and interleaved decode
    cells == oracle/kv_state_oracle.py on the same cells (this pins the restatement) == the product's writers fed the same rows
    in their own order; and the reference's state_read accepts the product's single-sequence stream"""
    n_layer, hd, nkv, size = 2, 16, 2, 96
    k_row = row_bytes(F16, hd * nkv)
    r = RefKV(n_layer, hd, nkv, size)
    cells = [(-1, set()) for _ in range(size)]
    k0 = np.zeros((size, k_row), np.uint8)
    tok, lens = 0, {}

    def batch(seq):
        nonlocal tok
        head = int(r.cmd("B %d %s" % (len(seq), " ".join(map(str, seq)))).split()[1])
        for i, s in enumerate(seq):
            cells[head + i] = (lens.get(s, 0), {s}); lens[s] = lens.get(s, 0) + 1
            k0[head + i, :4] = np.frombuffer(struct.pack("<i", tok), np.uint8); tok += 1

    batch([0] * 21)
    r.cmd("C 0 4 0 13")
    for c in range(13):
        cells[c] = (cells[c][0], {0, 4})
    lens[4] = 13
    for _ in range(9):
        batch([4, 0, 7])
    k_l = [k0, np.zeros_like(k0)]
    v_l = [np.zeros_like(k0), np.zeros_like(k0)]
    seq = -1 if whole else 4
    got = bytes.fromhex(r.cmd(f"S {seq}").split()[1])
    assert got == unified_state_write(cells, seq, k_l, v_l, F16, F16)
    # the product: the same contents in a paged cache
    pc = pkg.PagedKVCache(64, 4)
    pc.append(0, 21); pc.seq_cp(0, 4, 0, 13)
    where = {}                                                  # (seq, pos) -> flat cell of the reference
    for c, (pos, ids) in enumerate(cells):
        for s in ids:
            where[(s, pos)] = c
    for _ in range(9):
        pc.prepare_batch([4, 0, 7])
    if whole:
        # the paged cache numbers its cells differently (and copies the shared tail block on write): compare what each
        # sequence sees after reading the product's stream back with the reference's reader
        used = pc.used_cells()
        owner = {}
        for s in (0, 4, 7):
            for p, c in enumerate(pc.slots_of(s)):
                owner[c] = (s, p)
        rows = np.stack([np.stack([kl[where[owner[c]]] for c in used]) for kl in k_l])
        mine = pkg.state_all_pack(pc, rows, np.zeros_like(rows), F16, F16)
        r2 = RefKV(n_layer, hd, nkv, size)
        assert r2.cmd(f"L -1 {mine.hex()}") == "ok"
        for s in (0, 4, 7):
            assert r2.visible(s) == r.visible(s)
        r2.close()
        # the other direction: the reference's own whole-cache stream restored into blocks
        pc2 = pkg.PagedKVCache(64, 4)
        src, dst, k_rows, _ = pkg.state_all_restore(pc2, got, n_layer, F16, k_row, F16, k_row)
        flat = np.zeros((64 * 4, k_row), np.uint8)
        flat[dst] = k_rows[0][src]
        for s in (0, 4, 7):
            ids = [int(np.frombuffer(flat[c, :4].tobytes(), "<i4")[0]) for c in pc2.slots_of(s)]
            assert ids == [t for _, t in r.visible(s)]
        assert sum(pc2.ref_count(b) == 2 for b in range(64)) == 13 // 4         # the shared prompt's whole blocks stay shared
    else:
        rows = np.stack([np.stack([kl[where[(4, p)]] for p in range(lens[4])]) for kl in k_l])
        mine = pkg.state_seq_pack(rows, np.zeros_like(rows), F16, F16)
        assert mine == got                                      # byte for byte the reference's stream
        r2 = RefKV(n_layer, hd, nkv, size)
        assert r2.cmd(f"L 9 {mine.hex()}") == "ok"              # and the reference restores it as sequence 9
        assert [t for _, t in r2.visible(9)] == [t for _, t in r.visible(4)]
        r2.close()
    r.close()


@pytest.mark.parametrize("seed", range(5))
def test_context_shift_against_the_reference_pair_of_calls(pkg, seed):
    """llama_kv_self_seq_rm(seq, p0, p1) + llama_kv_self_seq_add(seq, p1, -1, -(p1 - p0)) on the reference's cache (what
    examples/main does on context overflow) against PagedKVCache.context_shift: the surviving tokens and their new positions,
    and the set of cells whose K rows need re-rotation with the delta the reference records (cell.delta, has_shift)"""
    rng = np.random.default_rng(900 + seed)
    bs = int(rng.choice([1, 4, 16]))
    size = 256
    r, pc = RefKV(1, 16, 2, size), pkg.PagedKVCache(size // bs + 16, bs)
    content = np.full((size // bs + 16) * bs, -1, dtype=np.int64)
    n = int(rng.integers(40, 120))
    assert r.cmd(f"A 0 {n}").startswith("slot")
    slots, _ = pc.append(0, n)
    content[slots] = np.arange(n)
    fork = 0
    if rng.random() < 0.5:                                    # a fork first: shared blocks must be copied before they are written
        fork = int(rng.integers(1, n + 1))
        r.cmd(f"C 0 1 0 {fork}"); assert pc.seq_cp(0, 1, 0, fork)
    p0 = int(rng.integers(0, n - 2)); p1 = int(rng.integers(p0 + 1, n))
    assert r.cmd(f"R 0 {p0} {p1}") == "1"
    r.cmd(f"T 0 {p1} -1 {-(p1 - p0)}")
    got = pc.context_shift(0, p0, p1)
    assert got is not None
    cow, moves, shift_cells = got
    for a, b in cow:
        content[b * bs:(b + 1) * bs] = content[a * bs:(a + 1) * bs]
    if moves:
        src, dst = [m[0] for m in moves], [m[1] for m in moves]
        content[dst] = content[src].copy()                     # gather all, then scatter all
    vis = r.visible(0)
    assert [p for p, _ in vis] == list(range(n - (p1 - p0)))
    assert content[pc.slots_of(0)].tolist() == [t for _, t in vis]
    # the rotation work: the reference marks every moved cell with delta = -(p1 - p0)
    e = r.cmd("E 0").split()
    assert e[1] == "1"
    marked = [tuple(map(int, x.split(":"))) for x in e[2:]]
    assert marked == [(p, -(p1 - p0)) for p in range(p0, n - (p1 - p0))]
    cell_of = pc.slots_of(0)
    assert sorted(shift_cells) == sorted(cell_of[p] for p, _ in marked)
    if fork:
        # the fork keeps its own view in the block tables (shared blocks were copied before they were written)
        assert content[pc.slots_of(1)].tolist() == list(range(fork))
        if fork <= p1:                                         # none of its cells was moved: the reference agrees
            assert [t for _, t in r.visible(1)] == list(range(fork))
        # (fork > p1: the unified cache has ONE position per cell, so seq_add moves the shared cells under the fork as well,
        #  src/llama-kv-cache.cpp:347-366 — a side effect the block tables do not reproduce)
    r.close()


@pytest.mark.parametrize("seed", range(5))
def test_prepare_batch_against_the_reference_mask_fill(pkg, seed):
    """The equivalence the whole paged path rests on, against the reference's own code: for every token row of a ubatch the
    cells that llm_graph_input_attn_kv_unified::set_input (src/llama-graph.cpp:403-477) leaves unmasked — after the reference's
    find_slot placed the ubatch and llama_context::decode's kv_self->n was set — hold exactly the tokens that the block-table
    row + q_pos of PagedKVCache::prepare_batch name, in the same order.  Interleaved sequences, forks, truncations."""
    rng = np.random.default_rng(1300 + seed)
    bs = int(rng.choice([1, 4, 16]))
    size, n_seq = 768, 5
    r, pc = RefKV(1, 16, 2, size), pkg.PagedKVCache(size // bs + 64, bs)
    content = np.full((size // bs + 64) * bs, -1, dtype=np.int64)
    next_tok, rows = 0, 0
    for _ in range(60):
        op = rng.choice(["batch", "batch", "batch", "cp", "trunc", "rm"])
        live = [s for s in range(n_seq) if pc.seq_len(s) > 0]
        if op == "batch":
            n = int(rng.integers(1, 40))
            seq = [int(s) for s in rng.integers(0, n_seq, n)] if rng.random() < 0.7 else [int(rng.integers(0, n_seq))] * n
            ans = r.cmd("M %d %s" % (n, " ".join(map(str, seq)))).split(" ")
            if ans[0] == "noslot":
                continue
            got = pc.prepare_batch(seq)
            assert isinstance(got, dict)
            for a, b in got["cow_pairs"]:
                content[b * bs:(b + 1) * bs] = content[a * bs:(a + 1) * bs]
            content[got["slot_mapping"]] = np.arange(next_tok, next_tok + n); next_tok += n
            stride = max((pc.seq_len(s) + bs - 1) // bs for s in got["seqs"])
            bt, lens = pc.export(got["seqs"], stride)
            assert len(ans) == 2 + n
            for i in range(n):
                want = [int(t) for t in ans[2 + i].split(",")]
                row = np.asarray(bt[got["q_seq"][i]])
                pos = np.arange(min(lens[got["q_seq"][i]], got["q_pos"][i] + 1))
                assert content[row[pos // bs] * bs + pos % bs].tolist() == want
                rows += 1
        elif op == "cp" and live:
            src, dst = int(rng.choice(live)), int(rng.integers(0, n_seq))
            if dst != src:
                p1 = int(rng.integers(1, pc.seq_len(src) + 1))
                r.cmd(f"R {dst} -1 -1"); r.cmd(f"C {src} {dst} 0 {p1}")
                assert pc.seq_cp(src, dst, 0, p1)
        elif op == "trunc" and live:
            s = int(rng.choice(live)); p0 = int(rng.integers(0, pc.seq_len(s) + 1))
            assert r.cmd(f"R {s} {p0} -1") == "1" and pc.seq_rm(s, p0, -1)
        elif op == "rm" and live:
            s = int(rng.choice(live))
            assert r.cmd(f"R {s} -1 -1") == "1" and pc.seq_rm(s)
    assert rows > 200
    r.close()


def test_stream_rejections_against_the_reference_reader(pkg):
    """every malformed single-sequence stream the product's parser rejects (by code) is rejected by the reference's own
    state_read ("failed to restore kv cache", src/llama-kv-cache.cpp:886-903), and the well-formed one is accepted by both"""
    from workloads import Q8_0
    rng = np.random.default_rng(4)
    n_layer, hd, nkv, n = 2, 16, 2, 5
    k_row, v_row = row_bytes(Q8_0, hd * nkv), row_bytes(F16, hd * nkv)
    rows = lambda w: rng.integers(0, 256, size=(n_layer, n, w), dtype=np.uint8)
    good = pkg.state_seq_pack(rows(k_row), rows(v_row), Q8_0, F16)
    meta_end = 4 + 8 * n
    k0 = meta_end + 8
    v0 = k0 + n_layer * (12 + n * k_row)

    def patched(off, fmt, val):
        b = bytearray(good)
        struct.pack_into(fmt, b, off, val)
        return bytes(b)

    cases = {
        "invalid seq_id-agnostic kv cell": patched(4 + 8 * 2 + 4, "<I", 1),
        "mismatched layer count": patched(meta_end + 4, "<I", 3),
        "incompatible V transposition": patched(meta_end, "<I", 1),
        "mismatched key/value type": patched(k0, "<i", F16),
        "mismatched key/value row size": patched(v0 + 4, "<Q", v_row - 2),
        "truncated stream": good[:-7],
    }
    r = RefKV(n_layer, hd, nkv, 64, Q8_0, F16)
    assert r.cmd(f"L 3 {good.hex()}") == "ok" and r.cmd("P 3") == str(n - 1)
    pkg.state_seq_unpack(good, n_layer, Q8_0, k_row, F16, v_row)
    for reason, blob in cases.items():
        assert r.cmd(f"L 3 {blob.hex()}").startswith("fail"), reason
        with pytest.raises(pkg.B200Error, match=reason):
            pkg.state_seq_unpack(blob, n_layer, Q8_0, k_row, F16, v_row)
    r.close()


@pytest.mark.parametrize("seed", range(6))
def test_llama_kv_cache_adapter_against_the_reference_call_for_call(pkg, seed):
    """llama_kv_cache_paged (through b200_kvc_*) and the reference's llama_kv_cache_unified receive the same llama_kv_self_*
    calls, name for name — decode ubatches, seq_cp, tail seq_rm, seq_keep, and the context shift as the reference's callers
    issue it (seq_rm of a middle range, then seq_add of the tail by the negative length) — with kv_self_update (b200_kvc_update)
    before every decode as llama_context::decode runs it.  After every call every sequence must show the same token ids in
    position order; get_n_tokens and seq_pos_max must agree.  (Shifts are only issued on sequences that share no cell: the
    unified cache moves a shared cell for every owner, see the context-shift pin above.)"""
    rng = np.random.default_rng(2100 + seed)
    bs = int(rng.choice([1, 4, 16]))
    size, n_seq = 768, 4
    r, kv = RefKV(1, 16, 2, size), pkg.LlamaKVCachePaged(size // bs + 64, bs)
    content = np.full((size // bs + 64) * bs, -1, dtype=np.int64)
    next_tok, shifts = 0, 0
    shared = set()                                            # sequences that took part in a seq_cp and still hold tokens

    def run_update():
        w = kv.update()
        assert w is not None
        for a, b in w["cow_pairs"]:
            content[b * bs:(b + 1) * bs] = content[a * bs:(a + 1) * bs]
        if w["row_moves"]:
            content[[m[1] for m in w["row_moves"]]] = content[[m[0] for m in w["row_moves"]]].copy()
        if w["defrag_moves"]:
            snap = content.copy()
            for a, b in w["defrag_moves"]:
                content[b * bs:(b + 1) * bs] = snap[a * bs:(a + 1) * bs]

    for _ in range(120):
        run_update()                                          # kv_self_update before the decode
        lens = {s: len(kv.slots_of(s) or []) for s in range(n_seq)}
        live = [s for s in range(n_seq) if lens[s] > 0]
        op = rng.choice(["decode", "decode", "decode", "cp", "shift", "shift", "trunc", "keep", "defrag"])
        if op == "decode":
            n = int(rng.integers(1, 30))
            seq = [int(s) for s in rng.integers(0, n_seq, n)]
            ans = r.cmd("B %d %s" % (n, " ".join(map(str, seq))))
            if ans == "noslot":
                continue
            got = kv.find_slot(seq)
            assert isinstance(got, dict)
            for a, b in got["cow_pairs"]:
                content[b * bs:(b + 1) * bs] = content[a * bs:(a + 1) * bs]
            content[got["slot_mapping"]] = np.arange(next_tok, next_tok + n); next_tok += n
        elif op == "cp" and live:
            src, dst = int(rng.choice(live)), int(rng.integers(0, n_seq))
            if dst == src:
                continue
            p1 = int(rng.integers(1, lens[src] + 1))
            r.cmd(f"R {dst} -1 -1"); r.cmd(f"C {src} {dst} 0 {p1}")
            kv.seq_cp(src, dst, 0, p1)
            shared |= {src, dst}
        elif op == "shift" and [s for s in live if s not in shared and lens[s] > 3]:
            s = int(rng.choice([s for s in live if s not in shared and lens[s] > 3]))
            p0 = int(rng.integers(0, lens[s] - 2)); p1 = int(rng.integers(p0 + 1, lens[s]))
            assert r.cmd(f"R {s} {p0} {p1}") == "1" and kv.seq_rm(s, p0, p1)
            r.cmd(f"T {s} {p1} -1 {p0 - p1}"); kv.seq_add(s, p1, -1, p0 - p1)
            assert not kv.unsupported()
            shifts += 1
            run_update()
        elif op == "trunc" and live:
            s = int(rng.choice(live)); p0 = int(rng.integers(0, lens[s] + 1))
            assert r.cmd(f"R {s} {p0} -1") == "1" and kv.seq_rm(s, p0, -1)
            if p0 == 0:
                shared.discard(s)
        elif op == "keep" and live and rng.random() < 0.25:
            s = int(rng.choice(live))
            r.cmd(f"K {s}"); kv.seq_keep(s)
            shared &= {s}
        elif op == "defrag":
            kv.defrag()                                       # the reference's defrag needs its graph: visibility must not change
            run_update()
        n_tokens = 0
        for s in range(n_seq):
            vis = r.visible(s)
            cells = kv.slots_of(s) or []
            assert [p for p, _ in vis] == list(range(len(vis))) and len(cells) == len(vis)
            assert content[cells].tolist() == [t for _, t in vis]
            assert kv.seq_pos_max(s) == int(r.cmd(f"P {s}"))
            n_tokens += len(cells)
        assert kv.get_n_tokens() == n_tokens == int(r.cmd("N").split()[0])
    assert shifts >= 5                                        # the pair of calls really ran
    r.close()
