"""K-shift and context shift on the paged layout.

Reference: llama_kv_self_seq_rm + llama_kv_self_seq_add mark cells with a position delta (src/llama-kv-cache.cpp:173-242,
:327-383); the next kv_self_update runs build_kv_self_shift (src/llama-context.cpp:524-571): rope over the K cache with
the deltas as positions, through a dequantise -> rope -> requantise round trip for Q8_0 / Q4_0 caches.

  * oracle.k_shift restates that graph; it is pinned byte for byte to the reference's own graph (oracle/_ref) and to
    the committed vectors the reference produced (tests/golden/k_shift_vectors.npz);
  * PagedKVCache.context_shift (host, integers) must give every sequence exactly the token order the cell-per-token
    model of the reference gives;
  * b200_k_shift (CUDA) is compared with the oracle: cos/sin come from CUDA's libm instead of glibc's (<= 2 ulp apart),
    so an F16 value may land on the neighbouring half and a quant code on the neighbouring level in rare cases — the
    test bounds both the size (one step) and the rate (1e-3) of such differences, and the NMSE after dequantisation
    (1e-7, the reference's own bar for rope: tests/test-backend-ops.cpp test_rope uses the default max_nmse_err)."""
import importlib.util
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
F16, Q4_0, Q8_0 = 1, 2, 8


def _mg():
    spec = importlib.util.spec_from_file_location("make_golden_kshift", os.path.join(HERE, "golden", "make_golden_kshift.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


MG = _mg()
CASES = [(t, tn, name, kw) for t, tn in MG.TYPES for name, kw in MG.ROPE_CASES]
IDS = [f"{tn}-{name}" for _, tn, name, _ in CASES]


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "k_shift_vectors.npz"))


def _budget(t, ref_bytes, got_bytes, oracle, n_head_kv, d):
    """differences the libm gap may cause: rare, one step in size, invisible after dequantisation"""
    n_cells = ref_bytes.size // (oracle_row(t, d) * n_head_kv)
    a = oracle.to_float(t, ref_bytes.reshape(n_cells, n_head_kv, -1), d)
    b = oracle.to_float(t, got_bytes.reshape(n_cells, n_head_kv, -1), d)
    assert oracle.nmse(a, b) <= 1e-7
    if t == F16:
        ha, hb = ref_bytes.view(np.int16).astype(np.int32), got_bytes.view(np.int16).astype(np.int32)
        diff = np.abs(ha - hb)
        assert diff.max(initial=0) <= 1, "an F16 value is more than one half-precision step away"
        assert (diff != 0).mean() <= 1e-3
    else:
        assert (ref_bytes != got_bytes).mean() <= 1e-3


def oracle_row(t, n):
    return {F16: 2 * n, Q8_0: n // 32 * 34, Q4_0: n // 32 * 18}[t]


# ---------------------------------------------------------------- the oracle is the reference's graph

@pytest.mark.parametrize("t,tn,name,kw", CASES, ids=IDS)
def test_oracle_k_shift_matches_committed_reference_bytes(oracle, gold, t, tn, name, kw):
    x, shift = MG.inputs()
    c = oracle.from_float(t, x).reshape(-1).copy()
    oracle.k_shift(c, t, MG.N_HEAD_KV, MG.D, shift, **kw)
    want = gold[f"{tn}_{name}"]
    if t == Q8_0:
        # the reference's ggml_cpy back into a Q8_0 cache runs the CPU backend's SIMD quantiser, which differs from the
        # deterministic _ref quantiser on ties (tests/test_oracle_vs_ref.py::test_cpu_backend_cpy_vs_ref_quantiser)
        assert (c != want).mean() <= 1e-4
        fa = oracle.to_float(t, want.reshape(MG.N_CELLS, MG.N_HEAD_KV, -1), MG.D)
        fb = oracle.to_float(t, c.reshape(MG.N_CELLS, MG.N_HEAD_KV, -1), MG.D)
        assert oracle.nmse(fa, fb) <= 1e-7
    else:
        assert np.array_equal(c, want)


@pytest.mark.parametrize("t,tn,name,kw", CASES, ids=IDS)
def test_oracle_k_shift_matches_reference_graph(oracle, ref, t, tn, name, kw):
    rng = np.random.default_rng(11)
    n_cells, hkv, d = 40, 4, 128
    x = rng.uniform(-1, 1, (n_cells, hkv, d)).astype(np.float32)
    shift = rng.integers(-4096, 4096, n_cells).astype(np.int32)
    base = oracle.from_float(t, x).reshape(-1)
    a, b = base.copy(), base.copy()
    oracle.k_shift(a, t, hkv, d, shift, **kw)
    ref.k_shift(b, t, hkv, d, shift, **kw)
    if t == Q8_0:
        assert (a != b).mean() <= 1e-4
    else:
        assert np.array_equal(a, b)


def test_oracle_k_shift_zero_is_identity_and_cells_select(oracle):
    rng = np.random.default_rng(5)
    n_cells, hkv, d = 24, 2, 128
    x = rng.standard_normal((n_cells, hkv, d)).astype(np.float32)
    for t in (F16, Q8_0, Q4_0):
        base = oracle.from_float(t, x).reshape(-1)
        c = base.copy()
        oracle.k_shift(c, t, hkv, d, np.zeros(n_cells, np.int32))
        assert np.array_equal(c, base)                     # cos = 1, sin = 0; re-quantisation of dequantised blocks is idempotent
        shift = np.full(n_cells, -7, np.int32)
        full, part = base.copy(), base.copy()
        oracle.k_shift(full, t, hkv, d, shift)
        cells = np.array([3, 17, 4, -1, 99], np.int32)     # out-of-range cells are skipped
        oracle.k_shift(part, t, hkv, d, shift, cells)
        rb = oracle_row(t, d) * hkv
        full, part, base2 = full.reshape(n_cells, rb), part.reshape(n_cells, rb), base.reshape(n_cells, rb)
        for cidx in range(n_cells):
            assert np.array_equal(part[cidx], full[cidx] if cidx in (3, 17, 4) else base2[cidx])


def test_rotation_composes(oracle):
    """rope(rope(k, a), b) == rope(k, a + b) up to rounding: what makes a delta-only K-shift valid at all"""
    rng = np.random.default_rng(8)
    n_cells, hkv, d = 16, 2, 128
    x = rng.standard_normal((n_cells, hkv, d)).astype(np.float32)
    F32 = 0
    c1 = x.reshape(-1).view(np.uint8).copy()
    c2 = c1.copy()
    oracle.k_shift(c1, F32, hkv, d, np.full(n_cells, 100, np.int32))
    oracle.k_shift(c1, F32, hkv, d, np.full(n_cells, -164, np.int32))
    oracle.k_shift(c2, F32, hkv, d, np.full(n_cells, -64, np.int32))
    assert oracle.nmse(c2.view(np.float32), c1.view(np.float32)) <= 1e-10


# ---------------------------------------------------------------- host: context shift on the block table

class CellModel:
    """cell-per-token model of llama_kv_cache_unified restricted to what a context shift touches:
    seq_rm(seq, p0, p1) erases the id from cells in range (src/llama-kv-cache.cpp:199-240), seq_add(seq, p1, inf, -(p1-p0))
    moves the later positions down (:327-383; a cell shared with another sequence moves for both there — the paged
    cache copies the block instead, so the OTHER sequence must keep its order: that is the stronger property tested)"""

    def __init__(self):
        self.seqs = {}

    def append(self, seq, toks):
        self.seqs.setdefault(seq, []).extend(toks)

    def seq_cp(self, src, dst, p1):
        self.seqs[dst] = list(self.seqs[src][:p1])

    def context_shift(self, seq, p0, p1):
        s = self.seqs[seq]
        self.seqs[seq] = s[:p0] + s[p1:]
        if not self.seqs[seq]:
            del self.seqs[seq]


def _apply_device_plan(store, bs, cow, moves):
    """what the device does with the plan, on a [n_cells] array of token ids"""
    for s, d in cow:
        store[d * bs:(d + 1) * bs] = store[s * bs:(s + 1) * bs]
    if moves:
        src = np.array([m[0] for m in moves]); dst = np.array([m[1] for m in moves])
        staged = store[src].copy()
        store[dst] = staged


@pytest.mark.parametrize("seed", range(12))
def test_context_shift_matches_the_cell_model(pkg, seed):
    rng = np.random.default_rng(seed)
    bs = int(rng.choice([1, 4, 16]))
    n_blocks = 64
    pc = pkg.PagedKVCache(n_blocks, bs)
    model = CellModel()
    store = np.full(n_blocks * bs, -1, dtype=np.int64)
    next_tok = [0]
    shifted_total = 0

    def append(seq, n):
        r = pc.append(seq, n)
        if r is None:
            return
        slots, cow = r
        _apply_device_plan(store, bs, cow, [])
        toks = list(range(next_tok[0], next_tok[0] + n)); next_tok[0] += n
        store[slots] = toks
        model.append(seq, toks)

    for _ in range(60):
        op = rng.integers(0, 4)
        live = sorted(model.seqs)
        if op == 0 or not live:
            append(int(rng.integers(0, 5)), int(rng.integers(1, 40)))
        elif op == 1:
            src = int(rng.choice(live)); dst = int(rng.integers(0, 5))
            if dst != src:
                p1 = int(rng.integers(1, len(model.seqs[src]) + 1))
                assert pc.seq_cp(src, dst, 0, p1)
                model.seq_cp(src, dst, p1)
        else:
            seq = int(rng.choice(live)); n = len(model.seqs[seq])
            p0 = int(rng.integers(0, n)); p1 = int(rng.integers(p0 + 1, n + 1))
            if rng.random() < 0.5:                      # the aligned fast path
                p0 = p0 // bs * bs; p1 = min(n // bs * bs, p0 + max(bs, (p1 - p0) // bs * bs))
                if p1 <= p0:
                    continue
            free_before = pc.n_free()
            plan = pc.context_shift(seq, p0, p1)
            if plan is None:
                assert free_before < n_blocks             # only ever refused for lack of copy-on-write room
                continue
            cow, moves, cells = plan
            _apply_device_plan(store, bs, cow, moves)
            model.context_shift(seq, p0, p1)
            new_len = n - (p1 - p0)
            assert len(cells) == new_len - p0
            if (p0 % bs == 0) and ((p1 - p0) % bs == 0):
                assert moves == []
            if cells:
                assert cells == pc.slots_of(seq)[p0:]
                for c in cells:
                    assert pc.ref_count(c // bs) == 1     # nothing shared gets rotated
            shifted_total += len(cells)
        # every sequence reads back its tokens in order; ref-counts add up
        refs = np.zeros(n_blocks, dtype=np.int64)
        for s, toks in model.seqs.items():
            assert pc.seq_len(s) == len(toks)
            slots = pc.slots_of(s)
            assert list(store[slots]) == toks, (seed, s)
            for b in sorted({c // bs for c in slots}):
                refs[b] += 1
        assert [pc.ref_count(b) for b in range(n_blocks)] == list(refs)
        assert pc.n_free() == int((refs == 0).sum())
    assert shifted_total > 0


def test_context_shift_refusals_and_all_or_nothing(pkg):
    pc = pkg.PagedKVCache(3, 4)
    assert pc.append(0, 8) is not None                 # blocks 0, 1
    assert pc.seq_cp(0, 1, 0, 8)
    assert pc.context_shift(0, 3, 3) is None and pc.context_shift(0, -1, 2) is None and pc.context_shift(0, 2, 9) is None
    assert pc.context_shift(5, 0, 1) is None
    # dropping [1, 3) rewrites blocks 0 and 1, both shared: two copies needed, one block free -> refused, nothing changed
    assert pc.context_shift(0, 1, 3) is None
    assert pc.seq_len(0) == 8 and pc.slots_of(0) == list(range(8)) and pc.n_free() == 1
    assert [pc.ref_count(b) for b in range(3)] == [2, 2, 0]
    # dropping the first block needs one copy (the survivor is shared and gets rotated)
    cow, moves, cells = pc.context_shift(0, 0, 4)
    assert cow == [(1, 2)] and moves == [] and cells == [8, 9, 10, 11]
    assert pc.slots_of(1) == list(range(8)) and [pc.ref_count(b) for b in range(3)] == [1, 1, 1]
    # dropping the tail is a plain truncation: nothing to move or rotate
    assert pc.context_shift(1, 5, 8) == ([], [], []) and pc.seq_len(1) == 5
    # dropping everything removes the sequence
    assert pc.context_shift(0, 0, 4) == ([], [], []) and pc.seq_len(0) == 0 and pc.slots_of(0) is None


# ---------------------------------------------------------------- CUDA

def _dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.gpu
@pytest.mark.parametrize("t,tn,name,kw", CASES, ids=IDS)
def test_gpu_k_shift_matches_oracle_and_reference_vectors(pkg, oracle, cuda, gold, t, tn, name, kw):
    import torch
    x, shift = MG.inputs()
    base = oracle.from_float(t, x).reshape(-1)
    want = base.copy()
    oracle.k_shift(want, t, MG.N_HEAD_KV, MG.D, shift, **kw)
    lay = pkg.KVLayout(t, MG.N_HEAD_KV, MG.D, 16, MG.N_CELLS // 16)
    layers = [_dev(torch, base), _dev(torch, base)]                      # two layers, one launch
    kw_dev = dict(kw)
    if "freq_factors" in kw_dev:
        kw_dev["freq_factors"] = _dev(torch, kw_dev["freq_factors"])
    n0 = pkg.lib.b200_launch_count()
    pkg.k_shift(layers, lay, _dev(torch, shift), **kw_dev)
    assert pkg.lib.b200_launch_count() - n0 == 1
    torch.cuda.synchronize()
    for c in layers:
        got = c.cpu().numpy()
        _budget(t, want, got, oracle, MG.N_HEAD_KV, MG.D)
        _budget(t, gold[f"{tn}_{name}"], got, oracle, MG.N_HEAD_KV, MG.D)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [0, 2], ids=["norm", "neox"])
@pytest.mark.parametrize("t", [F16, Q8_0, Q4_0])
def test_gpu_k_shift_register_form_equals_general_form(pkg, oracle, cuda, t, mode):
    """head_dim 128 runs on the register form (a lane = 4 elements, NEOX partners by shuffle, no F32 row in shared memory);
    B200_KSHIFT_GENERAL=1 forces the general kernel: same arithmetic sequence, so the cache bytes must be identical"""
    import os
    import torch
    rng = np.random.default_rng(50 + t)
    n_cells, hkv, d, L = 96, 8, 128, 3
    lay = pkg.KVLayout(t, hkv, d, 16, n_cells // 16)
    base = [oracle.from_float(t, rng.uniform(-1, 1, (n_cells, hkv, d)).astype(np.float32)).reshape(-1) for _ in range(L)]
    shift = rng.integers(-4000, 1, n_cells).astype(np.int32)
    outs = []
    for general in ("", "1"):
        if general:
            os.environ["B200_KSHIFT_GENERAL"] = general
        else:
            os.environ.pop("B200_KSHIFT_GENERAL", None)
        try:
            layers = [_dev(torch, b) for b in base]
            pkg.k_shift(layers, lay, _dev(torch, shift), mode=mode, freq_base=500000.0)
            torch.cuda.synchronize()
            outs.append([c.cpu().numpy() for c in layers])
        finally:
            os.environ.pop("B200_KSHIFT_GENERAL", None)
    for a, b, b0 in zip(outs[0], outs[1], base):
        assert np.array_equal(a, b), "register form and general form disagree"
        assert not np.array_equal(a, b0)


@pytest.mark.gpu
@pytest.mark.parametrize("t", [F16, Q8_0, Q4_0])
def test_gpu_k_shift_cell_list_and_strided_heads(pkg, oracle, cuda, t):
    """only the listed cells change; zero shift leaves the bytes alone; Llama-3-8B row shape (8 kv heads)"""
    import torch
    rng = np.random.default_rng(21)
    n_cells, hkv, d = 64, 8, 128
    x = rng.uniform(-1, 1, (n_cells, hkv, d)).astype(np.float32)
    base = oracle.from_float(t, x).reshape(-1)
    shift = np.zeros(n_cells, np.int32); shift[10:30] = -16
    cells = np.array(list(range(8, 32)) + [-3, 1000], np.int32)
    want = base.copy()
    oracle.k_shift(want, t, hkv, d, shift, cells)
    lay = pkg.KVLayout(t, hkv, d, 16, n_cells // 16)
    c = _dev(torch, base)
    pkg.k_shift([c], lay, _dev(torch, shift), _dev(torch, cells))
    torch.cuda.synchronize()
    got = c.cpu().numpy()
    rb = lay.token_bytes
    g2, b2 = got.reshape(n_cells, rb), base.reshape(n_cells, rb)
    untouched = [i for i in range(n_cells) if not (10 <= i < 30)]
    assert np.array_equal(g2[untouched], b2[untouched])
    _budget(t, want, got, oracle, hkv, d)


@pytest.mark.gpu
@pytest.mark.parametrize("t", [F16, Q4_0])
@pytest.mark.parametrize("p0,p1", [(16, 48), (5, 23)], ids=["aligned", "ragged"])
def test_gpu_context_shift_end_to_end(pkg, oracle, cuda, t, p0, p1):
    """two sequences sharing a prefix; sequence 0 drops [p0, p1): block copies, row moves and K rotation on the device
    over 2 layers == the same plan applied on the CPU with the oracle; V rows are moved byte for byte; sequence 1 intact"""
    import torch
    rng = np.random.default_rng(33)
    bs, n_blocks, hkv, d, L = 16, 24, 8, 128, 2
    lay = pkg.KVLayout(t, hkv, d, bs, n_blocks)
    pc = pkg.PagedKVCache(n_blocks, bs)
    slots0, _ = pc.append(0, 100)
    assert pc.seq_cp(0, 1, 0, 40)
    slots1_before = pc.slots_of(1)
    k_host = [oracle.from_float(t, rng.uniform(-1, 1, (lay.n_cells, hkv, d)).astype(np.float32)).reshape(-1) for _ in range(L)]
    v_host = [oracle.from_float(t, rng.uniform(-1, 1, (lay.n_cells, hkv, d)).astype(np.float32)).reshape(-1) for _ in range(L)]
    kd, vd = [_dev(torch, a) for a in k_host], [_dev(torch, a) for a in v_host]
    # the expected result: same plan, CPU
    probe = pkg.PagedKVCache(n_blocks, bs)
    probe.append(0, 100); probe.seq_cp(0, 1, 0, 40)
    cow, moves, cells = probe.context_shift(0, p0, p1)
    rb = lay.token_bytes
    want_k, want_v = [a.copy() for a in k_host], [a.copy() for a in v_host]
    for a in want_k + want_v:
        a2 = a.reshape(lay.n_cells, rb)
        for s, dd in cow:
            a2[dd * bs:(dd + 1) * bs] = a2[s * bs:(s + 1) * bs]
        if moves:
            src = np.array([m[0] for m in moves]); dst = np.array([m[1] for m in moves])
            a2[dst] = a2[src].copy()
    shift = np.zeros(lay.n_cells, np.int32); shift[cells] = p0 - p1
    for a in want_k:
        oracle.k_shift(a, t, hkv, d, shift, np.array(cells, np.int32))
    assert pkg.context_shift(pc, 0, p0, p1, kd, vd, lay)
    torch.cuda.synchronize()
    assert pc.slots_of(0) == probe.slots_of(0) and pc.seq_len(0) == 100 - (p1 - p0)
    assert pc.slots_of(1) == slots1_before
    for l in range(L):
        assert np.array_equal(vd[l].cpu().numpy(), want_v[l])
        gk = kd[l].cpu().numpy()
        _budget(t, want_k[l], gk, oracle, hkv, d)
        keep = np.array(slots1_before)                  # the other sequence's rows: bytes untouched
        assert np.array_equal(gk.reshape(lay.n_cells, rb)[keep], k_host[l].reshape(lay.n_cells, rb)[keep])
