"""KV session state and defrag on the GPU, through the C ABI.

save:  b200_kv_rows_gather pulls a sequence's rows out of its scattered blocks -> the reference's stream format; the
       stream must equal the reference writer's output for the same tokens in a flat cache, byte for byte (the rows
       are the oracle's own kv_write bytes, so the whole chain write -> gather -> frame is pinned);
load:  parse -> seq_rm + reserve -> b200_kv_rows_scatter into DIFFERENT blocks; attention over the restored sequence
       is bit-identical to before (same bytes, same kernel), like examples/save-load-state checks generation;
defrag: PagedKVCache::defrag plan + b200_block_copy; every sequence's attention output stays bit-identical.
"""
import numpy as np
import pytest

from oracle.kv_state_oracle import unified_state_write
from test_gpu_parity import _dev
from workloads import F16, Q4_0, Q8_0

pytestmark = pytest.mark.gpu
N_LAYER = 3


def _setup(pkg, oracle, torch, t, seq_lens, n_blocks=96, bs=16, n_head_kv=2, d=128, seed=5):
    """N_LAYER caches filled through PagedKVCache + b200_kv_write, interleaved appends so blocks are scattered"""
    lay = pkg.KVLayout(t, n_head_kv, d, bs, n_blocks)
    pc = pkg.PagedKVCache(n_blocks, bs)
    rng = np.random.default_rng(seed)
    kcs, vcs = [lay.alloc() for _ in range(N_LAYER)], [lay.alloc() for _ in range(N_LAYER)]
    hk = [np.zeros((lay.n_cells, lay.token_bytes), dtype=np.uint8) for _ in range(N_LAYER)]
    hv = [np.zeros((lay.n_cells, lay.token_bytes), dtype=np.uint8) for _ in range(N_LAYER)]
    left = dict(enumerate(seq_lens))
    while any(left.values()):
        for s in list(left):
            n = min(left[s], int(rng.integers(1, 24)))
            if n == 0:
                continue
            left[s] -= n
            slots, cow = pc.append(s, n)
            assert cow == []
            sm = np.asarray(slots, dtype=np.int32)
            for il in range(N_LAYER):
                ks = rng.uniform(-1, 1, (n, n_head_kv*d)).astype(np.float32)
                vs = rng.uniform(-1, 1, (n, n_head_kv*d)).astype(np.float32)
                pkg.kv_write(_dev(torch, ks), _dev(torch, vs), _dev(torch, sm), kcs[il], vcs[il], lay)
                oracle.kv_write(ks, sm, hk[il], t, n_head_kv, d, lay.token_bytes, lay.head_bytes, lay.n_cells)
                oracle.kv_write(vs, sm, hv[il], t, n_head_kv, d, lay.token_bytes, lay.head_bytes, lay.n_cells)
    return lay, pc, kcs, vcs, hk, hv, rng


def _attend(pkg, torch, pc, lay, kc, vc, seqs, q, stride=16):
    bt, lens = pc.export(seqs, stride)
    out = pkg.paged_attention(_dev(torch, q), kc, vc, lay, _dev(torch, np.asarray(bt, dtype=np.int32)),
                              _dev(torch, np.asarray(lens, dtype=np.int32)), 0.0884)
    torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.mark.parametrize("t", [F16, Q8_0, Q4_0])
def test_state_seq_save_load_round_trip(pkg, oracle, cuda, t):
    import torch
    lay, pc, kcs, vcs, hk, hv, rng = _setup(pkg, oracle, torch, t, [70, 133, 41])
    n_head = 8
    q = rng.uniform(-1, 1, (1, n_head, lay.head_dim)).astype(np.float32)
    before = [_attend(pkg, torch, pc, lay, kcs[il], vcs[il], [1], q) for il in range(N_LAYER)]

    n0 = pkg.lib.b200_launch_count()
    blob = pkg.state_seq_save(pc, 1, kcs, vcs, lay)
    assert pkg.lib.b200_launch_count() - n0 == 2                       # all layers' K in one launch, V in another
    # the reference writer over a flat unified cache that holds the same token rows (append-only: cell i = token i)
    cells_of = pc.slots_of(1)
    flat_k = [h[cells_of] for h in hk]; flat_v = [h[cells_of] for h in hv]
    want = unified_state_write([(p, {1}) for p in range(len(cells_of))], 1, flat_k, flat_v, t, t)
    assert blob == want

    # disturb the allocator, then restore into different blocks
    old_blocks = set(c // lay.block_size for c in cells_of)
    assert pc.seq_rm(1)
    pc.append(7, 37)                                                   # takes some of the freed blocks
    for il in range(N_LAYER):                                          # and scribble over everything seq 1 used to own
        kcs[il].view(-1, lay.token_bytes)[cells_of] = 0xAB
    n = pkg.state_seq_load(pc, 1, blob, kcs, vcs, lay)
    assert n == 133 and pc.seq_len(1) == 133
    new_blocks = set(c // lay.block_size for c in pc.slots_of(1))
    assert new_blocks != old_blocks
    after = [_attend(pkg, torch, pc, lay, kcs[il], vcs[il], [1], q) for il in range(N_LAYER)]
    for a, b in zip(before, after):
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), "restored sequence attends differently"
    # a second save reproduces the stream
    assert pkg.state_seq_save(pc, 1, kcs, vcs, lay) == blob
    # streams of another cache shape are refused before anything is touched
    lay2 = pkg.KVLayout(F16 if t != F16 else Q8_0, lay.n_head_kv, lay.head_dim, lay.block_size, lay.n_blocks)
    with pytest.raises(pkg.B200Error, match="failed to restore kv cache"):
        pkg.state_seq_load(pc, 1, blob, kcs, vcs, lay2)
    assert pc.seq_len(1) == 133


def test_state_of_empty_and_unknown_sequences(pkg, oracle, cuda):
    import torch
    lay, pc, kcs, vcs, *_ = _setup(pkg, oracle, torch, F16, [20])
    blob = pkg.state_seq_save(pc, 42, kcs, vcs, lay)                   # unknown sequence: cell_count 0, like the reference
    assert len(blob) == pkg.lib.b200_kv_state_seq_size(0, N_LAYER, lay.token_bytes, lay.token_bytes)
    assert pkg.state_seq_load(pc, 3, blob, kcs, vcs, lay) == 0 and pc.seq_len(3) == 0


@pytest.mark.parametrize("t", [F16, Q8_0])
def test_defrag_moves_blocks_without_changing_results(pkg, oracle, cuda, t):
    import torch
    lay, pc, kcs, vcs, hk, hv, rng = _setup(pkg, oracle, torch, t, [90, 60, 120, 33, 75], n_blocks=64)
    assert pc.seq_cp(2, 9, 0, 100)                                     # a fork: shared blocks must move once, for both
    pc.seq_rm(1); pc.seq_rm(3)                                         # holes
    seqs = [0, 2, 4, 9]
    q = rng.uniform(-1, 1, (len(seqs), 8, lay.head_dim)).astype(np.float32)
    before = [_attend(pkg, torch, pc, lay, kcs[il], vcs[il], seqs, q) for il in range(N_LAYER)]
    n_used = lay.n_blocks - pc.n_free()
    moves = pc.defrag()
    assert moves and all(s >= n_used > d for s, d in moves)
    n0 = pkg.lib.b200_launch_count()
    pkg.block_copy(kcs + vcs, lay.block_bytes, lay.n_blocks, _dev(torch, np.asarray(moves, dtype=np.int32)))
    assert pkg.lib.b200_launch_count() - n0 == 1                       # every layer's K and V in one launch
    for s, d in moves:                                                 # poison the vacated blocks: nothing may still read them
        for c in kcs + vcs:
            c.view(lay.n_blocks, lay.block_bytes)[s] = 0xEE
    after = [_attend(pkg, torch, pc, lay, kcs[il], vcs[il], seqs, q) for il in range(N_LAYER)]
    for a, b in zip(before, after):
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    bt, _ = pc.export(seqs, 16)
    assert max(max(r) for r in bt) < n_used
