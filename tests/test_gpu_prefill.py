"""GPU parity of the chunked-prefill tile kernel (tcgen05 / TMEM, csrc/attn_prefill.cu) against the CPU oracle:
many causal query rows per sequence appended into the paged cache (BASELINE config 3).  Same bar as decode:
max relative error <= 2e-3 for F16 KV against the F32 CPU result, NMSE <= 5e-4 (the reference's own bar)."""
import numpy as np
import pytest

from test_gpu_parity import TOL, _run_paged
from workloads import F16, Q4_0, Q8_0, make_workload, oracle_attention

pytestmark = pytest.mark.gpu
PREFILL = 3


def _chunk(seq, p0, p1):
    return [(seq, p) for p in range(p0, p1)]


CASES = {
    # name: (seq_lens, n_head, n_head_kv, rows, block_size)
    "one_tile_gqa4":      ([300], 8, 2, _chunk(0, 268, 300), 16),
    "chunk128_gqa4":      ([300], 8, 2, _chunk(0, 172, 300), 16),
    "first_chunk":        ([200], 8, 2, _chunk(0, 0, 200), 16),                       # positions 0.. : rows that see 1 cell
    "two_seqs_ragged":    ([90, 41], 8, 2, _chunk(0, 58, 90) + _chunk(1, 30, 41) + [(0, 0), (1, 0)], 16),
    "interleaved_rows":   ([77, 130], 8, 2, [(i % 2, 40 + i // 2) for i in range(60)], 16),    # every row its own tile
    "gqa8_70b":           ([1000], 64, 8, _chunk(0, 744, 1000), 16),
    "mha_gqa1":           ([260], 4, 4, _chunk(0, 0, 260), 16),
    "block32":            ([513, 64], 8, 2, _chunk(0, 385, 513) + _chunk(1, 0, 64), 32),
    "chunk512_ctx2k":     ([2048, 700], 32, 8, _chunk(0, 1536, 2048) + _chunk(1, 188, 700), 16),   # config-3 shape, 2 sequences
    "not_causal_limit":   ([150], 8, 2, [(0, 149)] * 40, 16),                          # q_pos beyond: every row sees all 150
    # KV split of the tail items (fewer work items than SMs, or a last wave that is not full): pieces + pf_combine_kernel
    "one_seq_chunk256_ctx2k": ([2048], 8, 2, _chunk(0, 1792, 2048), 16),              # 8 items of 32 kv tiles -> 2 pieces each (a piece is >= 16 tiles)
    "one_seq_chunk128_ctx8k": ([8192], 8, 2, _chunk(0, 8064, 8192), 16),              # 4 items of 128 kv tiles -> 8 pieces each
    "one_seq_first_chunk512": ([512], 32, 8, _chunk(0, 0, 512), 16),                  # early rows see nothing of the later pieces
    # 80 one-row tiles x 2 kv heads = 160 items: 148 whole ones + the 12 of the six long sequences (first in row order = last
    # in work order): 3 pieces for the 3300-cell ones, 2 for the 2100-cell ones, the 900-cell one runs whole
    "tail_wave_160_items":    ([3300, 3300, 2100, 2100, 2100, 900] + [40 + 3*i for i in range(74)], 8, 2,
                               [(s, n - 1) for s, n in enumerate([3300, 3300, 2100, 2100, 2100, 900])] + [(6 + i, 39 + 3*i) for i in range(74)], 16),
}


@pytest.mark.parametrize("name", list(CASES))
def test_prefill_tile_kernel(pkg, oracle, cuda, name):
    import torch
    seq_lens, n_head, n_head_kv, rows, bs = CASES[name]
    w = make_workload(oracle, seq_lens, n_head, n_head_kv, 128, F16, seed=31, q_rows=rows, block_size=bs)
    n0 = pkg.lib.b200_launch_count()
    got = _run_paged(pkg, torch, w, PREFILL)
    # plan + main, and the merge of split items only when some sequence is long enough for two pieces (2 x 16 kv tiles)
    assert pkg.lib.b200_launch_count() - n0 == (3 if max(seq_lens) >= 2048 else 2)
    want = oracle_attention(oracle, w)
    assert not np.isnan(got).any()
    err, nm = oracle.max_rel_err(want, got), oracle.nmse(want, got)
    assert err <= TOL[F16], f"max rel err {err:.3e} (nmse {nm:.3e})"
    assert nm <= 5e-4


@pytest.mark.parametrize("tk,tv", [(Q8_0, Q8_0), (Q4_0, Q4_0), (Q8_0, F16), (F16, Q4_0)], ids=["q8", "q4", "q8k_f16v", "f16k_q4v"])
@pytest.mark.parametrize("name", ["two_seqs_ragged", "chunk512_ctx2k", "block32", "gqa8_70b", "one_seq_chunk256_ctx2k"])
def test_prefill_quantised_cache(pkg, oracle, cuda, name, tk, tv):
    """quantised caches: sequences expanded once to the F16 scratch, then the same tile kernel"""
    import torch
    seq_lens, n_head, n_head_kv, rows, bs = CASES[name]
    w = make_workload(oracle, seq_lens, n_head, n_head_kv, 128, tk, tv, seed=32, q_rows=rows, block_size=bs)
    got = _run_paged(pkg, torch, w, PREFILL)
    want = oracle_attention(oracle, w)
    assert not np.isnan(got).any()
    err, nm = oracle.max_rel_err(want, got), oracle.nmse(want, got)
    assert err <= max(TOL[tk], TOL[tv]), f"max rel err {err:.3e} (nmse {nm:.3e})"
    assert nm <= 5e-4


def test_prefill_matches_streamk_rows(pkg, oracle, cuda):
    """the two kernels implement the same op: same inputs, results within twice the oracle bar of each other"""
    import torch
    rows = _chunk(0, 100, 612) + _chunk(1, 0, 300)
    w = make_workload(oracle, [612, 300], 32, 8, 128, F16, seed=5, q_rows=rows, dist="normal")
    a = _run_paged(pkg, torch, w, PREFILL)
    b = _run_paged(pkg, torch, w, 2)
    assert oracle.max_rel_err(b, a) <= 2*TOL[F16]


def test_auto_picks_prefill_for_chunks(pkg, oracle, cuda):
    """AUTO routes a 512-row causal chunk to the tile kernel (2 launches: plan + main; the context is too short for
    KV-split pieces, so no merge launch) and still matches the oracle"""
    import torch
    rows = _chunk(0, 512, 1024)
    w = make_workload(oracle, [1024], 32, 8, 128, F16, seed=6, q_rows=rows)
    n0 = pkg.lib.b200_launch_count()
    got = _run_paged(pkg, torch, w, 0)
    assert pkg.lib.b200_launch_count() - n0 == 2
    assert oracle.max_rel_err(oracle_attention(oracle, w), got) <= TOL[F16]
