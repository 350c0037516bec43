"""GPU parity tests proper: the CUDA path, called through the C ABI, against the CPU oracle on the same
seeded inputs.  Bars (north_star): KV write / block copy / block tables bit-exact; attention max relative
error <= 2e-3 for F16 KV and <= 1e-2 for Q8_0/Q4_0 against the F32 CPU result, where the relative error
of an element is |out-ref| / max(|ref|, rms(ref row)) (oracle.max_rel_err); NMSE <= 5e-4 is the
reference's own bar (tests/test-backend-ops.cpp:3216-3218)."""
import numpy as np
import pytest

from workloads import F16, F32, Q4_0, Q8_0, make_workload, mixed_ctx, oracle_attention, row_bytes

pytestmark = pytest.mark.gpu

TOL = {F32: 2e-3, F16: 2e-3, Q8_0: 1e-2, Q4_0: 1e-2}
TYPES = [F16, Q8_0, Q4_0]


def _dev(torch, a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return t if dtype is None else t.to(dtype)


def _run_paged(pkg, torch, w, kernel=0, hints=True):
    lay_k = pkg.KVLayout(w.type_k, w.n_head_kv, w.d, w.block_size, w.n_blocks)
    lay_v = pkg.KVLayout(w.type_v, w.n_head_kv, w.d, w.block_size, w.n_blocks)
    q = _dev(torch, w.q)
    kc, vc = _dev(torch, w.k_cache.reshape(-1)), _dev(torch, w.v_cache.reshape(-1))
    bt, sl = _dev(torch, w.block_table), _dev(torch, w.seq_lens)
    qs = None if w.q_seq is None else _dev(torch, w.q_seq)
    qp = None if w.q_pos is None else _dev(torch, w.q_pos)
    out = torch.full_like(q, float("nan"))
    call = pkg.AttnCall(q, kc, vc, lay_k, out, w.scale, layout_v=lay_v, block_table=bt, seq_lens=sl, q_seq=qs, q_pos=qp,
                        max_seq_len=int(w.seq_lens.max()) if hints else 0,
                        total_tokens=int(w.seq_lens.sum()) if hints else 0, kernel=kernel)
    if not call.supported():
        pytest.skip("kernel does not take this shape")
    before = pkg.lib.b200_launch_count()
    call.run()
    torch.cuda.synchronize()
    assert pkg.lib.b200_launch_count() > before, "no CUDA kernel was launched"
    return out.cpu().numpy()


# ---------------------------------------------------------------- KV write / block copy: bit-exact

@pytest.mark.parametrize("t", [F32, F16, Q8_0, Q4_0])
@pytest.mark.parametrize("shape", [(8, 128), (2, 64), (1, 32)])
def test_kv_write_bit_exact(pkg, oracle, cuda, t, shape):
    import torch
    n_head_kv, d = shape
    rng = np.random.default_rng(100 + t)
    n_tokens, n_cells = 300, 512
    src_k = rng.uniform(-1, 1, (n_tokens, n_head_kv * d)).astype(np.float32)
    src_v = rng.standard_normal((n_tokens, n_head_kv * d)).astype(np.float32) * 3
    src_k[0, :32] = 0.0
    src_k[1, :32] = -0.75
    src_k[2, :32] = np.linspace(-1, 1, 32); src_k[2, 31] = -1.0
    src_k[3, 5] = 6.0e4
    slot = rng.permutation(n_cells)[:n_tokens].astype(np.int32)
    slot[7] = -1                      # skipped (writeToCache ignores invalid slots)
    slot[9] = n_cells + 3             # out of range: skipped
    lay = pkg.KVLayout(t, n_head_kv, d, 16, n_cells // 16)
    kc = torch.full((lay.total_bytes,), 0x5a, dtype=torch.uint8, device=cuda)
    vc = torch.full((lay.total_bytes,), 0x5a, dtype=torch.uint8, device=cuda)
    pkg.kv_write(_dev(torch, src_k), _dev(torch, src_v), _dev(torch, slot), kc, vc, lay)
    torch.cuda.synchronize()
    for src, dev in ((src_k, kc), (src_v, vc)):
        want = np.full((n_cells, lay.token_bytes), 0x5a, dtype=np.uint8)
        oracle.kv_write(src, slot, want, t, n_head_kv, d, lay.token_bytes, lay.head_bytes, n_cells)
        got = dev.cpu().numpy().reshape(n_cells, lay.token_bytes)
        assert np.array_equal(got, want), f"{int((got != want).sum())} bytes differ"


@pytest.mark.parametrize("t", [F32, F16, Q8_0, Q4_0])
@pytest.mark.parametrize("shape", [(8, 128), (4, 64)])
def test_kv_write_many_rows_bit_exact(pkg, oracle, cuda, t, shape):
    """a prefill-sized store (more one-row CTAs than fit the machine at once) takes the several-rows-per-CTA form of the
    kernel: a token count that is not a multiple of its row group, skipped slots inside a group"""
    import torch
    n_head_kv, d = shape
    rng = np.random.default_rng(300 + t)
    n_tokens, n_cells = 1501, 2048
    src_k = rng.uniform(-1, 1, (n_tokens, n_head_kv * d)).astype(np.float32)
    src_v = rng.standard_normal((n_tokens, n_head_kv * d)).astype(np.float32) * 3
    src_k[4, :32] = 0.0
    src_k[5, 7] = 6.0e4
    slot = rng.permutation(n_cells)[:n_tokens].astype(np.int32)
    slot[6] = -1
    slot[1499] = n_cells + 9
    slot[1500] = 3 if 3 not in slot[:1500] else slot[1500]
    lay = pkg.KVLayout(t, n_head_kv, d, 16, n_cells // 16)
    kc = torch.full((lay.total_bytes,), 0x5a, dtype=torch.uint8, device=cuda)
    vc = torch.full((lay.total_bytes,), 0x5a, dtype=torch.uint8, device=cuda)
    pkg.kv_write(_dev(torch, src_k), _dev(torch, src_v), _dev(torch, slot), kc, vc, lay)
    torch.cuda.synchronize()
    for src, dev in ((src_k, kc), (src_v, vc)):
        want = np.full((n_cells, lay.token_bytes), 0x5a, dtype=np.uint8)
        oracle.kv_write(src, slot, want, t, n_head_kv, d, lay.token_bytes, lay.head_bytes, n_cells)
        got = dev.cpu().numpy().reshape(n_cells, lay.token_bytes)
        assert np.array_equal(got, want), f"{int((got != want).sum())} bytes differ"


def test_kv_write_mixed_types_and_strided_heads(pkg, oracle, cuda):
    """type_k != type_v (-ctk q8_0 -ctv f16) and a head stride wider than the row (generic store path)"""
    import torch
    n_head_kv, d, n_tokens, n_cells = 4, 128, 40, 64
    rng = np.random.default_rng(5)
    src = rng.uniform(-1, 1, (n_tokens, n_head_kv * d)).astype(np.float32)
    slot = rng.permutation(n_cells)[:n_tokens].astype(np.int32)
    d_ = pkg.KVWriteDesc()
    nbh_k, nbh_v = row_bytes(Q8_0, d) + 6, row_bytes(F16, d)
    nbt_k, nbt_v = nbh_k * n_head_kv + 10, nbh_v * n_head_kv
    kc = torch.zeros(n_cells * nbt_k, dtype=torch.uint8, device=cuda)
    vc = torch.zeros(n_cells * nbt_v, dtype=torch.uint8, device=cuda)
    s, sm = _dev(torch, src), _dev(torch, slot)
    d_.k_src = d_.v_src = s.data_ptr(); d_.src_nb1 = src.strides[0]; d_.n_tokens = n_tokens
    d_.slot_mapping = sm.data_ptr(); d_.k_cache, d_.v_cache = kc.data_ptr(), vc.data_ptr()
    d_.type_k, d_.type_v, d_.n_head_kv, d_.head_dim = Q8_0, F16, n_head_kv, d
    d_.nbk_token, d_.nbk_head, d_.nbv_token, d_.nbv_head, d_.n_cells = nbt_k, nbh_k, nbt_v, nbh_v, n_cells
    import ctypes as C
    pkg.check(pkg.lib.b200_kv_write(C.byref(d_), C.c_void_p(pkg.current_stream())), "kv_write")
    torch.cuda.synchronize()
    wk = np.zeros(n_cells * nbt_k, dtype=np.uint8); wv = np.zeros(n_cells * nbt_v, dtype=np.uint8)
    oracle.kv_write(src, slot, wk, Q8_0, n_head_kv, d, nbt_k, nbh_k, n_cells)
    oracle.kv_write(src, slot, wv, F16, n_head_kv, d, nbt_v, nbh_v, n_cells)
    assert np.array_equal(kc.cpu().numpy(), wk) and np.array_equal(vc.cpu().numpy(), wv)


@pytest.mark.parametrize("t", TYPES)
def test_block_copy_bit_exact(pkg, oracle, cuda, t):
    import torch
    lay = pkg.KVLayout(t, 8, 128, 16, 40)
    rng = np.random.default_rng(t)
    caches = [rng.integers(0, 256, lay.total_bytes, dtype=np.uint8) for _ in range(5)]
    pairs = np.array([[3, 20], [4, 21], [39, 0], [7, 7], [11, 30]], dtype=np.int32)
    dev = [_dev(torch, c) for c in caches]
    pkg.block_copy(dev, lay.block_bytes, lay.n_blocks, _dev(torch, pairs))
    torch.cuda.synchronize()
    for c, d in zip(caches, dev):
        want = c.copy()
        oracle.block_copy(want, lay.block_bytes, lay.n_blocks, pairs)
        assert np.array_equal(d.cpu().numpy(), want)


# ---------------------------------------------------------------- attention vs the F32 CPU result

DECODE_CASES = [
    # seq_lens, n_head, n_head_kv, d
    ([1, 15, 16, 17, 100, 333], 32, 8, 128),        # ragged, block-edge lengths
    ([1024] * 8, 32, 8, 128),                        # BASELINE config 1 shape
    ([700, 64, 2049], 64, 8, 128),                   # 70B GQA ratio 8
    ([129, 40], 8, 8, 128),                          # MHA (ratio 1): CUDA-core single-query path
    ([77, 300], 12, 4, 64),                          # d = 64, ratio 3
    ([600, 17, 2100], 64, 4, 128),                   # ratio 16 (405B-class): two virtual kv heads of 8 per cache head on stream-K
    ([333, 1500], 48, 2, 128),                       # ratio 24: three virtual kv heads
]


@pytest.mark.parametrize("kernel", [1, 2], ids=["generic", "streamk"])
@pytest.mark.parametrize("t", TYPES)
@pytest.mark.parametrize("case", DECODE_CASES, ids=lambda c: f"{len(c[0])}x{max(c[0])}_h{c[1]}_{c[2]}_d{c[3]}")
def test_paged_decode(pkg, oracle, cuda, case, t, kernel):
    import torch
    seq_lens, n_head, n_head_kv, d = case
    w = make_workload(oracle, seq_lens, n_head, n_head_kv, d, t, seed=1234 + t)
    got = _run_paged(pkg, torch, w, kernel)
    want = oracle_attention(oracle, w)
    assert not np.isnan(got).any()
    err, nm = oracle.max_rel_err(want, got), oracle.nmse(want, got)
    assert err <= TOL[t], f"max rel err {err:.3e} (nmse {nm:.3e})"
    assert nm <= 5e-4


@pytest.mark.parametrize("t", TYPES)
def test_paged_alibi_equals_the_dense_mask_form(pkg, oracle, cuda, t):
    """ALiBi through a block table (max_bias > 0 without a mask): the kernel derives the bias the reference's F16 mask
    carries, -|pos(cell) - pos(query)| (src/llama-graph.cpp:403-477), from the cell's index in its sequence.  Checked
    against the oracle's dense-mask form fed with exactly that mask; decode rows and rows in the middle of a sequence."""
    import torch
    lens = [300, 17, 2600, 64]
    rows = [(0, 299), (1, 16), (2, 2599), (3, 63), (2, 1000), (0, 5)]
    w = make_workload(oracle, lens, 16, 4, 128, t, seed=310 + t, q_rows=rows)
    mask = w.dense_mask().astype(np.float32)
    for i, (s_, pos) in enumerate(rows):
        p = np.arange(pos + 1)
        cells = w.block_table[s_, p // w.block_size].astype(np.int64) * w.block_size + p % w.block_size
        mask[i, cells] = -np.abs(p - pos)
    want = oracle.attention(w.q, w.k_cache, w.v_cache, t, t, w.n_head_kv, w.scale, mask=mask.astype(np.float16), n_kv=w.n_cells, max_bias=8.0)
    lay = pkg.KVLayout(t, w.n_head_kv, w.d, w.block_size, w.n_blocks)
    q = _dev(torch, w.q)
    out = torch.full_like(q, float("nan"))
    call = pkg.AttnCall(q, _dev(torch, w.k_cache.reshape(-1)), _dev(torch, w.v_cache.reshape(-1)), lay, out, w.scale,
                        block_table=_dev(torch, w.block_table), seq_lens=_dev(torch, w.seq_lens), q_seq=_dev(torch, w.q_seq),
                        q_pos=_dev(torch, w.q_pos), max_bias=8.0)
    assert call.supported()
    call.run(); torch.cuda.synchronize()
    got = out.cpu().numpy()
    assert not np.isnan(got).any()
    assert oracle.max_rel_err(want, got) <= TOL[t] and oracle.nmse(want, got) <= 5e-4


KV_SHARD_CASES = [
    # a KV-head shard of a GQA model (config 4 sharded over 8 / 4 / 2 GPUs): 1, 2 or 4 kv heads in the call
    ([700, 64, 2049, 0, 33, 1], 8, 1, 128),          # one kv head, gqa 8: the 8 warps of a CTA are 8 stream-K items
    ([9000, 300, 5000], 8, 2, 128),                  # two kv heads, gqa 4; rows long enough to span several CTAs
    ([129, 4097, 16, 777], 32, 4, 128),              # four kv heads, gqa 8
    ([40000], 8, 1, 128),                            # one long row over the whole machine
]


@pytest.mark.parametrize("t", TYPES)
@pytest.mark.parametrize("case", KV_SHARD_CASES, ids=lambda c: f"{len(c[0])}x{max(c[0])}_h{c[1]}_{c[2]}")
def test_paged_decode_kv_head_shard(pkg, oracle, cuda, case, t):
    """fewer than 8 kv heads in the call (what one rank of the KV-head-sharded config 4 computes): the warps of a CTA are then
    several stream-K items instead of the 8 heads of one; same bars as every other decode case, and agreement with the
    CUDA-core kernel"""
    import torch
    seq_lens, n_head, n_head_kv, d = case
    w = make_workload(oracle, seq_lens, n_head, n_head_kv, d, t, seed=77 + t)
    got = _run_paged(pkg, torch, w, pkg.KERNEL_STREAMK)
    want = oracle_attention(oracle, w)
    live = np.array(seq_lens) > 0
    assert np.isnan(got[~live]).all() and not np.isnan(got[live]).any()
    err, nm = oracle.max_rel_err(want[live], got[live]), oracle.nmse(want[live], got[live])
    assert err <= TOL[t], f"max rel err {err:.3e} (nmse {nm:.3e})"
    assert nm <= 5e-4
    gen = _run_paged(pkg, torch, w, pkg.KERNEL_GENERIC)
    assert oracle.max_rel_err(gen[live], got[live]) <= 2e-3


@pytest.mark.parametrize("kernel", [1, 2], ids=["generic", "streamk"])
@pytest.mark.parametrize("t", [F16, Q4_0])
def test_paged_decode_empty_sequences(pkg, oracle, cuda, t, kernel):
    """a row whose sequence holds no token: the reference divides 0 by a zero row sum (ggml-cpu.c:12418) -> NaN row;
    its neighbours are untouched by it (stream-K: such rows own no segment; the combine kernel writes them)"""
    import torch
    w = make_workload(oracle, [0, 33, 0, 700, 0], 32, 8, 128, t, seed=5)
    got = _run_paged(pkg, torch, w, kernel)
    want = oracle_attention(oracle, w)
    for r in (0, 2, 4):
        assert np.isnan(got[r]).all() and np.isnan(want[r]).all()
    live = [1, 3]
    assert not np.isnan(got[live]).any()
    assert oracle.max_rel_err(want[live], got[live]) <= TOL[t]


@pytest.mark.parametrize("kernel", [1, 2], ids=["generic", "streamk"])
@pytest.mark.parametrize("dist", ["uniform", "normal"])
def test_paged_decode_mixed_types_and_normal_inputs(pkg, oracle, cuda, dist, kernel):
    import torch
    w = make_workload(oracle, [513, 90, 1200], 32, 8, 128, Q8_0, F16, seed=77, dist=dist)
    got = _run_paged(pkg, torch, w, kernel)
    want = oracle_attention(oracle, w)
    assert oracle.max_rel_err(want, got) <= 1e-2


MIXED = [(F16, Q8_0), (F16, Q4_0), (Q8_0, Q4_0), (Q4_0, Q8_0), (Q4_0, F16)]


@pytest.mark.parametrize("tk,tv", MIXED, ids=lambda x: str(x))
def test_paged_decode_every_mixed_type_pair(pkg, oracle, cuda, tk, tv):
    """-ctk / -ctv chosen independently (common/arg.cpp cache-type-k / cache-type-v): every pair has its own tile ring
    (stage size, stage count); long and short rows so the ring wraps and rows split across warps"""
    import torch
    w = make_workload(oracle, [2500, 17, 640, 1, 333], 32, 8, 128, tk, tv, seed=101)
    got = _run_paged(pkg, torch, w, 2)
    want = oracle_attention(oracle, w)
    assert oracle.max_rel_err(want, got) <= 1e-2


@pytest.mark.parametrize("kernel", [1, 2], ids=["generic", "streamk"])
@pytest.mark.parametrize("t", TYPES)
def test_chunked_prefill_rows(pkg, oracle, cuda, t, kernel):
    """multi-token rows: each query sees positions <= its own (causal through q_pos), two sequences interleaved"""
    import torch
    seq_lens = [90, 41]
    rows = [(0, p) for p in range(58, 90)] + [(1, p) for p in range(30, 41)] + [(0, 0), (1, 0)]
    w = make_workload(oracle, seq_lens, 8, 2, 128, t, seed=9, q_rows=rows)
    got = _run_paged(pkg, torch, w, kernel)
    want = oracle_attention(oracle, w)
    assert oracle.max_rel_err(want, got) <= TOL[t]


def test_without_host_hints(pkg, oracle, cuda):
    """max_seq_len / total_tokens are only scheduling hints: results must not depend on them"""
    import torch
    w = make_workload(oracle, [300, 5, 1000], 32, 8, 128, Q4_0, seed=3)
    a = _run_paged(pkg, torch, w, 0, hints=True)
    b = _run_paged(pkg, torch, w, 0, hints=False)
    want = oracle_attention(oracle, w)
    assert oracle.max_rel_err(want, a) <= 1e-2 and oracle.max_rel_err(want, b) <= 1e-2


@pytest.mark.parametrize("t", TYPES)
def test_dense_mask_mode_matches_reference_semantics(pkg, oracle, cuda, t):
    """block_table == NULL: flat cache + F16 mask, finite mask values added, -inf cells skipped"""
    import torch
    rng = np.random.default_rng(21)
    n_q, n_head, n_head_kv, d, n_kv = 5, 16, 4, 128, 512
    q = rng.uniform(-1, 1, (n_q, n_head, d)).astype(np.float32)
    kf = rng.uniform(-1, 1, (n_kv * n_head_kv, d)).astype(np.float32)
    vf = rng.uniform(-1, 1, (n_kv * n_head_kv, d)).astype(np.float32)
    k = oracle.from_float(t, kf).reshape(-1); v = oracle.from_float(t, vf).reshape(-1)
    mask = rng.uniform(-1, 1, (64, n_kv)).astype(np.float16)
    mask[:, 400:] = -np.inf
    mask[2, 10:200] = -np.inf
    lay = pkg.KVLayout(t, n_head_kv, d, 16, n_kv // 16)
    for softcap in (0.0, 10.0):
        got = pkg.dense_attention(_dev(torch, q), _dev(torch, k), _dev(torch, v), lay, n_kv, _dev(torch, mask), 0.0884,
                                  softcap=softcap).cpu().numpy()
        want = oracle.attention(q, k, v, t, t, n_head_kv, 0.0884, mask=mask, n_kv=n_kv, softcap=softcap)
        assert oracle.max_rel_err(want, got) <= TOL[t]


def test_dense_alibi(pkg, oracle, cuda):
    import torch
    rng = np.random.default_rng(22)
    n_q, n_head, n_head_kv, d, n_kv = 3, 8, 8, 128, 256
    q = rng.uniform(-1, 1, (n_q, n_head, d)).astype(np.float32)
    k = oracle.from_float(F16, rng.uniform(-1, 1, (n_kv * n_head_kv, d)).astype(np.float32)).reshape(-1)
    v = oracle.from_float(F16, rng.uniform(-1, 1, (n_kv * n_head_kv, d)).astype(np.float32)).reshape(-1)
    mask = rng.uniform(-1, 1, (64, n_kv)).astype(np.float16)
    lay = pkg.KVLayout(F16, n_head_kv, d, 16, n_kv // 16)
    got = pkg.dense_attention(_dev(torch, q), _dev(torch, k), _dev(torch, v), lay, n_kv, _dev(torch, mask), 0.0884,
                              max_bias=8.0).cpu().numpy()
    want = oracle.attention(q, k, v, F16, F16, n_head_kv, 0.0884, mask=mask, n_kv=n_kv, max_bias=8.0)
    assert oracle.max_rel_err(want, got) <= 2e-3


@pytest.mark.parametrize("t", TYPES)
def test_dense_mask_on_the_tensor_core_kernel_equals_paged(pkg, oracle, cuda, t):
    """An unmodified ggml graph (flat cache + the dense F16 mask libllama fills) runs on the stream-K kernel through
    per-row tile lists scanned from the mask: same cache bytes, same rows as the block-table form -> same result as
    the oracle, and as the CUDA-core kernel; the launch count shows which kernels ran (scan + plan + main + combine)."""
    import torch
    w = make_workload(oracle, [1, 15, 16, 17, 100, 333, 2049, 700], 32, 8, 128, t, seed=31)
    mask = w.dense_mask()                                   # [pad64(n_q), n_cells]: 0 on the row's own cells, -inf elsewhere
    lay = pkg.KVLayout(t, w.n_head_kv, w.d, w.block_size, w.n_blocks)
    q, kc, vc, m = _dev(torch, w.q), _dev(torch, w.k_cache.reshape(-1)), _dev(torch, w.v_cache.reshape(-1)), _dev(torch, mask)
    want = oracle_attention(oracle, w)
    outs = {}
    for kernel, n_launch in ((pkg.KERNEL_AUTO, 4), (pkg.KERNEL_STREAMK, 4), (pkg.KERNEL_GENERIC, None)):
        out = torch.full_like(q, float("nan"))
        call = pkg.AttnCall(q, kc, vc, lay, out, w.scale, mask=m, n_kv=w.n_cells, kernel=kernel)
        assert call.supported()
        before = pkg.lib.b200_launch_count()
        call.run()
        torch.cuda.synchronize()
        if n_launch is not None:
            assert pkg.lib.b200_launch_count() - before == n_launch, "the dense-mask call did not run on the stream-K kernels"
        outs[kernel] = out.cpu().numpy()
        assert oracle.max_rel_err(want, outs[kernel]) <= TOL[t], f"kernel {kernel}"
    assert np.array_equal(outs[pkg.KERNEL_AUTO], outs[pkg.KERNEL_STREAMK])


def test_dense_mask_edge_rows(pkg, oracle, cuda):
    """a row whose mask hides every cell gives NaN like the reference (1/0 * 0, ggml-cpu.c:12418); no mask at all: every
    row sees every cell; cells hidden inside a visible tile and biases (finite mask values) are honoured"""
    import torch
    rng = np.random.default_rng(33)
    n_q, n_head, n_head_kv, d, n_kv = 6, 32, 8, 128, 1024
    q = rng.uniform(-1, 1, (n_q, n_head, d)).astype(np.float32)
    k = oracle.from_float(Q8_0, rng.uniform(-1, 1, (n_kv * n_head_kv, d)).astype(np.float32)).reshape(-1)
    v = oracle.from_float(Q8_0, rng.uniform(-1, 1, (n_kv * n_head_kv, d)).astype(np.float32)).reshape(-1)
    lay = pkg.KVLayout(Q8_0, n_head_kv, d, 16, n_kv // 16)
    mask = np.full((64, n_kv), -np.inf, dtype=np.float16)
    mask[0, 5:6] = 0.0                                       # one cell
    mask[1, ::3] = 0.0                                       # every tile mixed
    mask[2, 100:900] = rng.uniform(-2, 0, 800).astype(np.float16)   # biases
    mask[3, :] = 0.0                                         # everything, all tiles clean
    mask[4, 16:32] = 0.0; mask[4, 512:1024] = 0.0            # clean tiles only, with a gap
    # row 5 stays hidden entirely
    got = pkg.dense_attention(_dev(torch, q), _dev(torch, k), _dev(torch, v), lay, n_kv, _dev(torch, mask), 0.0884,
                              kernel=pkg.KERNEL_STREAMK).cpu().numpy()
    want = oracle.attention(q, k, v, Q8_0, Q8_0, n_head_kv, 0.0884, mask=mask, n_kv=n_kv)
    assert np.isnan(got[5]).all() and np.isfinite(got[:5]).all()
    assert oracle.max_rel_err(want[:5], got[:5]) <= 1e-2
    # no mask: mixed K/V types are allowed too
    vf = oracle.from_float(F16, rng.uniform(-1, 1, (n_kv * n_head_kv, d)).astype(np.float32)).reshape(-1)
    out = torch.empty((n_q, n_head, d), dtype=torch.float32, device="cuda")
    call = pkg.AttnCall(_dev(torch, q), _dev(torch, k), _dev(torch, vf), lay, out, 0.0884, layout_v=pkg.KVLayout(F16, n_head_kv, d, 16, n_kv // 16),
                        n_kv=n_kv, kernel=pkg.KERNEL_STREAMK)
    assert call.supported()
    call.run()
    torch.cuda.synchronize()
    want = oracle.attention(q, k, vf, Q8_0, F16, n_head_kv, 0.0884, n_kv=n_kv)
    assert oracle.max_rel_err(want, out.cpu().numpy()) <= 1e-2


# ---------------------------------------------------------------- end to end: write -> fork/CoW -> attend

@pytest.mark.parametrize("t", TYPES)
def test_write_fork_cow_attend(pkg, oracle, cuda, t):
    """config 3 in miniature: chunked prefill into the paged cache through PagedKVCache + b200_kv_write, fork a
    shared prefix at a non-block-aligned point, append to both branches (copy-on-write through b200_block_copy),
    attend; cache bytes must equal the oracle's replay bit for bit and outputs meet the tolerance."""
    import torch
    n_head, n_head_kv, d, bs, n_blocks = 8, 2, 128, 16, 64
    lay = pkg.KVLayout(t, n_head_kv, d, bs, n_blocks)
    pc = pkg.PagedKVCache(n_blocks, bs)
    rng = np.random.default_rng(31)
    kc, vc = lay.alloc(), lay.alloc()
    hk = np.zeros((lay.n_cells, lay.token_bytes), dtype=np.uint8); hv = hk.copy()

    def write(seq, n):
        slots, cow = pc.append(seq, n)
        if cow:
            pairs = np.asarray(cow, dtype=np.int32)
            pkg.block_copy([kc, vc], lay.block_bytes, n_blocks, _dev(torch, pairs))
            oracle.block_copy(hk, lay.block_bytes, n_blocks, pairs); oracle.block_copy(hv, lay.block_bytes, n_blocks, pairs)
        ks = rng.uniform(-1, 1, (n, n_head_kv * d)).astype(np.float32)
        vs = rng.uniform(-1, 1, (n, n_head_kv * d)).astype(np.float32)
        sm = np.asarray(slots, dtype=np.int32)
        pkg.kv_write(_dev(torch, ks), _dev(torch, vs), _dev(torch, sm), kc, vc, lay)
        oracle.kv_write(ks, sm, hk, t, n_head_kv, d, lay.token_bytes, lay.head_bytes, lay.n_cells)
        oracle.kv_write(vs, sm, hv, t, n_head_kv, d, lay.token_bytes, lay.head_bytes, lay.n_cells)
        return cow

    for chunk in (32, 32, 36):                       # 100-token prompt in chunks
        assert write(0, chunk) == []
    assert pc.seq_cp(0, 1, 0, 90)                    # fork at token 90: block 5 (tokens 80..95) is shared, partly
    assert pc.ref_count(pc.export([0], 8)[0][0][5]) == 2
    cow1 = write(1, 7)                               # branch appends -> copy-on-write of the shared tail block
    assert len(cow1) == 1
    cow0 = write(0, 3)                               # parent's tail block (block 6) was never shared
    assert cow0 == []
    torch.cuda.synchronize()
    assert np.array_equal(kc.cpu().numpy().reshape(hk.shape), hk)
    assert np.array_equal(vc.cpu().numpy().reshape(hv.shape), hv)
    (bt, lens) = pc.export([0, 1], 8)
    assert lens == [103, 97]
    q = rng.uniform(-1, 1, (2, n_head, d)).astype(np.float32)
    btn, sln = np.asarray(bt, dtype=np.int32), np.asarray(lens, dtype=np.int32)
    got = pkg.paged_attention(_dev(torch, q), kc, vc, lay, _dev(torch, btn), _dev(torch, sln), 0.0884).cpu().numpy()
    want = oracle.attention(q, hk, hv, t, t, n_head_kv, 0.0884, block_table=btn, seq_lens=sln, block_size=bs)
    assert oracle.max_rel_err(want, got) <= TOL[t]


# ---------------------------------------------------------------- full-size properties (no oracle needed)

@pytest.mark.parametrize("t", [F16, Q8_0, Q4_0])
def test_full_size_properties_config2(pkg, oracle, cuda, t):
    """BASELINE config 2 at full size (64 seqs x U{512..8192}): (a) output is invariant under a re-permutation
    of physical blocks, (b) the two kernels agree, (c) sampled rows meet the oracle tolerance,
    (d) V == const gives exactly that const (softmax weights sum to 1)."""
    import torch
    lens = mixed_ctx()
    n_head, n_head_kv, d, bs = 32, 8, 128, 16
    rng = np.random.default_rng(8)
    total = int(lens.sum())
    nb = int(((lens + bs - 1) // bs).sum())
    lay = pkg.KVLayout(t, n_head_kv, d, bs, nb)
    kc, vc = lay.alloc(), lay.alloc()

    def build(perm_seed):
        ids = np.random.default_rng(perm_seed).permutation(nb).astype(np.int32)
        bt = np.zeros((len(lens), int(((lens + bs - 1) // bs).max())), dtype=np.int32)
        slot = np.empty(total, dtype=np.int32); o = 0; ob = 0
        for s, n in enumerate(lens):
            n = int(n); b = (n + bs - 1) // bs
            bt[s, :b] = ids[ob:ob + b]; ob += b
            p = np.arange(n); slot[o:o + n] = bt[s, p // bs] * bs + p % bs; o += n
        return bt, slot

    gen = torch.Generator(device="cuda"); gen.manual_seed(5)
    ks = torch.rand((total, n_head_kv * d), generator=gen, device=cuda) * 2 - 1
    vs = torch.rand((total, n_head_kv * d), generator=gen, device=cuda) * 2 - 1
    q = torch.rand((len(lens), n_head, d), generator=gen, device=cuda) * 2 - 1
    sl = _dev(torch, lens)
    outs = []
    for seed in (1, 2):
        bt, slot = build(seed)
        pkg.kv_write(ks, vs, _dev(torch, slot), kc, vc, lay)
        for kernel in (1, 2):
            out = torch.empty_like(q)
            call = pkg.AttnCall(q, kc, vc, lay, out, 0.0884, block_table=_dev(torch, bt), seq_lens=sl,
                                max_seq_len=int(lens.max()), total_tokens=total, kernel=kernel)
            if not call.supported():
                continue
            call.run(); torch.cuda.synchronize()
            outs.append(out.cpu().numpy())
    # generic (FP32 CUDA cores) vs stream-K (half-precision mma operands, FP32 accumulation): the half rounding of
    # Q, P and the dequantised V is ~2.4e-4 rms per element, so kernel-to-kernel agreement is bounded by the F16 bar
    for o in outs[1:]:
        assert oracle.max_rel_err(outs[0], o) <= 2e-3
    # sampled rows against the oracle (3 shortest sequences keep the CPU time in seconds)
    bt, slot = build(2)
    rows = np.argsort(lens)[:3]
    hk, hv = kc.cpu().numpy().reshape(lay.n_cells, -1), vc.cpu().numpy().reshape(lay.n_cells, -1)
    want = oracle.attention(q.cpu().numpy()[rows], hk, hv, t, t, n_head_kv, 0.0884, block_table=bt[rows],
                            seq_lens=lens[rows], block_size=bs)
    assert oracle.max_rel_err(want, outs[-1][rows]) <= TOL[t]
    # constant V
    pkg.kv_write(None, torch.full_like(vs, 0.5), _dev(torch, slot), None, vc, lay)
    out = pkg.paged_attention(q, kc, vc, lay, _dev(torch, bt), sl, 0.0884, max_seq_len=int(lens.max()), total_tokens=total)
    torch.cuda.synchronize()
    c = oracle.to_float(t, oracle.from_float(t, np.full((1, d), 0.5, dtype=np.float32)), d)[0, 0]
    assert np.abs(out.cpu().numpy() - c).max() <= 2e-3 * abs(c)


def test_full_size_properties_config4(pkg, oracle, cuda):
    """BASELINE config 4 at its real shape on one GPU: Llama-3-70B heads (64 q / 8 kv, d 128), 32 sequences x 32 768
    tokens, F16 (2 x 2 GiB of cache per layer).  Every row splits into dozens of stream-K segments per kv head here, so
    this is where the combine's numerics matter most: (a) bit-identical output under a re-permutation of the physical
    blocks, (b) agreement with the FP32 CUDA-core kernel (unsplit rows), (c) sampled rows against the oracle,
    (d) V == const gives that const."""
    import torch
    n_seq, ctx, n_head, n_head_kv, d, bs = 32, 32768, 64, 8, 128, 16
    bps = ctx // bs
    nb = n_seq * bps
    lay = pkg.KVLayout(F16, n_head_kv, d, bs, nb)
    kc, vc = lay.alloc(), lay.alloc()
    lens = np.full(n_seq, ctx, dtype=np.int32)
    sl = _dev(torch, lens)
    gen = torch.Generator(device="cuda"); gen.manual_seed(44)
    q = torch.rand((n_seq, n_head, d), generator=gen, device=cuda) * 2 - 1
    scale = 1.0 / np.sqrt(d)

    def fill(perm_seed, const_v=None):
        ids = np.random.default_rng(perm_seed).permutation(nb).astype(np.int32).reshape(n_seq, bps)
        g = torch.Generator(device="cuda")
        pos = np.arange(ctx)
        for s in range(n_seq):                                # one sequence at a time: 2 x 128 MiB of F32 rows
            g.manual_seed(1000 + s)                           # the same rows whatever the permutation
            ks = torch.rand((ctx, n_head_kv * d), generator=g, device=cuda) * 2 - 1
            vs = torch.rand((ctx, n_head_kv * d), generator=g, device=cuda) * 2 - 1
            if const_v is not None:
                vs.fill_(const_v)
            slot = ids[s, pos // bs].astype(np.int64) * bs + pos % bs
            pkg.kv_write(ks, vs, _dev(torch, slot.astype(np.int32)), kc, vc, lay)
        return ids

    outs = {}
    for seed in (1, 2):
        bt = fill(seed)
        for kernel in ((pkg.KERNEL_STREAMK, pkg.KERNEL_GENERIC) if seed == 2 else (pkg.KERNEL_STREAMK,)):
            out = torch.full_like(q, float("nan"))
            call = pkg.AttnCall(q, kc, vc, lay, out, scale, block_table=_dev(torch, bt), seq_lens=sl, max_seq_len=ctx,
                                total_tokens=n_seq * ctx, kernel=kernel)
            assert call.supported()
            call.run(); torch.cuda.synchronize()
            outs[(seed, kernel)] = out.cpu().numpy()
    a, b, gnr = outs[(1, pkg.KERNEL_STREAMK)], outs[(2, pkg.KERNEL_STREAMK)], outs[(2, pkg.KERNEL_GENERIC)]
    assert np.isfinite(a).all() and np.array_equal(a, b), "stream-K output depends on the physical block placement"
    assert oracle.max_rel_err(gnr, b) <= 2e-3, "split rows (stream-K + combine) disagree with unsplit rows (CUDA cores)"
    # sampled rows against the oracle: that sequence's cells gathered into a compact cache with an identity table
    for row in (0, 17):
        cells = (torch.from_numpy(bt[row].astype(np.int64)).to(cuda)[:, None] * bs + torch.arange(bs, device=cuda)[None, :]).reshape(-1)
        hk = kc.view(lay.n_cells, -1)[cells].cpu().numpy(); hv = vc.view(lay.n_cells, -1)[cells].cpu().numpy()
        want = oracle.attention(q[row:row + 1].cpu().numpy(), hk, hv, F16, F16, n_head_kv, scale,
                                block_table=np.arange(bps, dtype=np.int32)[None, :], seq_lens=np.array([ctx], dtype=np.int32), block_size=bs)
        assert oracle.max_rel_err(want, b[row:row + 1]) <= TOL[F16]
    fill(2, const_v=0.5)
    out = pkg.paged_attention(q, kc, vc, lay, _dev(torch, bt), sl, scale, max_seq_len=ctx, total_tokens=n_seq * ctx)
    torch.cuda.synchronize()
    assert np.abs(out.cpu().numpy() - 0.5).max() <= 2e-3 * 0.5


@pytest.mark.parametrize("t", [Q8_0, Q4_0])
def test_one_32k_sequence_quantised_70b_heads(pkg, oracle, cuda, t):
    """one sequence of 32 768 tokens at the 70B head shape with a quantised cache: a single row spread over the whole
    machine (hundreds of segments merged by the combine kernel) against the oracle"""
    import torch
    w = make_workload(oracle, [32768], 64, 8, 128, t, seed=45)
    got = _run_paged(pkg, torch, w, pkg.KERNEL_STREAMK)
    assert oracle.max_rel_err(oracle_attention(oracle, w), got) <= TOL[t]


def test_removed_kernel_id_is_refused(pkg, oracle, cuda):
    """kernel id 4 (the tcgen05 formulation of Q8_0 decode, removed in round 2) is refused, not silently re-routed"""
    import torch
    w = make_workload(oracle, [128], 32, 8, 128, Q8_0, seed=2)
    lay = pkg.KVLayout(Q8_0, 8, 128, 16, w.n_blocks)
    q = _dev(torch, w.q)
    call = pkg.AttnCall(q, _dev(torch, w.k_cache.reshape(-1)), _dev(torch, w.v_cache.reshape(-1)), lay, torch.empty_like(q), w.scale,
                        block_table=_dev(torch, w.block_table), seq_lens=_dev(torch, w.seq_lens), kernel=pkg.KERNEL_STREAMK_DQ)
    assert not call.supported()
    with pytest.raises(pkg.B200Error):
        call.run()


# ---------------------------------------------------------------- the layers of a step share one schedule

@pytest.mark.parametrize("t", [F16, Q8_0])
def test_decode_step_reuses_the_plan_across_layers(pkg, oracle, cuda, t):
    """b200_decode_step over L layers that share seq_lens / block table / workspace launches ONE plan kernel (layer 0)
    and gives bit-identical results to L independent calls; an explicit B200_ATTN_REUSE_PLAN call does the same."""
    import torch
    L = 3
    ws_ = [make_workload(oracle, [300, 17, 1000, 64], 32, 8, 128, t, seed=40 + l) for l in range(L)]
    w0 = ws_[0]
    lay = pkg.KVLayout(t, w0.n_head_kv, w0.d, w0.block_size, w0.n_blocks)
    bt, sl = _dev(torch, w0.block_table), _dev(torch, w0.seq_lens)       # same lengths / tables for every layer
    qs = [_dev(torch, w.q) for w in ws_]
    kcs = [_dev(torch, w.k_cache.reshape(-1)) for w in ws_]
    vcs = [_dev(torch, w.v_cache.reshape(-1)) for w in ws_]

    def mk(l, out, **kw):
        return pkg.AttnCall(qs[l], kcs[l], vcs[l], lay, out, w0.scale, block_table=bt, seq_lens=sl,
                            max_seq_len=int(w0.seq_lens.max()), total_tokens=int(w0.seq_lens.sum()), kernel=2, **kw)

    ref_out = [torch.full_like(qs[l], float("nan")) for l in range(L)]
    n0 = pkg.lib.b200_launch_count()
    for l in range(L):
        mk(l, ref_out[l]).run()
    assert pkg.lib.b200_launch_count() - n0 == 3*L                        # plan + main + combine each
    out = [torch.full_like(qs[l], float("nan")) for l in range(L)]
    first = mk(0, out[0])
    calls = [first] + [mk(l, out[l], workspace=first.workspace) for l in range(1, L)]
    n0 = pkg.lib.b200_launch_count()
    pkg.StepCall(calls).run()
    assert pkg.lib.b200_launch_count() - n0 == 1 + 2*L
    out2 = [torch.full_like(qs[l], float("nan")) for l in range(L)]
    first2 = mk(0, out2[0])
    n0 = pkg.lib.b200_launch_count()
    first2.run()
    for l in range(1, L):
        mk(l, out2[l], workspace=first2.workspace, reuse_plan=True).run()
    assert pkg.lib.b200_launch_count() - n0 == 1 + 2*L
    torch.cuda.synchronize()
    for l in range(L):
        assert torch.equal(out[l], ref_out[l]) and torch.equal(out2[l], ref_out[l])
        want = oracle_attention(oracle, ws_[l])
        assert oracle.max_rel_err(want, out[l].cpu().numpy()) <= TOL[t]
