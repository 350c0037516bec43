"""The ggml op-node boundary (include/b200_ggml_abi.h) on the GPU: b200_flash_attn_ext / b200_cpy / b200_supports_op
driven with `struct ggml_tensor` nodes built the way the reference's own operator tests build them.

Mirrors tests/test-backend-ops.cpp of the reference:
  * test_flash_attn_ext (:3197-3272; grid :4367-4391): q F32 [hs, nb, nh*nr], k/v type_KV [hs, kv, nh], mask F16
    [kv, PAD(nb, 64)], out = ggml_flash_attn_ext(q, k, v, m, 1/sqrt(hs), max_bias, logit_softcap); the same tensors
    permuted {0, 2, 1, 3} (the strided views a KV cache gives, llama-graph.cpp:1215-1245); bar NMSE <= 5e-4 (:3216-3218);
  * test_cpy (:1443-1491): F32 -> type_dst, bar NMSE <= 1e-6 (:1469-1471) — here byte-exact against the _ref quantisers.
The reference compares a backend with its CPU backend on the same random tensors; the CPU side here is the oracle, which
tests/test_oracle_vs_ref.py pins to that CPU backend."""
import ctypes as C

import numpy as np
import pytest

from workloads import F16, F32, Q4_0, Q8_0

pytestmark = pytest.mark.gpu

TYPES = [F16, Q8_0, Q4_0]
TOL = {F16: 2e-3, Q8_0: 1e-2, Q4_0: 1e-2}
ELEM = {F32: 4, F16: 2}


def _node(pkg, type_id, ne, nb, data, op=0):
    t = pkg.GgmlTensor()
    t.type, t.op = type_id, op
    for i in range(4):
        t.ne[i] = ne[i] if i < len(ne) else 1
        t.nb[i] = nb[i] if i < len(nb) else (t.nb[i - 1] * t.ne[i - 1])
    t.data = data
    return t


def _f32bits(x):
    return int(np.float32(x).view(np.int32))


class FlashAttnCase:
    """one test_flash_attn_ext instance: host arrays, the op node over device copies, the oracle's answer"""

    def __init__(self, pkg, oracle, torch, dev, hs, nh, nr, kv, nb, mask, max_bias, softcap, t, permuted, seed):
        rng = np.random.default_rng(seed)
        self.pkg, self.t = pkg, t
        n_head, rb = nh * nr, pkg.row_bytes(t, hs)
        q = rng.uniform(-1, 1, (nb, n_head, hs)).astype(np.float32)            # init_tensor_uniform (:70-120)
        kf = rng.uniform(-1, 1, (kv * nh, hs)).astype(np.float32)
        vf = rng.uniform(-1, 1, (kv * nh, hs)).astype(np.float32)
        k = oracle.from_float(t, kf).reshape(kv, nh, rb)                       # token-major bytes [kv][nh][row]
        v = oracle.from_float(t, vf).reshape(kv, nh, rb)
        pad = (nb + 63) // 64 * 64                                             # GGML_KQ_MASK_PAD
        m = None
        if mask:
            m = rng.uniform(-1, 1, (pad, kv)).astype(np.float16)
            m[0, kv // 3: kv // 2] = -np.inf                                   # what a causal / per-sequence mask holds
        self.want = oracle.attention(q, k.reshape(kv, -1), v.reshape(kv, -1), t, t, nh, 1.0 / np.sqrt(hs),
                                     mask=m, n_kv=kv, max_bias=max_bias, softcap=softcap)
        # device tensors in the storage order the reference test allocates them
        if permuted:                                                            # storage [hs, nh, kv] / [hs, n_head, nb]
            qd, q_nb = q, (4, n_head * hs * 4, hs * 4)
            kd, vd, kv_nb = k, v, (None, nh * rb, rb)
        else:                                                                   # storage [hs, kv, nh] / [hs, nb, n_head]
            qd, q_nb = np.ascontiguousarray(q.transpose(1, 0, 2)), (4, hs * 4, nb * hs * 4)
            kd, vd = np.ascontiguousarray(k.transpose(1, 0, 2)), np.ascontiguousarray(v.transpose(1, 0, 2))
            kv_nb = (None, rb, kv * rb)
        blk_nb0 = {F16: 2, Q8_0: 34, Q4_0: 18}[t]
        self.bufs = [torch.from_numpy(a.view(np.uint8).reshape(-1)).to(dev) for a in (qd, kd, vd)]
        self.out = torch.full((nb, n_head, hs), float("nan"), dtype=torch.float32, device=dev)
        self.q = _node(pkg, F32, (hs, nb, n_head, 1), q_nb, self.bufs[0].data_ptr())
        self.k = _node(pkg, t, (hs, kv, nh, 1), (blk_nb0,) + kv_nb[1:], self.bufs[1].data_ptr())
        self.v = _node(pkg, t, (hs, kv, nh, 1), (blk_nb0,) + kv_nb[1:], self.bufs[2].data_ptr())
        self.dst = _node(pkg, F32, (hs, n_head, nb, 1), (4, hs * 4, n_head * hs * 4), self.out.data_ptr(),
                         op=pkg.GGML_OP_FLASH_ATTN_EXT)
        self.dst.src[0], self.dst.src[1], self.dst.src[2] = C.pointer(self.q), C.pointer(self.k), C.pointer(self.v)
        if mask:
            self.mbuf = torch.from_numpy(m.view(np.uint8).reshape(-1)).to(dev)
            self.m = _node(pkg, F16, (kv, pad, 1, 1), (2, kv * 2), self.mbuf.data_ptr())
            self.dst.src[3] = C.pointer(self.m)
        self.dst.op_params[0] = _f32bits(1.0 / np.sqrt(hs))
        self.dst.op_params[1] = _f32bits(max_bias)
        self.dst.op_params[2] = _f32bits(softcap)
        self.dst.op_params[3] = 1                                              # GGML_PREC_F32 as llama-graph.cpp:1240 sets it

    def run(self, torch):
        lib = self.pkg.lib
        need = lib.b200_flash_attn_ext_workspace(C.byref(self.dst))
        ws = torch.empty(max(need, 16), dtype=torch.uint8, device=self.out.device)
        rc = lib.b200_flash_attn_ext(C.byref(self.dst), ws.data_ptr(), need, None)
        torch.cuda.synchronize()
        return rc, self.out.cpu().numpy()


def _grid(hs):
    """the reference's loop nest (tests/test-backend-ops.cpp:4367-4391) for one head size, BF16 left out
    (ggml-cuda.cu:3233-3236 refuses it as well)"""
    for mask in (True, False):
        for max_bias in (0.0, 8.0):
            if not mask and max_bias > 0.0:
                continue
            for softcap in (0.0, 10.0):
                if hs != 128 and softcap != 0.0:
                    continue
                for nr in (1, 4, 16):
                    if nr == 16 and hs != 128:
                        continue
                    for kv in (512, 1024):
                        if nr != 1 and kv != 512:
                            continue
                        for nb in (1, 3, 32, 35):
                            yield dict(mask=mask, max_bias=max_bias, softcap=softcap, nr=nr, kv=kv, nb=nb, permuted=False)
                            if mask and max_bias == 0.0 and softcap == 0.0 and kv == 512:
                                yield dict(mask=mask, max_bias=max_bias, softcap=softcap, nr=nr, kv=kv, nb=nb, permuted=True)


@pytest.mark.parametrize("t", TYPES)
@pytest.mark.parametrize("hs", [64, 80, 128, 256])
def test_flash_attn_ext_node_reference_grid(pkg, oracle, cuda, hs, t):
    import torch
    if hs % 32 and t != F16:
        hs = 96                  # GGML_PAD(hs, block size): what the reference's grid builds for a quantised type (test-backend-ops.cpp:3233)
    bad, n = [], 0
    for i, c in enumerate(_grid(hs)):
        case = FlashAttnCase(pkg, oracle, torch, cuda, hs, 4, c["nr"], c["kv"], c["nb"], c["mask"], c["max_bias"],
                             c["softcap"], t, c["permuted"], seed=1234 + i)
        assert pkg.lib.b200_supports_op(C.byref(case.dst)) == 1, c
        rc, got = case.run(torch)
        assert rc == 0, (c, pkg.lib.b200_last_error().decode())
        n += 1
        nmse, rel = oracle.nmse(case.want, got), oracle.max_rel_err(case.want, got)
        if not (nmse <= 5e-4 and rel <= TOL[t]) or np.isnan(got).any():
            bad.append((c, nmse, rel))
    assert not bad, f"{len(bad)} of {n}: {bad[:4]}"
    assert n == (108 if hs == 128 else 44)                   # every case of the grid ran


@pytest.mark.parametrize("hs", [72, 160])
def test_flash_attn_ext_node_unsupported_head_sizes_are_declined(pkg, oracle, cuda, hs):
    """supports_op says no and the entry returns a status instead of computing: the backend keeps its own kernel"""
    import torch
    case = FlashAttnCase(pkg, oracle, torch, cuda, hs, 4, 1, 512, 3, True, 0.0, 0.0, F16, False, seed=5)
    assert pkg.lib.b200_supports_op(C.byref(case.dst)) == 0
    rc, got = case.run(torch)
    assert rc != 0 and np.isnan(got).all()


def test_flash_attn_ext_node_paged_extension(pkg, oracle, cuda):
    """src[4..7] + op_params[4..6]: the same node with a block table instead of a mask equals the descriptor path"""
    import torch
    from workloads import make_workload, oracle_attention
    w = make_workload(oracle, [130, 7, 64], 32, 8, 128, Q8_0, seed=17)
    want = oracle_attention(oracle, w)
    dev = cuda
    n_q, n_head, d = w.q.shape
    rb = pkg.row_bytes(Q8_0, d)
    bufs = [torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
            for a in (w.q, w.k_cache, w.v_cache, w.block_table.astype(np.int32), w.seq_lens.astype(np.int32))]
    out = torch.full((n_q, n_head, d), float("nan"), dtype=torch.float32, device=dev)
    q = _node(pkg, F32, (d, n_q, n_head, 1), (4, n_head * d * 4, d * 4), bufs[0].data_ptr())
    k = _node(pkg, Q8_0, (d, w.n_cells, 8, 1), (34, 8 * rb, rb), bufs[1].data_ptr())
    v = _node(pkg, Q8_0, (d, w.n_cells, 8, 1), (34, 8 * rb, rb), bufs[2].data_ptr())
    bt = _node(pkg, pkg.GGML_TYPE_I32, (w.block_table.shape[1], w.block_table.shape[0], 1, 1),
               (4, w.block_table.shape[1] * 4), bufs[3].data_ptr())
    sl = _node(pkg, pkg.GGML_TYPE_I32, (len(w.seq_lens), 1, 1, 1), (4,), bufs[4].data_ptr())
    dst = _node(pkg, F32, (d, n_head, n_q, 1), (4, d * 4, n_head * d * 4), out.data_ptr(), op=pkg.GGML_OP_FLASH_ATTN_EXT)
    for i, s in ((0, q), (1, k), (2, v), (4, bt), (5, sl)):
        dst.src[i] = C.pointer(s)
    dst.op_params[0] = _f32bits(w.scale)
    dst.op_params[4] = w.block_size
    assert pkg.lib.b200_supports_op(C.byref(dst)) == 1
    need = pkg.lib.b200_flash_attn_ext_workspace(C.byref(dst))
    ws = torch.empty(max(need, 16), dtype=torch.uint8, device=dev)
    assert pkg.lib.b200_flash_attn_ext(C.byref(dst), ws.data_ptr(), need, None) == 0, pkg.lib.b200_last_error().decode()
    torch.cuda.synchronize()
    assert oracle.max_rel_err(want, out.cpu().numpy()) <= 1e-2
    # a paged node must not also carry a mask: it would be dropped silently (the block table only says "causal prefix")
    mbuf = torch.zeros(64 * w.n_cells, dtype=torch.float16, device=dev)
    m = _node(pkg, F16, (w.n_cells, 64, 1, 1), (2, w.n_cells * 2), mbuf.data_ptr())
    dst.src[3] = C.pointer(m)
    assert pkg.lib.b200_supports_op(C.byref(dst)) == 0
    assert pkg.lib.b200_flash_attn_ext(C.byref(dst), ws.data_ptr(), need, None) != 0
    assert "must not carry a mask" in pkg.lib.b200_last_error().decode()


@pytest.mark.parametrize("t", [F16, Q8_0, Q4_0])
def test_flash_attn_ext_node_paged_with_fused_kv_store(pkg, oracle, cuda, t):
    """the paged node as the libllama patch emits it (llama.cpp_test_b200/host/libllama_paged.patch): src[3] = I32 slot
    mapping, src[8] / src[9] = this step's K / V rows; the store runs first, then the attention sees the new token.  The
    cache starts WITHOUT the last token of every sequence; afterwards its bytes equal the oracle's full cache and the
    output equals the oracle's attention over it."""
    import torch
    from workloads import make_workload, oracle_attention
    w = make_workload(oracle, [130, 7, 64, 33], 32, 8, 128, t, seed=19)
    want = oracle_attention(oracle, w)
    dev = cuda
    n_q, n_head, d = w.q.shape
    rb = pkg.row_bytes(t, d)
    last = np.cumsum(w.seq_lens.astype(np.int64)) - 1                     # row of each sequence's last token in k_f32 / slot_mapping
    slots_last = w.slot_mapping[last].astype(np.int32)
    k_cache, v_cache = w.k_cache.copy(), w.v_cache.copy()
    k_cache[slots_last] = 0; v_cache[slots_last] = 0                      # the step's tokens are not in the cache yet
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
    bq, bk, bv = up(w.q), up(k_cache), up(v_cache)
    bbt, bsl, bslot = up(w.block_table.astype(np.int32)), up(w.seq_lens.astype(np.int32)), up(slots_last)
    bkn, bvn = up(w.k_f32[last]), up(w.v_f32[last])
    out = torch.full((n_q, n_head, d), float("nan"), dtype=torch.float32, device=dev)
    nb0 = {F16: 2, Q8_0: 34, Q4_0: 18}[t]
    q = _node(pkg, F32, (d, n_q, n_head, 1), (4, n_head * d * 4, d * 4), bq.data_ptr())
    k = _node(pkg, t, (d, w.n_cells, 8, 1), (nb0, 8 * rb, rb), bk.data_ptr())
    v = _node(pkg, t, (d, w.n_cells, 8, 1), (nb0, 8 * rb, rb), bv.data_ptr())
    bt = _node(pkg, pkg.GGML_TYPE_I32, (w.block_table.shape[1], w.block_table.shape[0], 1, 1), (4, w.block_table.shape[1] * 4), bbt.data_ptr())
    sl = _node(pkg, pkg.GGML_TYPE_I32, (len(w.seq_lens), 1, 1, 1), (4,), bsl.data_ptr())
    sm = _node(pkg, pkg.GGML_TYPE_I32, (n_q, 1, 1, 1), (4,), bslot.data_ptr())
    kn = _node(pkg, F32, (8 * d, n_q, 1, 1), (4, 8 * d * 4), bkn.data_ptr())
    vn = _node(pkg, F32, (8 * d, n_q, 1, 1), (4, 8 * d * 4), bvn.data_ptr())
    dst = _node(pkg, F32, (d, n_head, n_q, 1), (4, d * 4, n_head * d * 4), out.data_ptr(), op=pkg.GGML_OP_FLASH_ATTN_EXT)
    for i, s_ in ((0, q), (1, k), (2, v), (3, sm), (4, bt), (5, sl), (8, kn), (9, vn)):
        dst.src[i] = C.pointer(s_)
    dst.op_params[0] = _f32bits(w.scale)
    dst.op_params[4] = w.block_size
    assert pkg.lib.b200_supports_op(C.byref(dst)) == 1, pkg.lib.b200_last_error().decode()
    need = pkg.lib.b200_flash_attn_ext_workspace(C.byref(dst))
    ws = torch.empty(max(need, 16), dtype=torch.uint8, device=dev)
    assert pkg.lib.b200_flash_attn_ext(C.byref(dst), ws.data_ptr(), need, None) == 0, pkg.lib.b200_last_error().decode()
    torch.cuda.synchronize()
    assert np.array_equal(bk.cpu().numpy().reshape(w.k_cache.shape), w.k_cache), "the fused store did not write the oracle's K bytes"
    assert np.array_equal(bv.cpu().numpy().reshape(w.v_cache.shape), w.v_cache)
    assert oracle.max_rel_err(want, out.cpu().numpy()) <= TOL[t]
    # a slot mapping without rows, or rows without a slot mapping, is malformed
    dst.src[8] = dst.src[9] = None
    assert pkg.lib.b200_supports_op(C.byref(dst)) == 0


@pytest.mark.parametrize("t", [F32, F16, Q8_0, Q4_0])
@pytest.mark.parametrize("ne", [(256, 4, 4, 4), (1024, 37, 1, 1), (128, 2, 3, 1)])
def test_cpy_node_f32_to_cache_types(pkg, oracle, cuda, t, ne):
    """test_cpy(GGML_TYPE_F32, type_dst, {256, 4, 4, 4}) and the KV-store shape of llama-graph.cpp:1375-1397
    (F32 [n_embd_k_gqa, n_tokens] into a 1-D view of the cache): bytes equal the _ref quantisers'"""
    import torch
    rng = np.random.default_rng(ne[0] + ne[1])
    n = int(np.prod(ne))
    x = rng.uniform(-1, 1, (n // ne[0], ne[0])).astype(np.float32)
    want = x.view(np.uint8).reshape(-1) if t == F32 else oracle.from_float(t, x).reshape(-1)
    src_b = torch.from_numpy(x.view(np.uint8).reshape(-1)).to(cuda)
    dst_b = torch.zeros(want.size, dtype=torch.uint8, device=cuda)
    rb = pkg.row_bytes(t, ne[0])
    src = _node(pkg, F32, ne, (4, ne[0] * 4), src_b.data_ptr())
    if ne[2] == 1 and ne[1] > 4:                                               # the KV-store form: 1-D destination view
        dst = _node(pkg, t, (n, 1, 1, 1), ({F32: 4, F16: 2, Q8_0: 34, Q4_0: 18}[t], pkg.row_bytes(t, n)), dst_b.data_ptr(),
                    op=pkg.GGML_OP_CPY)
    else:
        dst = _node(pkg, t, ne, ({F32: 4, F16: 2, Q8_0: 34, Q4_0: 18}[t], rb), dst_b.data_ptr(), op=pkg.GGML_OP_CPY)
    dst.src[0], dst.src[1] = C.pointer(src), C.pointer(dst)
    assert pkg.lib.b200_supports_op(C.byref(dst)) == 1
    assert pkg.lib.b200_cpy(C.byref(src), C.byref(dst), None) == 0, pkg.lib.b200_last_error().decode()
    torch.cuda.synchronize()
    assert np.array_equal(dst_b.cpu().numpy(), want)


@pytest.mark.parametrize("perm", [(0, 2, 1, 3), (1, 0, 2, 3)], ids=["by_rows", "not_contiguous"])
@pytest.mark.parametrize("ts,td", [(F32, F32), (F32, F16), (F16, F16), (F16, F32), (F32, Q8_0), (F32, Q4_0), (F16, Q8_0)])
def test_cpy_node_permuted_sources(pkg, oracle, cuda, ts, td, perm):
    """test_cpy(type_src, type_dst, {256, 2, 3, 4}, permute) of tests/test-backend-ops.cpp:3992-4009: a permuted view of a
    contiguous F32 / F16 tensor copied into a contiguous destination of the view's shape; the bytes must equal the CPU
    semantics (element i of the view in its own order -> element i of the destination, converted with the _ref quantisers)"""
    import torch
    ne = (256, 2, 3, 4)
    rng = np.random.default_rng(7)
    base = rng.uniform(-1, 1, ne[::-1]).astype(np.float32 if ts == F32 else np.float16)     # numpy axes are (3, 2, 1, 0)
    esz = base.itemsize
    nb = (esz, ne[0] * esz, ne[0] * ne[1] * esz, ne[0] * ne[1] * ne[2] * esz)
    # ggml_permute(a, p0..p3): axis i of a becomes axis p_i of the result
    vne, vnb = [0] * 4, [0] * 4
    for i in range(4):
        vne[perm[i]], vnb[perm[i]] = ne[i], nb[i]
    if td in (Q8_0, Q4_0) and vne[0] % 32:
        pytest.skip("row length of the view is not a multiple of the quant block")
    view = np.lib.stride_tricks.as_strided(base, shape=tuple(vne[::-1]), strides=tuple(vnb[::-1]))
    flat = np.ascontiguousarray(view).astype(np.float32).reshape(-1, vne[0])
    if td == F32:
        want = flat.view(np.uint8).reshape(-1)
    elif td == F16:
        want = flat.astype(np.float16).view(np.uint8).reshape(-1)
    else:
        want = oracle.from_float(td, flat).reshape(-1)
    src_b = torch.from_numpy(base.view(np.uint8).reshape(-1)).to(cuda)
    dst_b = torch.zeros(want.size, dtype=torch.uint8, device=cuda)
    src = _node(pkg, ts, tuple(vne), tuple(vnb), src_b.data_ptr())
    dsz = {F32: 4, F16: 2, Q8_0: 34, Q4_0: 18}[td]
    rb = pkg.row_bytes(td, vne[0])
    dst = _node(pkg, td, tuple(vne), (dsz, rb, rb * vne[1], rb * vne[1] * vne[2]), dst_b.data_ptr(), op=pkg.GGML_OP_CPY)
    dst.src[0], dst.src[1] = C.pointer(src), C.pointer(dst)
    assert pkg.lib.b200_supports_op(C.byref(dst)) == 1, pkg.lib.b200_last_error().decode()
    assert pkg.lib.b200_cpy(C.byref(src), C.byref(dst), None) == 0, pkg.lib.b200_last_error().decode()
    torch.cuda.synchronize()
    assert np.array_equal(dst_b.cpu().numpy(), want)


@pytest.mark.parametrize("ts", [Q8_0, Q4_0])
def test_cpy_node_quantised_to_f32(pkg, oracle, cuda, ts):
    """the cast of build_rope_shift (src/llama-context.cpp:476-480): Q8_0 / Q4_0 rows -> F32 = dequantize_row_*"""
    import torch
    rng = np.random.default_rng(8)
    x = rng.uniform(-1, 1, (24, 256)).astype(np.float32)
    q = oracle.from_float(ts, x)
    want = oracle.to_float(ts, q, 256).astype(np.float32)
    src_b = torch.from_numpy(q.reshape(-1)).to(cuda)
    dst_b = torch.zeros(want.size * 4, dtype=torch.uint8, device=cuda)
    rb = pkg.row_bytes(ts, 256)
    src = _node(pkg, ts, (256, 4, 3, 2), ({Q8_0: 34, Q4_0: 18}[ts], rb, rb * 4, rb * 12), src_b.data_ptr())
    dst = _node(pkg, F32, (256, 4, 3, 2), (4, 1024, 4096, 12288), dst_b.data_ptr(), op=pkg.GGML_OP_CPY)
    dst.src[0], dst.src[1] = C.pointer(src), C.pointer(dst)
    assert pkg.lib.b200_supports_op(C.byref(dst)) == 1
    assert pkg.lib.b200_cpy(C.byref(src), C.byref(dst), None) == 0, pkg.lib.b200_last_error().decode()
    torch.cuda.synchronize()
    assert np.array_equal(dst_b.cpu().numpy().view(np.float32).reshape(want.shape), want)


@pytest.mark.parametrize("t", [F16, Q8_0, Q4_0])
def test_cpy_node_same_type_block_move(pkg, oracle, cuda, t):
    """build_kv_self_defrag's moves (llama-context.cpp:670-707): same-type 2-D row views, bytes copied"""
    import torch
    rng = np.random.default_rng(3)
    x = oracle.from_float(t, rng.uniform(-1, 1, (48, 1024)).astype(np.float32))          # 48 cells of a 1024-wide layer
    src_b = torch.from_numpy(x.reshape(-1)).to(cuda)
    dst_b = torch.zeros_like(src_b)
    rb = pkg.row_bytes(t, 1024)
    nb0 = {F16: 2, Q8_0: 34, Q4_0: 18}[t]
    src = _node(pkg, t, (1024, 16, 1, 1), (nb0, rb), src_b.data_ptr() + 8 * rb)                 # cells 8..23
    dst = _node(pkg, t, (1024, 16, 1, 1), (nb0, rb), dst_b.data_ptr() + 30 * rb, op=pkg.GGML_OP_CPY)   # -> cells 30..45
    dst.src[0], dst.src[1] = C.pointer(src), C.pointer(dst)
    assert pkg.lib.b200_supports_op(C.byref(dst)) == 1
    assert pkg.lib.b200_cpy(C.byref(src), C.byref(dst), None) == 0, pkg.lib.b200_last_error().decode()
    torch.cuda.synchronize()
    got = dst_b.cpu().numpy().reshape(48, rb)
    assert np.array_equal(got[30:46], x.reshape(48, rb)[8:24]) and not got[:30].any() and not got[46:].any()
