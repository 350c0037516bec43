"""b200_paged_attention_static_ws (include/b200_paged_attn.h): the stream-K decode path on the library's static
__device__ workspace — the entry point for callers that cannot call cudaMalloc.  It has no counterpart in the
reference, so its checker is the existing entry point:
b200_paged_attention with a caller-provided workspace runs the same kernels on the same plan, hence the outputs must
be bit-identical (asserted), which is inside the tolerance the parity tests use against the oracle (also asserted,
TOL below).  Two calls in a row with different shapes share the one array: the second must not see the first's
leftovers.  Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

from workloads import F16, Q4_0, Q8_0, make_workload, mixed_ctx, oracle_attention

pytestmark = pytest.mark.gpu

TOL = {F16: 2e-3, Q8_0: 1e-2, Q4_0: 1e-2}       # same bars as tests/test_gpu_parity.py


def _dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _both(pkg, torch, w):
    """(output of b200_paged_attention with a torch-allocated workspace, output of the static-workspace entry)"""
    lay = pkg.KVLayout(w.type_k, w.n_head_kv, w.d, w.block_size, w.n_blocks)
    q = _dev(torch, w.q)
    kc, vc = _dev(torch, w.k_cache.reshape(-1)), _dev(torch, w.v_cache.reshape(-1))
    bt, sl = _dev(torch, w.block_table), _dev(torch, w.seq_lens)
    want = torch.full_like(q, float("nan"))
    call = pkg.AttnCall(q, kc, vc, lay, want, w.scale, block_table=bt, seq_lens=sl, max_seq_len=int(w.seq_lens.max()),
                        total_tokens=int(w.seq_lens.sum()), kernel=pkg.KERNEL_STREAMK)
    assert call.supported()
    call.run()
    torch.cuda.synchronize()

    got = torch.full_like(q, float("nan"))
    desc = pkg.AttnDesc.from_buffer_copy(call.desc)          # same descriptor ...
    desc.dst = got.data_ptr()
    desc.workspace, desc.workspace_bytes = None, 0           # ... without a workspace: the entry point must ignore these
    desc.kernel = pkg.KERNEL_AUTO
    need = pkg.lib.b200_paged_attention_workspace(C.byref(call.desc))
    assert 0 < need <= pkg.lib.b200_static_workspace_bytes()
    before = pkg.lib.b200_launch_count()
    pkg.check(pkg.lib.b200_paged_attention_static_ws(C.byref(desc), C.c_void_p(pkg.current_stream())),
              "b200_paged_attention_static_ws")
    torch.cuda.synchronize()
    assert pkg.lib.b200_launch_count() >= before + 3, "plan + main + combine were not launched"
    return want.cpu().numpy(), got.cpu().numpy()


def test_static_workspace_two_calls_in_a_row(pkg, oracle, cuda):
    import torch
    # call 1: ragged Q8_0 batch with rows long enough to be split into several stream-K segments (partials in the
    # array); call 2: another type and shape on the same array, right behind it
    cases = [make_workload(oracle, [4097, 33, 1, 16, 2500, 777], 32, 8, 128, Q8_0, seed=11),
             make_workload(oracle, [1500, 1500, 3000, 64], 64, 8, 128, F16, seed=12)]
    for i, w in enumerate(cases):
        want, got = _both(pkg, torch, w)
        assert np.isfinite(got).all()
        assert np.array_equal(want, got), f"call {i + 1}: static-workspace output differs from b200_paged_attention's"
        err = oracle.max_rel_err(oracle_attention(oracle, w), got)
        assert err <= TOL[w.type_k], f"call {i + 1}: max rel err {err:.3e} against the oracle"


def test_static_workspace_config2_size_fits(pkg, oracle, cuda):
    """BASELINE config 2's row and head counts (64 sequences x 32 heads, Q4_0) fit the stated maximum: the scratch
    depends on n_q and n_head, not on the context lengths, which are kept short here"""
    import torch
    w = make_workload(oracle, mixed_ctx(64, 64, 1024, seed=7), 32, 8, 128, Q4_0, seed=13)
    want, got = _both(pkg, torch, w)
    assert np.array_equal(want, got)


def test_static_workspace_rejections(pkg, oracle, cuda):
    import torch
    w = make_workload(oracle, [100, 200], 32, 8, 128, F16, seed=14)
    lay = pkg.KVLayout(w.type_k, w.n_head_kv, w.d, w.block_size, w.n_blocks)
    q = _dev(torch, w.q)
    kc, vc = _dev(torch, w.k_cache.reshape(-1)), _dev(torch, w.v_cache.reshape(-1))
    bt, sl = _dev(torch, w.block_table), _dev(torch, w.seq_lens)
    out = torch.zeros_like(q)
    call = pkg.AttnCall(q, kc, vc, lay, out, w.scale, block_table=bt, seq_lens=sl)
    st = C.c_void_p(pkg.current_stream())
    # beyond the maximum: n_q * n_head partial slots that need more than the static array -> B200_ERR_ALLOC, clear text
    big = pkg.AttnDesc.from_buffer_copy(call.desc)
    big.n_q = 4096
    n_seqs_big = torch.zeros(4096, dtype=torch.int32, device=cuda)
    bt_big = torch.zeros((4096, bt.shape[1]), dtype=torch.int32, device=cuda)
    big.seq_lens, big.block_table, big.n_seqs = n_seqs_big.data_ptr(), bt_big.data_ptr(), 4096
    assert pkg.lib.b200_paged_attention_workspace(C.byref(big)) > pkg.lib.b200_static_workspace_bytes()
    assert pkg.lib.b200_paged_attention_static_ws(C.byref(big), st) == -2
    msg = pkg.lib.b200_last_error().decode()
    assert "static workspace holds" in msg and str(pkg.lib.b200_static_workspace_bytes()) in msg
    # dense-mask mode and the other kernels are not offered
    dense = pkg.AttnDesc.from_buffer_copy(call.desc)
    dense.block_table = None
    assert pkg.lib.b200_paged_attention_static_ws(C.byref(dense), st) == -4
    gen = pkg.AttnDesc.from_buffer_copy(call.desc)
    gen.kernel = pkg.KERNEL_GENERIC
    assert pkg.lib.b200_paged_attention_static_ws(C.byref(gen), st) == -3
    assert pkg.lib.b200_paged_attention_static_ws(None, st) == -4
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0.0, "a rejected call wrote to dst"
