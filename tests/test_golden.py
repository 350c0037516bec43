"""Committed golden vectors produced by the reference itself (tests/golden/make_golden.py ran the reference's
ggml CPU code, compiled from /root/reference, in the build container).  They travel to machines that have no
reference tree: the oracle is pinned to them on the CPU, and the CUDA kernels are checked against them with -m gpu."""
import os

import numpy as np
import pytest

from workloads import F16, Q4_0, Q8_0, make_workload, oracle_attention

HERE = os.path.dirname(os.path.abspath(__file__))
TOL = {F16: 2e-3, Q8_0: 1e-2, Q4_0: 1e-2}


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "ref_vectors.npz"))


def _cases():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


MG = _cases()
TYPES = [(F16, "f16"), (Q8_0, "q8_0"), (Q4_0, "q4_0")]


@pytest.mark.parametrize("t,name", TYPES)
def test_oracle_quantisers_match_reference_bytes(oracle, gold, t, name):
    x = MG.quant_rows()
    assert np.array_equal(oracle.from_float(t, x), gold[f"quant_ref_{name}"])
    assert np.array_equal(oracle.to_float(t, gold[f"quant_ref_{name}"], 128).view(np.uint32), gold[f"dequant_{name}"].view(np.uint32))
    # the CPU backend's GGML_OP_CPY (SIMD quantiser) may differ from the _ref quantiser by single codes on ties;
    # the reference's own bar is NMSE <= 1e-6 after dequantisation (tests/test-backend-ops.cpp:1469-1471)
    a = oracle.to_float(t, gold[f"cpy_cpu_{name}"], 128)
    b = oracle.to_float(t, oracle.from_float(t, x), 128)
    assert oracle.nmse(a, b) <= 1e-6


@pytest.mark.parametrize("case", MG.ATTN_CASES, ids=lambda c: c[0])
def test_oracle_attention_matches_reference_outputs(oracle, gold, case):
    name, lens, nh, nkv, d, tk, tv, seed = case
    w = make_workload(oracle, lens, nh, nkv, d, tk, tv, seed=seed)
    chk = np.array([w.q.sum(dtype=np.float64), w.k_cache.astype(np.uint64).sum(), w.v_cache.astype(np.uint64).sum()])
    assert np.allclose(chk, gold[f"attn_in_{name}"], rtol=1e-12), "regenerated inputs differ from the ones the goldens were made with"
    got = oracle_attention(oracle, w)
    assert oracle.max_rel_err(gold[f"attn_f32_{name}"], got) < 2e-5            # O1: the F32 CPU result
    assert oracle.nmse(gold[f"attn_fa_{name}"].astype(np.float32), got) < 5e-4  # O2: the reference op's own bar


@pytest.mark.gpu
@pytest.mark.parametrize("t,name", TYPES)
def test_gpu_kv_write_matches_reference_bytes(pkg, cuda, gold, t, name):
    import torch
    x = MG.quant_rows()
    n = x.shape[0]
    lay = pkg.KVLayout(t, 1, 128, 16, (n + 15) // 16)
    kc = lay.alloc()
    pkg.kv_write(torch.from_numpy(x).cuda(), None, None, kc, None, lay)
    torch.cuda.synchronize()
    got = kc.cpu().numpy()[: n * lay.token_bytes].reshape(n, -1)
    assert np.array_equal(got, gold[f"quant_ref_{name}"])


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", [1, 2], ids=["generic", "streamk"])
@pytest.mark.parametrize("case", MG.ATTN_CASES, ids=lambda c: c[0])
def test_gpu_attention_matches_reference_outputs(pkg, oracle, cuda, gold, case, kernel):
    import torch
    name, lens, nh, nkv, d, tk, tv, seed = case
    w = make_workload(oracle, lens, nh, nkv, d, tk, tv, seed=seed)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    lay_k = pkg.KVLayout(tk, nkv, d, w.block_size, w.n_blocks)
    lay_v = pkg.KVLayout(tv, nkv, d, w.block_size, w.n_blocks)
    out = torch.full(w.q.shape, float("nan"), dtype=torch.float32, device="cuda")
    call = pkg.AttnCall(dev(w.q), dev(w.k_cache.reshape(-1)), dev(w.v_cache.reshape(-1)), lay_k, out, w.scale, layout_v=lay_v,
                        block_table=dev(w.block_table), seq_lens=dev(w.seq_lens), max_seq_len=int(max(lens)),
                        total_tokens=int(sum(lens)), kernel=kernel)
    if not call.supported():
        pytest.skip("kernel does not take this shape")
    call.run()
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    tol = max(TOL[tk], TOL[tv])
    assert oracle.max_rel_err(gold[f"attn_f32_{name}"], got) <= tol
    assert oracle.nmse(gold[f"attn_fa_{name}"].astype(np.float32), got) <= 5e-4
