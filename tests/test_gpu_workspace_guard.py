"""The workspace a call is given is exactly b200_paged_attention_workspace() bytes inside a larger buffer whose surroundings
hold a pattern: no kernel of the call may write outside its workspace (the stream-K plan / arrival counters / partials,
the prefill plan / KV-split piece buffers / F16 scratch), and the call must still match the oracle."""
import numpy as np
import pytest

from test_gpu_parity import TOL, _dev
from workloads import F16, Q8_0, make_workload, oracle_attention

pytestmark = pytest.mark.gpu
GUARD = 1 << 16


def _chunk(seq, p0, p1):
    return [(seq, p) for p in range(p0, p1)]


CASES = {
    # name: (seq_lens, n_head, n_head_kv, q_rows, type, kernel)
    "streamk_split_rows":   ([300, 17, 5000, 64, 1, 2049], 32, 8, None, Q8_0, 2),
    "streamk_gqa16":        ([600, 17, 2100], 64, 4, None, F16, 2),
    "streamk_one_kv_head":  ([9000, 300], 8, 1, None, F16, 2),
    "prefill_split_pieces": ([2048], 8, 2, _chunk(0, 1792, 2048), F16, 3),
    "prefill_quantised":    ([2048, 700], 8, 2, _chunk(0, 1792, 2048) + _chunk(1, 600, 700), Q8_0, 3),
    "generic_splits":       ([77, 3000], 12, 4, None, Q8_0, 1),
}


@pytest.mark.parametrize("name", list(CASES))
def test_no_write_outside_the_workspace(pkg, oracle, cuda, name):
    import torch
    seq_lens, n_head, n_head_kv, rows, t, kernel = CASES[name]
    w = make_workload(oracle, seq_lens, n_head, n_head_kv, 128, t, seed=410, q_rows=rows)
    lay = pkg.KVLayout(t, w.n_head_kv, w.d, w.block_size, w.n_blocks)
    q = _dev(torch, w.q)
    out = torch.full_like(q, float("nan"))
    kw = dict(block_table=_dev(torch, w.block_table), seq_lens=_dev(torch, w.seq_lens), max_seq_len=int(w.seq_lens.max()),
              total_tokens=int(w.seq_lens.sum()), kernel=kernel)
    if rows is not None:
        kw.update(q_seq=_dev(torch, w.q_seq), q_pos=_dev(torch, w.q_pos))
    kc, vc = _dev(torch, w.k_cache.reshape(-1)), _dev(torch, w.v_cache.reshape(-1))
    probe = pkg.AttnCall(q, kc, vc, lay, out, w.scale, **kw)             # only to learn the size
    if not probe.supported():
        pytest.skip("kernel does not take this shape")
    need = (probe.desc.workspace_bytes + 255) // 256 * 256
    big = torch.full((GUARD + need + GUARD,), 0xA5, dtype=torch.uint8, device=cuda)
    call = pkg.AttnCall(q, kc, vc, lay, out, w.scale, workspace=big[GUARD:GUARD + need], **kw)
    call.desc.workspace_bytes = probe.desc.workspace_bytes               # exactly what the library asked for
    for _ in range(2):                                                   # twice: the second call finds the first one's leftovers
        call.run()
    torch.cuda.synchronize()
    assert bool((big[:GUARD] == 0xA5).all()) and bool((big[GUARD + need:] == 0xA5).all()), "a kernel wrote outside its workspace"
    got = out.cpu().numpy()
    live = np.array([(w.seq_lens[i if w.q_seq is None else w.q_seq[i]] > 0) for i in range(w.n_q)])
    assert oracle.max_rel_err(oracle_attention(oracle, w)[live], got[live]) <= TOL[t]
