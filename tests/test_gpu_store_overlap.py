"""Store overlap (attn_streamk.cu): in b200_decode_step / b200_store_attention the stream-K main kernel of a pure decode
layer is started UNDER that layer's kv_write and waits for it only in front of each row's last tile.  Results must be
those of "store, synchronise, attend" — against the oracle on the finished cache, and bit for bit against the same
library run step by step — and the device-side check of the slot mapping must switch the overlap off when a stored row
does not go to its own sequence's last cell."""
import ctypes as C

import numpy as np
import pytest

from workloads import F16, Q4_0, Q8_0, make_workload, oracle_attention

pytestmark = pytest.mark.gpu

TOL = {F16: 2e-3, Q8_0: 1e-2, Q4_0: 1e-2}
# lengths: one cell, a tile boundary either side, rows split over many warps, the newest cell first / last in its tile
LENS = [300, 17, 1000, 64, 1, 16, 33, 4096, 2049]


def _dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


class _Layers:
    """L layers of one decode step: caches whose newest cell per sequence still holds poison, the F32 rows that belong
    there, and the descriptors to store them and attend."""

    def __init__(self, pkg, oracle, torch, t, L, lens=LENS, slots=None):
        self.ws_ = [make_workload(oracle, lens, 32, 8, 128, t, seed=70 + l) for l in range(L)]
        w0 = self.ws_[0]
        self.lay = pkg.KVLayout(t, w0.n_head_kv, w0.d, w0.block_size, w0.n_blocks)
        self.bt, self.sl = _dev(torch, w0.block_table), _dev(torch, w0.seq_lens)
        last = np.cumsum(w0.seq_lens.astype(np.int64)) - 1                   # token index of each sequence's newest row
        self.last_slots = w0.slot_mapping[last].astype(np.int32)
        self.slots = _dev(torch, self.last_slots if slots is None else slots)
        self.full_k = [_dev(torch, w.k_cache) for w in self.ws_]             # the finished caches
        self.full_v = [_dev(torch, w.v_cache) for w in self.ws_]
        self.src_k = [_dev(torch, w.k_f32[last]) for w in self.ws_]
        self.src_v = [_dev(torch, w.v_f32[last]) for w in self.ws_]
        self.qs = [_dev(torch, w.q) for w in self.ws_]
        self.kc = [torch.empty_like(c) for c in self.full_k]
        self.vc = [torch.empty_like(c) for c in self.full_v]
        self.poison_idx = _dev(torch, self.last_slots.astype(np.int64))
        self.pkg, self.torch, self.L, self.w0 = pkg, torch, L, w0

    def reset(self):
        """stream-ordered: finished caches with the newest cells poisoned again"""
        for l in range(self.L):
            self.kc[l].copy_(self.full_k[l]); self.vc[l].copy_(self.full_v[l])
            self.kc[l][self.poison_idx] = 0x7b; self.vc[l][self.poison_idx] = 0x7b

    def calls(self, outs, q_seq=None, q_pos=None):
        pkg, w0 = self.pkg, self.w0
        first = None
        calls, writes = [], []
        for l in range(self.L):
            c = pkg.AttnCall(self.qs[l], self.kc[l].view(-1), self.vc[l].view(-1), self.lay, outs[l], w0.scale, block_table=self.bt,
                             seq_lens=self.sl, q_seq=q_seq, q_pos=q_pos, max_seq_len=int(w0.seq_lens.max()), total_tokens=int(w0.seq_lens.sum()),
                             kernel=pkg.KERNEL_STREAMK, workspace=None if first is None else first.workspace)
            first = first or c
            calls.append(c)
            writes.append(pkg.kv_write_desc(self.src_k[l], self.src_v[l], self.slots, self.kc[l].view(-1), self.vc[l].view(-1), self.lay))
        return calls, writes


@pytest.mark.parametrize("t", [F16, Q8_0, Q4_0])
def test_decode_step_overlapping_the_store_equals_store_then_attend(pkg, oracle, cuda, t):
    import torch
    L = 4
    ly = _Layers(pkg, oracle, torch, t, L)
    # step by step: store, wait, attend, wait
    want = [torch.full_like(q, float("nan")) for q in ly.qs]
    calls, writes = ly.calls(want)
    ly.reset()
    for l in range(L):
        assert calls[l].supported()
        pkg.check(pkg.lib.b200_kv_write(C.byref(writes[l]), C.c_void_p(pkg.current_stream())), "b200_kv_write")
        torch.cuda.synchronize()
        calls[l].run()
        torch.cuda.synchronize()
        assert torch.equal(ly.kc[l], ly.full_k[l]) and torch.equal(ly.vc[l], ly.full_v[l]), "the stored rows are not the workload's"
        assert oracle.max_rel_err(oracle_attention(oracle, ly.ws_[l]), want[l].cpu().numpy()) <= TOL[t]
    # one call, eight times back to back with the newest cells poisoned again in between (all stream-ordered, no host wait)
    outs = [[torch.full_like(q, float("nan")) for q in ly.qs] for _ in range(8)]
    steps = []
    for rep in range(8):
        calls, writes = ly.calls(outs[rep])
        steps.append(pkg.StepCall(calls, writes))
    for rep in range(8):
        ly.reset()
        steps[rep].run()
    torch.cuda.synchronize()
    plan = steps[-1]._keep[0][0].workspace.view(torch.int32)[:4].cpu().numpy()
    assert plan[3] == 1, "the plan kernel did not accept the slot mapping: the overlap never ran"
    for rep in range(8):
        for l in range(L):
            assert torch.equal(outs[rep][l], want[l]), f"repetition {rep} layer {l}: differs from store-then-attend"
    for l in range(L):
        assert torch.equal(ly.kc[l], ly.full_k[l]) and torch.equal(ly.vc[l], ly.full_v[l])


def test_store_attention_single_call_and_skipped_rows(pkg, oracle, cuda):
    """b200_store_attention (one layer; what a fused paged FLASH_ATTN_EXT node calls) with rows kv_write skips (slot -1 and
    out of range): those sequences read what the cache holds"""
    import torch
    ly = _Layers(pkg, oracle, torch, Q8_0, 1)
    slots = ly.last_slots.copy()
    slots[1] = -1; slots[3] = ly.w0.n_cells + 5
    ly.slots = _dev(torch, slots)
    keep = torch.ones(len(slots), dtype=torch.bool); keep[1] = keep[3] = False
    ly.poison_idx = ly.poison_idx[keep.cuda()]                       # rows 1 and 3 are complete already
    out = [torch.full_like(ly.qs[0], float("nan"))]
    calls, writes = ly.calls(out)
    ly.reset()
    pkg.check(pkg.lib.b200_store_attention(C.byref(writes[0]), C.byref(calls[0].desc), C.c_void_p(pkg.current_stream())), "b200_store_attention")
    torch.cuda.synchronize()
    assert calls[0].workspace.view(torch.int32)[3].item() == 1
    assert torch.equal(ly.kc[0], ly.full_k[0])
    assert oracle.max_rel_err(oracle_attention(oracle, ly.ws_[0]), out[0].cpu().numpy()) <= TOL[Q8_0]


def test_store_into_other_cells_switches_the_overlap_off(pkg, oracle, cuda):
    """rows stored into cells that are NOT their sequence's newest one (here: the first cell of the next sequence): the
    plan kernel's check fails, the main kernel waits in front of everything, and every sequence sees the overwritten cell"""
    import torch
    t = F16
    lens = [300, 170, 1000, 64, 33]
    w = make_workload(oracle, lens, 32, 8, 128, t, seed=91)
    first = np.concatenate([[0], np.cumsum(w.seq_lens.astype(np.int64))[:-1]])
    slots = np.roll(w.slot_mapping[first], -1).astype(np.int32)                # row r -> first cell of sequence r+1
    ly = _Layers(pkg, oracle, torch, t, 2, lens=lens, slots=slots)
    ly.poison_idx = _dev(torch, slots.astype(np.int64))
    outs = [torch.full_like(q, float("nan")) for q in ly.qs]
    calls, writes = ly.calls(outs)
    ly.reset()
    pkg.StepCall(calls, writes).run()
    torch.cuda.synchronize()
    assert calls[0].workspace.view(torch.int32)[3].item() == 0, "the plan kernel accepted a slot mapping it must refuse"
    for l in range(2):
        wl = ly.ws_[l]
        hk, hv = wl.k_cache.copy(), wl.v_cache.copy()
        last = np.cumsum(wl.seq_lens.astype(np.int64)) - 1
        oracle.kv_write(wl.k_f32[last], slots, hk, t, wl.n_head_kv, wl.d, ly.lay.token_bytes, ly.lay.head_bytes, wl.n_cells)
        oracle.kv_write(wl.v_f32[last], slots, hv, t, wl.n_head_kv, wl.d, ly.lay.token_bytes, ly.lay.head_bytes, wl.n_cells)
        assert np.array_equal(ly.kc[l].cpu().numpy(), hk)
        want = oracle.attention(wl.q, hk, hv, t, t, wl.n_head_kv, wl.scale, block_table=wl.block_table, seq_lens=wl.seq_lens,
                                block_size=wl.block_size)
        assert oracle.max_rel_err(want, outs[l].cpu().numpy()) <= TOL[t]


@pytest.mark.parametrize("identity", [True, False], ids=["q_seq_identity", "q_seq_permuted"])
def test_decode_ubatch_with_q_seq_and_q_pos(pkg, oracle, cuda, identity):
    """the form a llama_kv_cache host emits for a decode ubatch: q_seq and q_pos given.  With q_seq the identity and every
    row at its sequence's newest position the overlap runs; with the rows in another order of sequences the plan kernel
    must refuse it (another row could read the cell under the store) — and both give store-then-attend's results"""
    import torch
    t = Q8_0
    lens = [300, 170, 1000, 64, 33, 2049]
    n = len(lens)
    perm = np.arange(n) if identity else np.array([1, 0, 2, 3, 5, 4])
    rows = [(int(s), lens[int(s)] - 1) for s in perm]
    ws_ = [make_workload(oracle, lens, 32, 8, 128, t, seed=120 + l, q_rows=rows) for l in range(2)]
    w0 = ws_[0]
    lay = pkg.KVLayout(t, 8, 128, 16, w0.n_blocks)
    last = np.cumsum(w0.seq_lens.astype(np.int64)) - 1
    slots = w0.slot_mapping[last][perm].astype(np.int32)                          # stored row r = query row r
    bt, sl = _dev(torch, w0.block_table), _dev(torch, w0.seq_lens)
    q_seq, q_pos = _dev(torch, w0.q_seq), _dev(torch, w0.q_pos)
    slots_d, poison = _dev(torch, slots), _dev(torch, slots.astype(np.int64))
    outs, calls, writes, caches = [], [], [], []
    for l, w in enumerate(ws_):
        kc, vc = _dev(torch, w.k_cache), _dev(torch, w.v_cache)
        kc[poison] = 0x7b; vc[poison] = 0x7b
        out = torch.full((n, 32, 128), float("nan"), device=cuda)
        c = pkg.AttnCall(_dev(torch, w.q), kc.view(-1), vc.view(-1), lay, out, w0.scale, block_table=bt, seq_lens=sl, q_seq=q_seq, q_pos=q_pos,
                         max_seq_len=max(lens), total_tokens=sum(lens), kernel=pkg.KERNEL_STREAMK, workspace=calls[0].workspace if calls else None)
        calls.append(c); outs.append(out); caches.append((kc, vc))
        src = (_dev(torch, w.k_f32[last][perm]), _dev(torch, w.v_f32[last][perm]))   # the descriptor holds raw pointers: keep the rows alive
        caches[-1] += src
        writes.append(pkg.kv_write_desc(src[0], src[1], slots_d, kc.view(-1), vc.view(-1), lay))
    pkg.StepCall(calls, writes).run()
    torch.cuda.synchronize()
    assert calls[0].workspace.view(torch.int32)[3].item() == (1 if identity else 0)
    for l, w in enumerate(ws_):
        assert np.array_equal(caches[l][0].cpu().numpy(), w.k_cache) and np.array_equal(caches[l][1].cpu().numpy(), w.v_cache)
        assert oracle.max_rel_err(oracle_attention(oracle, w), outs[l].cpu().numpy()) <= TOL[t]
