"""Multi-GPU host logic on CPU: world_size-2 gloo processes.  The per-shard attention is computed
by the CPU oracle here (no GPU in this container); what is under test is the partitioning and the gather:
  A. sequence-batch sharding: the union of the ranks' outputs equals the unsharded result, no collective used
  B. KV-head sharding: per-rank head slices + ONE all_gather == the unsharded result, bit for bit
"""
import os
import socket

import numpy as np
import pytest

from workloads import F16, Q8_0, make_workload, oracle_attention, row_bytes


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, mode, out_dir):
    import torch
    import torch.distributed as dist
    from conftest import load_pkg
    from oracle.loader import Oracle
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pkg, orc = load_pkg(), Oracle()
        lens = [70, 5, 33, 129, 16, 48, 1]
        w = make_workload(orc, lens, 16, 4, 128, Q8_0 if mode == "heads" else F16, seed=11)
        full = oracle_attention(orc, w)
        if mode == "heads":
            sh = pkg.head_shard(rank, world, w.n_head, w.n_head_kv)
            hb = row_bytes(w.type_k, w.d)
            # this rank's cache holds only its kv heads of every cell (what a KV-head-sharded GPU stores)
            kc = np.ascontiguousarray(w.k_cache.reshape(w.n_cells, w.n_head_kv, hb)[:, sh.kv0:sh.kv0 + sh.n_head_kv]).reshape(w.n_cells, -1)
            vc = np.ascontiguousarray(w.v_cache.reshape(w.n_cells, w.n_head_kv, hb)[:, sh.kv0:sh.kv0 + sh.n_head_kv]).reshape(w.n_cells, -1)
            q = np.ascontiguousarray(w.q[:, sh.q0:sh.q0 + sh.n_head])
            local = orc.attention(q, kc, vc, w.type_k, w.type_v, sh.n_head_kv, w.scale, block_table=w.block_table,
                                  seq_lens=w.seq_lens, block_size=w.block_size)
            gathered = pkg.gather_head_outputs(torch.from_numpy(local))
            assert tuple(gathered.shape) == (world, len(lens), w.n_head // world, w.d)
            merged = pkg.merge_gathered(gathered).numpy()
            assert np.array_equal(merged.view(np.uint32), full.view(np.uint32)), "head-sharded result differs"
        else:
            owned = pkg.shard_sequences(lens, world)
            mine = owned[rank]
            sub = orc.attention(np.ascontiguousarray(w.q[mine]), w.k_cache, w.v_cache, w.type_k, w.type_v, w.n_head_kv, w.scale,
                                block_table=np.ascontiguousarray(w.block_table[mine]), seq_lens=np.ascontiguousarray(w.seq_lens[mine]),
                                block_size=w.block_size)
            assert np.array_equal(sub.view(np.uint32), full[mine].view(np.uint32)), "sequence-sharded result differs"
            # the assignment is a partition and every rank derived the same one without talking
            flat = sorted(i for o in owned for i in o)
            assert flat == list(range(len(lens)))
            objs = [None] * world
            dist.all_gather_object(objs, owned)
            assert all(o == owned for o in objs)
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["sequences", "heads"])
def test_world2_gloo(tmp_path, mode):
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, mode, str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))


def test_shard_sequences_balance_and_determinism(pkg):
    rng = np.random.default_rng(7)
    lens = rng.integers(512, 8193, size=64)
    for world in (1, 2, 4, 8):
        owned = pkg.shard_sequences(lens, world)
        assert sorted(i for o in owned for i in o) == list(range(64))
        loads = [int(lens[o].sum()) for o in owned]
        assert max(loads) - min(loads) <= int(lens.max())          # LPT bound
        assert owned == pkg.shard_sequences(list(lens), world)
    with pytest.raises(ValueError):
        pkg.shard_sequences(lens, 0)


def test_head_shard_ranges(pkg):
    for world in (1, 2, 4, 8):
        seen_q, seen_kv = [], []
        for r in range(world):
            s = pkg.head_shard(r, world, 64, 8)
            assert s.n_head == 64 // world and s.n_head_kv == 8 // world
            assert s.q0 // 8 == s.kv0                                   # query head h -> kv head h // gqa stays on the rank
            seen_q += range(s.q0, s.q0 + s.n_head); seen_kv += range(s.kv0, s.kv0 + s.n_head_kv)
        assert seen_q == list(range(64)) and seen_kv == list(range(8))
    with pytest.raises(ValueError):
        pkg.head_shard(0, 3, 64, 8)
    with pytest.raises(ValueError):
        pkg.head_shard(2, 2, 64, 8)
