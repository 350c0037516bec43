"""The peer-memory exchange of the KV-head-sharded path (csrc/peer_gather.cu) on >= 2 GPUs of one box: two processes,
one per GPU; what every rank ends up with must equal NCCL's all_gather bit for bit, eagerly and as a replayed CUDA
graph (the epoch lives on the device).  Skipped on a single-GPU box; run with `gpurun --gpus 2`."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    from conftest import load_pkg
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        pkg = load_pkg()
        shape = (32, 8, 128)                                         # config 4 at G = 8: [n_seq, n_head/G, d]
        pg = pkg.PeerGather(shape, n_slots=3)
        gen = torch.Generator(device="cuda"); gen.manual_seed(100 + rank)
        for it in range(7):                                          # eager, every slot, every form, changing data
            local = torch.rand(shape, device="cuda", generator=gen)
            if it % 2:
                pg.publish(local, it % 3); pg.wait()
            else:
                pg.exchange_ll(local, it % 3)
            pg.check()
            want = pkg.gather_head_outputs(local)
            assert torch.equal(pg.gathered(it % 3), want), f"iteration {it}"
        with pytest.raises(ValueError):
            pg.publish(local, 6 % 3)                                 # same slot twice in a row is refused
        # captured: 3 exchanges per graph (slots 1, 2, 0), replayed with new data each time
        locals_ = [torch.zeros(shape, device="cuda") for _ in range(3)]
        outs = [torch.zeros((world,) + shape, device="cuda") for _ in range(3)]
        side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())

        def step():
            for i in range(3):
                if i == 1:
                    pg.publish(locals_[i], (i + 1) % 3); pg.wait()           # the two-launch form
                elif i == 0:
                    pg.exchange(locals_[i], (i + 1) % 3)                     # publish + wait in one launch
                else:
                    pg.exchange_ll(locals_[i], (i + 1) % 3)                  # payload + epoch in the same stores
                outs[i].copy_(pg.gathered((i + 1) % 3))
        e0 = pg.epoch
        with torch.cuda.stream(side):
            step()
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            step()
        pg.epoch = e0 + 3                                            # capture launched nothing
        for rep in range(5):
            for i in range(3):
                locals_[i].copy_(torch.rand(shape, device="cuda", generator=gen))
            g.replay(); pg.note_replay(3)
            pg.check()
            for i in range(3):
                assert torch.equal(outs[i], pkg.gather_head_outputs(locals_[i])), f"replay {rep} exchange {i}"
        pg.close()
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_peer_gather_equals_nccl_all_gather(tmp_path, cuda):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))
