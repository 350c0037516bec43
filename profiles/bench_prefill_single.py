#!/usr/bin/env python
"""profiles/bench_prefill_single.py — under gpurun: ONE sequence's prefill chunk (what llama.cpp's own prompt processing
submits: n_ubatch rows of one sequence) through b200_paged_attention with the tile kernel, CUDA-event timed.  The chunk is
16-64 work items on 148 SMs, the case the KV split of pf_main_kernel exists for (A/B: B200_LIB=<-DPF_SPLIT_MAX=1 build>).
Prints one JSON line per shape.  F16 cache, random contents (timing only; parity is tests/test_gpu_prefill.py)."""
import importlib.util
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("b200pkg", os.path.join(ROOT, "llama.cpp_test_b200", "__init__.py"),
                                              submodule_search_locations=[os.path.join(ROOT, "llama.cpp_test_b200")])
pkg = importlib.util.module_from_spec(spec); sys.modules["b200pkg"] = pkg; spec.loader.exec_module(pkg)


def run(ctx, chunk, n_head=32, n_head_kv=8, d=128, bs=16, iters=30):
    dev = torch.device("cuda:0")
    n_blocks = (ctx + bs - 1) // bs + 2
    lay = pkg.KVLayout(pkg.TYPE_F16 if hasattr(pkg, "TYPE_F16") else 1, n_head_kv, d, bs, n_blocks)
    g = torch.Generator(device=dev); g.manual_seed(1)
    kc = (torch.randn(lay.total_bytes // 2, device=dev, generator=g) * 0.5).half().view(torch.uint8)
    vc = (torch.randn(lay.total_bytes // 2, device=dev, generator=g) * 0.5).half().view(torch.uint8)
    bt = torch.randperm(n_blocks, device=dev, generator=g).to(torch.int32)[None, :].contiguous()
    sl = torch.tensor([ctx], dtype=torch.int32, device=dev)
    q = torch.randn(chunk, n_head, d, device=dev, generator=g)
    out = torch.empty_like(q)
    q_seq = torch.zeros(chunk, dtype=torch.int32, device=dev)
    q_pos = torch.arange(ctx - chunk, ctx, dtype=torch.int32, device=dev)
    call = pkg.AttnCall(q, kc, vc, lay, out, 1.0 / np.sqrt(d), block_table=bt, seq_lens=sl, q_seq=q_seq, q_pos=q_pos,
                        max_seq_len=ctx, total_tokens=ctx, kernel=3)
    assert call.supported()
    for _ in range(3):
        call.run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        call.run()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    vis = sum(min(ctx, p + 1) for p in range(ctx - chunk, ctx))
    flops = 4.0 * n_head * d * vis
    print(json.dumps({"ctx": ctx, "chunk": chunk, "n_head": n_head, "n_head_kv": n_head_kv, "us_per_op": round(us, 1),
                      "tflops": round(flops / us / 1e6, 1), "lib": os.environ.get("B200_LIB", "product")}))


if __name__ == "__main__":
    for ctx, chunk in ((4096, 512), (8192, 512), (32768, 512), (2048, 2048), (16384, 2048)):
        run(ctx, chunk)
