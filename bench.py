#!/usr/bin/env python
"""bench.py — paged-attn decode KV-read GB/s (BASELINE.json metric); default workload = BASELINE config 2.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload config2|config1|config4] [--kv-type q8_0|q4_0|f16]
                  [--impl b200|reference]

Workloads (all synthetic U(-1,1) data, 16-token blocks scattered by a seeded random permutation):
  config2  (default; the configuration the metric is quoted on) Llama-3-8B attention shape (32 q / 8 kv heads,
           d = 128, 32 layers), 64 sequences with context ~ U{512..8192} (seed 7 + rank), Q8_0 (default) / Q4_0 /
           F16 KV.  Multi-GPU: sequence-batch sharding, every rank holds its own 64 sequences => "weak".
  config1  same shape, 8 sequences x 1024 context, F16 (the reference's CPU-runnable case; latency-bound).
  config4  Llama-3-70B shape (64 q / 8 kv heads), 32 sequences x 32768 context, F16, 8 layers resident;
           KV heads sharded over the ranks + ONE NCCL all-gather of the outputs per layer => "strong".
  config3  chunked prefill: 8 sequences, each appending a 512-token chunk at position 3584 (4096 causal query rows per
           layer-op): slot-mapped KV write of the chunk + tile attention; metric = attention TFLOP/s (tensor roofline).
One STEP = one decode step of the hot path = the attention op of every layer, each over its own KV cache (GBs per
step, far beyond the 126 MB L2, so no flush is needed between iterations).

  value        whole-job KV GB/s with everything resident in HBM (CUDA-graph replay when no collective is inside)
  e2e          same metric through the C ABI the way a host uses it per step: pinned-host Q / new K,V rows / block
               table / seq_lens copied H2D (copy stream, overlapped group by group), one b200_decode_step call
               (b200_kv_write + b200_paged_attention per layer) per group of 8 layers, outputs D2H, host waits for the last byte
  roofline     op-level achieved GB/s = algorithmic KV bytes of one layer-op / CUDA-event time, vs MEASURED_PEAKS.json
  cpu_baseline the reference's own ggml CPU FLASH_ATTN_EXT (oracle/_ref) on a bounded sample of the same workload
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

TYPES = {"f16": 1, "q4_0": 2, "q8_0": 8}
D, BS = 128, 16
WORKLOADS = {
    "config2": dict(n_head=32, n_head_kv=8, n_layer=32, n_seqs=64, ctx="U{512..8192} seed 7+rank, union LPT-sharded over the ranks", shard="sequences",
                    scaling="weak", kv_type="q8_0", steps=100),
    "config1": dict(n_head=32, n_head_kv=8, n_layer=32, n_seqs=8, ctx="1024", shard="sequences", scaling="weak",
                    kv_type="f16", steps=500),
    "config3": dict(n_head=32, n_head_kv=8, n_layer=32, n_seqs=8, ctx="4096 (chunk of 512 appended at 3584)", shard="sequences",
                    scaling="weak", kv_type="f16", steps=20),
    "config4": dict(n_head=64, n_head_kv=8, n_layer=8, n_seqs=32, ctx="32768", shard="heads", scaling="strong",
                    kv_type="f16", steps=50),
    # the attention share of config 5 (llama-batched-bench -npl 256 -fa -ctk q4_0 -ctv q4_0, 32 sequences per GPU at ~1k
    # context): the model's matmuls are outside this path, so this is attention-only tok/s, not S_TG
    "config5_attn": dict(n_head=32, n_head_kv=8, n_layer=32, n_seqs=32, ctx="1024", shard="sequences", scaling="weak",
                         kv_type="q4_0", steps=300),
    # diagnostics, not bench lines: what ONE rank of config 4 computes at G = 8 / G = 4, without any exchange
    "config4_shard8": dict(n_head=8, n_head_kv=1, n_layer=8, n_seqs=32, ctx="32768", shard="sequences", scaling="weak",
                           kv_type="f16", steps=50),
    "config4_pair": dict(n_head=16, n_head_kv=2, n_layer=8, n_seqs=32, ctx="32768", shard="heads", scaling="strong",
                         kv_type="f16", steps=50),          # N = 2: each rank computes what a rank of config 4 at G = 8 computes
    "config4_shard4": dict(n_head=16, n_head_kv=2, n_layer=8, n_seqs=32, ctx="32768", shard="sequences", scaling="weak",
                           kv_type="f16", steps=50),
}


def row_bytes(t, n):
    return {1: n * 2, 8: n // 32 * 34, 2: n // 32 * 18}[t]


def ctx_lens(workload, rank, world=1):
    w = WORKLOADS[workload]
    if workload == "config2":
        draw = lambda r: np.random.default_rng(7 + r).integers(512, 8193, size=w["n_seqs"]).astype(np.int32)
        if world == 1:
            return draw(rank)
        # the job's batch is the union of every rank's 64 draws; sequence-batch sharding (partition A, sharding.py)
        # hands each rank the share that balances the summed context, so no rank waits for the unluckiest draw
        from __graft_entry__ import load
        allc = np.concatenate([draw(r) for r in range(world)])
        return allc[load().shard_sequences(allc, world)[rank]]
    return np.full(w["n_seqs"], int(w["ctx"].split()[0]), dtype=np.int32)


def block_tables(lens, seed=42):
    nb = (lens.astype(np.int64) + BS - 1) // BS
    ids = np.random.default_rng(seed).permutation(int(nb.sum())).astype(np.int32)
    bt = np.zeros((len(lens), int(nb.max())), dtype=np.int32)
    slot = np.empty(int(lens.sum()), dtype=np.int32)
    o = ob = 0
    for s, n in enumerate(lens):
        n = int(n); b = int(nb[s])
        bt[s, :b] = ids[ob:ob + b]; ob += b
        p = np.arange(n); slot[o:o + n] = bt[s, p // BS] * BS + p % BS; o += n
    return bt, slot, int(nb.sum())


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kv_type):
    """dram bytes read+written by the dominant kernel for one layer-op of config 2, from the committed
    `ncu --set full` capture (profiles/traffic.json, written by profiles/summarize.py) — None if absent"""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            return json.load(f).get(kv_type)
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)"""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        if self.index < 0:
            return
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        self.stop_flag = True
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if len(r) >= 6 and r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}




def cpu_sample(workload):
    """the bounded slice of the workload the CPU arm times: (ctx lens, n_head, n_head_kv, description)"""
    w = WORKLOADS[workload]
    lens = ctx_lens(workload, 0)
    if workload == "config4":          # 4 of the 32 sequences, 1 of the 8 kv heads (with its 8 query heads)
        return lens[:4], w["n_head"] // w["n_head_kv"], 1, "4 of 32 sequences x 1 of 8 kv heads of one layer-op"
    return lens, w["n_head"], w["n_head_kv"], f"one layer-op of the full workload (1/{w['n_layer']} of a step)"


def cpu_layer_op(workload, t, steps, warmup):
    """time the reference CPU implementation on the sample; returns (sample, GB/s, cores, kind, ms per op)"""
    from oracle.loader import Oracle, Ref
    from workloads import make_workload
    orc, ref = Oracle(), Ref.load()
    lens, n_head, n_head_kv, what = cpu_sample(workload)
    w = make_workload(orc, lens, n_head, n_head_kv, D, t, seed=1234, spare_blocks=0)
    kv_bytes = w.kv_bytes()
    if ref is not None:
        mask = w.dense_mask()
        n_iter = max(3, 1 + steps)                           # one warm-up run + the timed ones (the harness discards the first)
        _, us = ref.attention(0, w.q, w.k_cache, w.v_cache, t, t, n_head_kv, w.n_cells, mask, w.scale, n_iter=n_iter)
        kind, cores = "reference", ref.n_threads
        sample = (f"{what}: ggml CPU GGML_OP_FLASH_ATTN_EXT, flat cache of {w.n_cells} cells + dense F16 mask, "
                  f"{cores} threads, isa {ref.isa}, median of {n_iter - 1} runs after 1 warm-up run")
    else:
        t0 = time.perf_counter()
        rows = np.argsort(lens)[:4]
        orc.attention(w.q[rows], w.k_cache, w.v_cache, t, t, n_head_kv, w.scale, block_table=w.block_table[rows],
                      seq_lens=lens[rows], block_size=BS)
        us = (time.perf_counter() - t0) * 1e6
        kv_bytes = int(lens[rows].sum()) * n_head_kv * 2 * row_bytes(t, D)
        kind, cores = "port", 1
        sample = f"oracle port (scalar C), the 4 shortest sequences of: {what}"
    return sample, kv_bytes / us / 1e3, cores, kind, us / 1e3


def workload_config(workload, kv_type, world):
    w = WORKLOADS[workload]
    return {"workload": f"{workload}_{'llama3_70b' if w['n_head'] == 64 else 'llama3_8b'}_{kv_type}", "n_seqs": w["n_seqs"],
            "ctx": w["ctx"], "block_size": BS, "n_head": w["n_head"], "n_head_kv": w["n_head_kv"], "head_dim": D,
            "parallelism": (f"sequence-batch sharding x{world}, no collective" if w["shard"] == "sequences"
                            else f"KV-head sharding x{world} + one exchange of the outputs per layer")}


def reference_arm(args, rank, world):
    """the reference's own CPU implementation of the path (oracle/_ref: ggml FLASH_ATTN_EXT on the flat cache with
    the dense F16 mask, SURVEY.md §8d "CPU baseline timing"), all host threads, one bounded sample per step"""
    if rank != 0:
        return
    t = TYPES[args.kv_type]
    sample, value, cores, kind, ms = cpu_layer_op(args.workload, t, args.steps, args.warmup)
    line = {"impl": "reference", "metric": "paged-attn decode KV-read GB/s", "value": value, "unit": "GB/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "ms_per_layer_op": ms,
            "step": "ONE layer-op of the workload (the b200 arm's step is the attention op of all its layers); both values are GB/s over the same algorithmic bytes",
            "higher_is_better": True,
            "scaling": WORKLOADS[args.workload]["scaling"], "vs_baseline": None, "dtype": args.kv_type, "data": "synthetic",
            "config": workload_config(args.workload, args.kv_type, world),
            "cpu_baseline": {"value": value, "unit": "GB/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=0, help="timed steps (0: the workload's default)")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="config2", choices=list(WORKLOADS))
    ap.add_argument("--kv-type", default=None, choices=list(TYPES))
    ap.add_argument("--kernel", type=int, default=0, help="0 auto, 1 generic, 2 stream-K")
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--plan-per-layer", action="store_true", help="A/B: every layer recomputes the schedule (no B200_ATTN_REUSE_PLAN)")
    ap.add_argument("--exchange", choices=["peer", "peer-flags", "nccl"], default="peer",
                    help="config4 at N>1: output exchange through the library's own NVLink peer-memory kernels (default) or NCCL all-gather")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs only: skip the end-to-end leg")
    ap.add_argument("--no-head-sharded", action="store_true", help="default workload only: skip the config-4 KV-head-sharded sub-record")
    args = ap.parse_args()
    W = WORKLOADS[args.workload]
    args.kv_type = args.kv_type or W["kv_type"]
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        args.steps = args.steps or 5
        return reference_arm(args, rank, world)
    if args.workload == "config3":
        return run_config3(args, rank, world, local)
    args.warmup = max(args.warmup, 3)

    import copy
    import torch
    import torch.distributed as dist
    from __graft_entry__ import load
    pkg = load()
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    a = copy.copy(args)
    a.steps = args.steps or W["steps"]
    line = run_decode(a, rank, world, local, pkg)
    # The driver's scaling run launches the DEFAULT workload at N = 1, 2, 4, 8: config 2, sequence-sharded, no collective.
    # The path's other partitioning (BASELINE config 4: 70B heads, 32 x 32k, KV heads sharded over the
    # ranks + one exchange of the outputs per layer) rides along as a sub-record of the same JSON line, so that record
    # carries its strong-scaling curve too (N = 1: the whole problem on one GPU, the curve's baseline).
    if args.workload == "config2" and not args.no_head_sharded and not args.layers and args.kv_type == W["kv_type"] and not args.no_e2e:
        del a
        import gc
        gc.collect(); torch.cuda.empty_cache()
        b = copy.copy(args)
        b.workload, b.kv_type, b.layers, b.kernel, b.plan_per_layer = "config4", "f16", 0, 0, False
        # long enough for the clock sampler: ~0.8 s of timed steps whatever the N (5 ms per step on one GPU, 0.87 ms on eight)
        b.steps = {1: 160, 2: 300, 4: 500}.get(world, 900)
        b.no_cpu_baseline = True
        sub = run_decode(b, rank, world, local, pkg)
        if rank == 0:
            keep = ("value", "unit", "ms_per_step", "scaling", "dtype", "decode_tok_s", "gpu_launches", "roofline", "clocks", "e2e",
                    "exchange", "steps", "warmup")
            line["kv_head_sharded"] = {"workload": sub["config"]["workload"], "parallelism": sub["config"]["parallelism"],
                                       "n_layer": sub["config"]["n_layer"], **{k: sub[k] for k in keep if k in sub}}
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_decode(args, rank, world, local, pkg):
    """one decode workload (args.workload) on this process group; returns the JSON line on rank 0"""
    import torch
    import torch.distributed as dist
    W = WORKLOADS[args.workload]
    t = TYPES[args.kv_type]
    n_layer = args.layers or W["n_layer"]
    n_seqs = W["n_seqs"]
    by_heads = W["shard"] == "heads"
    sh = pkg.head_shard(rank if by_heads else 0, world if by_heads else 1, W["n_head"], W["n_head_kv"])
    n_head, n_head_kv = sh.n_head, sh.n_head_kv              # what THIS rank computes
    collective = by_heads and world > 1
    lens = ctx_lens(args.workload, rank, 1 if by_heads else world)
    n_seqs = len(lens)
    bt_h, slot_h, n_blocks = block_tables(lens)
    total = int(lens.sum())
    lay = pkg.KVLayout(t, n_head_kv, D, BS, n_blocks + 8)
    kv_bytes_op = total * n_head_kv * 2 * row_bytes(t, D)    # algorithmic bytes of one layer-op on this rank
    scale = float(1.0 / np.sqrt(D))
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()

    # ---- resident state: per-layer KV caches filled through the product's own kv_write kernel
    gen = torch.Generator(device="cuda"); gen.manual_seed(1234 + rank)
    slot_d, bt_d, sl_d = dev(slot_h), dev(bt_h), dev(lens)
    kcs, vcs = [], []
    CH = 65536
    for _ in range(n_layer):
        kc, vc = lay.alloc(), lay.alloc()
        for o in range(0, total, CH):
            n = min(CH, total - o)
            ks = torch.rand((n, n_head_kv * D), generator=gen, device="cuda") * 2 - 1
            vs = torch.rand((n, n_head_kv * D), generator=gen, device="cuda") * 2 - 1
            pkg.kv_write(ks, vs, slot_d[o:o + n], kc, vc, lay)
        kcs.append(kc); vcs.append(vc)
    del ks, vs
    q_d = torch.rand((n_layer, n_seqs, n_head, D), generator=gen, device="cuda") * 2 - 1
    out_d = torch.empty_like(q_d)
    gathered = torch.empty((n_layer, world, n_seqs, n_head, D), device="cuda") if collective else None
    # the layers of a step share seq_lens / block table, hence the device-side schedule: one workspace, layer 0 plans,
    # layers 1.. reuse it (B200_ATTN_REUSE_PLAN; b200_decode_step works that out itself, separate calls say so)
    calls = []
    for l in range(n_layer):
        share = not args.plan_per_layer and l > 0
        calls.append(pkg.AttnCall(q_d[l], kcs[l], vcs[l], lay, out_d[l], scale, block_table=bt_d, seq_lens=sl_d,
                                  max_seq_len=int(lens.max()), total_tokens=total, kernel=args.kernel,
                                  workspace=calls[0].workspace if share else None, reuse_plan=share and collective))
    step_all = None if collective else pkg.StepCall(calls)
    torch.cuda.synchronize()

    peer = None
    if collective and args.exchange != "nccl":
        peer = pkg.PeerGather((n_seqs, n_head, D), n_slots=max(2, n_layer) + 1)
        # once, before anything is timed: the peer-memory exchange delivers exactly what NCCL's all-gather delivers
        calls[0].run(); peer.exchange_ll(out_d[0], max(2, n_layer)); peer.check()
        ref_g = pkg.gather_head_outputs(out_d[0])
        if not torch.equal(peer.gathered(max(2, n_layer)), ref_g):
            raise SystemExit("peer-memory exchange differs from NCCL all_gather")

    def layer_op(l, keep=False):
        calls[l].run()
        if peer is not None:
            (peer.exchange_ll if args.exchange == "peer" else peer.exchange)(out_d[l], l if n_layer > 1 else peer.epoch & 1)   # slot = layer: gathered(l) is that layer's exchanged output
        elif collective:
            pkg.gather_head_outputs(out_d[l], gathered[l])

    def step_resident():
        if step_all is not None:
            return step_all.run()                            # one b200_decode_step call over all layers
        for l in range(n_layer):
            layer_op(l)

    l0 = pkg.lib.b200_launch_count(); step_resident(); launches_per_step = pkg.lib.b200_launch_count() - l0
    graph = None
    if not args.no_graph and (not collective or peer is not None):
        e_before = peer.epoch if peer is not None else 0
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            step_resident()
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            step_resident()
        if peer is not None:
            # capture launched nothing: the host epoch must only count publishes that ran (the warm-up on `side` did)
            peer.epoch = e_before + n_layer

    def run_step():
        if graph is None:
            return step_resident()
        graph.replay()
        if peer is not None:
            peer.note_replay(n_layer)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    sampler = ClockSampler(local if rank == 0 else -1); sampler.start()   # rank 0's GPU: eight concurrent nvidia-smi loops starve each other
    ms_total = timed(run_step, args.steps, args.warmup)
    ms_step = ms_total / args.steps
    bytes_step_rank = kv_bytes_op * n_layer
    exchange_us = None
    if collective:
        # what the exchange adds per layer: the same step without it (this rank's attention only), same replay form
        def attn_only():
            for c in calls:
                c.run()
        g2 = None
        if graph is not None:
            side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                attn_only()
            torch.cuda.current_stream().wait_stream(side)
            g2 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g2):
                attn_only()
        ms_attn = timed(g2.replay if g2 is not None else attn_only, args.steps, args.warmup) / args.steps
        exchange_us = round((ms_step - ms_attn) * 1e3 / n_layer, 2)
    # whole-job bytes = sum over ranks (weak: every rank holds its own sequences; strong: its own kv heads)
    tot = torch.tensor([float(bytes_step_rank), float(n_seqs)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tot)
    value = tot[0].item() / (ms_step * 1e-3) / 1e9
    tok_s = (tot[1].item() if not by_heads else float(n_seqs)) / (ms_step * 1e-3)

    # ---- end to end through the C ABI with HOST buffers: per step, H2D of the step's inputs (copy stream, layer by
    # layer, overlapped with compute), kv_write + attention (+ all-gather) per layer, D2H of every layer's output
    e2e = None
    if not args.no_e2e:
        # head-sharded: the exchanged result [n_layer, world, ...] is on every GPU; each rank brings ITS slice to its host
        # process — the union over the ranks is the whole result and every byte crosses a host link once (eight ranks each
        # downloading all of it would move it eight times through one host's memory)
        res_d = ((peer.all_slots()[:n_layer] if peer is not None else gathered)[:, rank:rank + 1]) if collective else out_d
        q_h = (torch.rand((n_layer, n_seqs, n_head, D)) * 2 - 1).pin_memory()
        knew_h = (torch.rand((n_layer, n_seqs, n_head_kv * D)) * 2 - 1).pin_memory()
        vnew_h = (torch.rand((n_layer, n_seqs, n_head_kv * D)) * 2 - 1).pin_memory()
        bt_hp, sl_hp = torch.from_numpy(bt_h).pin_memory(), torch.from_numpy(lens).pin_memory()
        # the step's new token of each sequence overwrites its last cached position (cache stays the same size)
        last = np.cumsum(lens.astype(np.int64)) - 1
        slot_new_h = torch.from_numpy(slot_h[last].copy()).pin_memory()
        out_h = torch.empty(tuple(res_d.shape)).pin_memory()
        knew_d, vnew_d = torch.empty_like(knew_h, device="cuda"), torch.empty_like(vnew_h, device="cuda")
        slot_new_d = torch.empty_like(slot_new_h, device="cuda")
        h2d = 4 * (q_h.numel() + knew_h.numel() + vnew_h.numel() + bt_hp.numel() + sl_hp.numel() + slot_new_h.numel())
        d2h = 4 * out_h.numel()
        s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
        # layers go through the C ABI in groups: one b200_decode_step call (kv_write + attention per layer) per group,
        # its inputs copied H2D on the copy stream while the previous group computes, its outputs D2H while the next does
        # group sizes ramp up and down (1,1,2,4,8,...,8,4,2,1,1): only the first group's H2D and the last group's D2H
        # are exposed, everything else overlaps with compute
        if collective:
            # the step's inputs are small here (Q and one new K/V row per sequence: ~1 MB): the compute of each half of the
            # layers (kv_write + attention + exchange per layer) is ONE captured graph, the first half's outputs go D2H
            # while the second half computes
            sizes = [n_layer - n_layer // 2, n_layer // 2] if n_layer >= 2 and peer is not None else [1]*n_layer
        else:
            ramp, sizes, left = [1, 1, 2, 4], [], n_layer
            big = bytes_step_rank > 2e9                      # small steps are host-call bound: keep few groups
            head = [r for r in ramp if (left := left - r) >= 0] if big and n_layer >= 2*sum(ramp) + 8 else []
            mid = n_layer - 2*sum(head)
            sizes = head + [8]*(mid // 8) + ([mid % 8] if mid % 8 else []) + head[::-1]
        starts = np.cumsum([0] + sizes[:-1]).tolist()
        groups = [(a0, a0 + n) for a0, n in zip(starts, sizes)]
        assert groups[-1][1] == n_layer
        ev_in = [torch.cuda.Event() for _ in groups]
        ev_c = [torch.cuda.Event() for _ in groups]
        group_calls = None
        if not collective:
            group_calls = [pkg.StepCall(calls[a0:a1], [pkg.kv_write_desc(knew_d[l], vnew_d[l], slot_new_d, kcs[l], vcs[l], lay)
                                                        for l in range(a0, a1)]) for a0, a1 in groups]
        group_graphs = None
        if not collective and not args.no_graph:
            # each group's b200_decode_step call (kv_write + plan/main/combine per layer) captured once and replayed per step,
            # like the resident path: the host-side launch gaps between ~100 kernels per step go away, the copies stay
            # ordinary stream operations around the replays
            group_graphs = []
            for gc in group_calls:
                side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    gc.run()
                torch.cuda.current_stream().wait_stream(side)
                torch.cuda.synchronize()
                gg = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gg):
                    gc.run()
                group_graphs.append(gg)
        if collective and peer is not None and not args.no_graph and len(groups) == 2:
            def group_body(a0, a1):
                for l in range(a0, a1):
                    pkg.kv_write(knew_d[l], vnew_d[l], slot_new_d, kcs[l], vcs[l], lay)
                    layer_op(l, keep=True)
            group_graphs = []
            for a0, a1 in groups:
                e_before = peer.epoch
                side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    group_body(a0, a1)                      # warm-up outside capture (really runs: the epoch advances)
                torch.cuda.current_stream().wait_stream(side)
                torch.cuda.synchronize()
                if world > 1:
                    dist.barrier()
                gg = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gg):
                    group_body(a0, a1)
                peer.epoch = e_before + (a1 - a0)           # capture launched nothing
                group_graphs.append(gg)

        # One sequence-sharded step as ONE CUDA graph: the H2D copies (copy stream), every group's b200_decode_step call and
        # the D2H copies (second copy stream) captured with their cross-stream dependencies, so a step is one graph launch —
        # the ten group replays of the form below each ended in a full drain of the GPU, and the programmatic launch chain
        # of the kernels broke at every group boundary.  The copies are still the step's own, from / to pinned host memory.
        step_graph = None
        # (small steps — config 1, config 5: ~0.5 ms — measured 2 % slower this way and keep the per-group form)
        if group_calls is not None and group_graphs is not None and bytes_step_rank > 2e9 and os.environ.get("B200_E2E_STEP_GRAPH", "1") != "0":
            def step_body():
                cur = torch.cuda.current_stream()
                s_in.wait_stream(cur); s_out.wait_stream(cur)
                with torch.cuda.stream(s_in):
                    bt_d.copy_(bt_hp, non_blocking=True); sl_d.copy_(sl_hp, non_blocking=True)
                    slot_new_d.copy_(slot_new_h, non_blocking=True)
                    for gi, (a0, a1) in enumerate(groups):
                        q_d[a0:a1].copy_(q_h[a0:a1], non_blocking=True)
                        knew_d[a0:a1].copy_(knew_h[a0:a1], non_blocking=True); vnew_d[a0:a1].copy_(vnew_h[a0:a1], non_blocking=True)
                        ev_in[gi].record(s_in)
                for gi, (a0, a1) in enumerate(groups):
                    cur.wait_event(ev_in[gi])
                    group_calls[gi].run()
                    ev_c[gi].record(cur)
                    s_out.wait_event(ev_c[gi])
                    with torch.cuda.stream(s_out):
                        out_h[a0:a1].copy_(res_d[a0:a1], non_blocking=True)
                cur.wait_stream(s_in); cur.wait_stream(s_out)
            try:
                torch.cuda.synchronize()
                gg = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gg):
                    step_body()
                step_graph = gg
            except Exception as e:                           # keep the per-group form
                print(f"bench: whole-step graph capture failed ({type(e).__name__}: {e}); per-group graphs stay", file=sys.stderr)
                torch.cuda.synchronize()
                step_graph = None

        def step_e2e():
            if step_graph is not None:
                step_graph.replay()
                torch.cuda.current_stream().synchronize()    # the host holds the step's result
                return
            cur = torch.cuda.current_stream()
            s_in.wait_stream(cur)
            with torch.cuda.stream(s_in):
                bt_d.copy_(bt_hp, non_blocking=True); sl_d.copy_(sl_hp, non_blocking=True)
                slot_new_d.copy_(slot_new_h, non_blocking=True)
                for gi, (a0, a1) in enumerate(groups):
                    q_d[a0:a1].copy_(q_h[a0:a1], non_blocking=True)
                    knew_d[a0:a1].copy_(knew_h[a0:a1], non_blocking=True); vnew_d[a0:a1].copy_(vnew_h[a0:a1], non_blocking=True)
                    ev_in[gi].record(s_in)
            for gi, (a0, a1) in enumerate(groups):
                cur.wait_event(ev_in[gi])
                if group_graphs is not None and group_calls is not None:
                    group_graphs[gi].replay()
                elif group_calls is not None:
                    group_calls[gi].run()
                elif group_graphs is not None:
                    group_graphs[gi].replay(); peer.note_replay(a1 - a0)
                else:
                    for l in range(a0, a1):
                        pkg.kv_write(knew_d[l], vnew_d[l], slot_new_d, kcs[l], vcs[l], lay)
                        layer_op(l, keep=True)
                ev_c[gi].record(cur)
                s_out.wait_event(ev_c[gi])
                with torch.cuda.stream(s_out):
                    out_h[a0:a1].copy_(res_d[a0:a1], non_blocking=True)
            s_out.synchronize()                              # the host holds the step's result

        def step_copies():                                   # the step's host<->device traffic alone, same streams and sizes
            cur = torch.cuda.current_stream()
            s_in.wait_stream(cur)
            with torch.cuda.stream(s_in):
                bt_d.copy_(bt_hp, non_blocking=True); sl_d.copy_(sl_hp, non_blocking=True)
                slot_new_d.copy_(slot_new_h, non_blocking=True)
                q_d.copy_(q_h, non_blocking=True); knew_d.copy_(knew_h, non_blocking=True); vnew_d.copy_(vnew_h, non_blocking=True)
            s_out.wait_stream(cur)
            with torch.cuda.stream(s_out):
                out_h.copy_(res_d, non_blocking=True)
            s_in.synchronize(); s_out.synchronize()

        e2e_steps = max(3, args.steps // 2)
        ms_e2e = timed(step_e2e, e2e_steps, 3) / e2e_steps
        ms_copies = timed(step_copies, e2e_steps, 3) / e2e_steps
        e2e = {"value": round(tot[0].item() / (ms_e2e * 1e-3) / 1e9, 2), "unit": "GB/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": round(ms_e2e, 4),
               # where an e2e step's time can go besides the kernels: both directions of the step's copies run alone (max over
               # ranks; H2D and D2H concurrently, as in the step) and the GB/s that is per rank
               "form": ("one CUDA graph per step: H2D copies, b200_decode_step per layer group, D2H copies" if step_graph is not None
                        else "copies on two copy streams around one b200_decode_step call (captured graph) per layer group"),
               "copies_only_ms_per_step": round(ms_copies, 4),
               "copies_only_gbs_per_rank": round((h2d + d2h) / (ms_copies * 1e-3) / 1e9, 1)}
    clocks = sampler.summary()

    line = None
    if rank == 0:
        peak, peak_src = peaks()
        achieved = kv_bytes_op / (ms_step / n_layer * 1e-3) / 1e9         # per GPU, per layer-op
        which = "stream-K" if (args.kernel != 1 and _uses_streamk(pkg, calls[0])) else "generic"
        cfg = workload_config(args.workload, args.kv_type, world)
        cfg.update({"n_layer": n_layer, "n_head_per_gpu": n_head, "n_head_kv_per_gpu": n_head_kv,
                    "kv_bytes_per_step_per_gpu": bytes_step_rank, "cuda_graph": graph is not None, "kernel": which,
                    "l2_flush": f"inputs (per-layer KV caches, {bytes_step_rank / 1e9:.1f} GB per step per GPU) are far larger than the 126 MB L2"})
        cfg["schedule"] = ("every layer plans" if args.plan_per_layer else
                           "layer 0 plans on the device, layers 1.. reuse it (B200_ATTN_REUSE_PLAN: same seq_lens within a step)")
        if collective:
            cfg["exchange"] = (("NVLink peer-memory push, payload + epoch in the same 8-byte stores, one launch (b200_peer_exchange_ll)" if args.exchange == "peer"
                                else "NVLink peer-memory push + epoch flags, one launch (b200_peer_exchange)") + ", verified equal to NCCL all_gather"
                               if peer is not None else "NCCL all_gather_into_tensor")
        line = {
            "metric": "paged-attn decode KV-read GB/s", "value": round(value, 2), "unit": "GB/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 4), "higher_is_better": True,
            "scaling": W["scaling"], "vs_baseline": None, "dtype": args.kv_type, "data": "synthetic", "config": cfg,
            "decode_tok_s": round(tok_s, 1), "pct_hbm_nominal_8TBs": round(100 * achieved / 8000.0, 2),
            "gpu_launches": int(launches_per_step * args.steps),
            "roofline": {"bound": "hbm", "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                         "traffic": ncu_traffic(args.kv_type) if args.workload == "config2" else None, "peak_source": peak_src,
                         "kernel": f"paged attention op ({which}: plan + main + combine), one layer",
                         "algorithmic_bytes_per_launch": kv_bytes_op},
            "clocks": clocks,
        }
        if collective:
            line["exchange"] = {"us_per_layer": exchange_us, "form": cfg["exchange"], "equals_nccl": peer is not None,
                                "what": "step time with the exchange minus the same ranks' attention-only step, per layer"}
        if e2e is not None:
            line["e2e"] = e2e
        if not args.no_cpu_baseline:
            sample, v, cores, kind, _ = cpu_layer_op(args.workload, t, 4, 1)
            line["cpu_baseline"] = {"value": round(v, 3), "unit": "GB/s", "cores": cores, "kind": kind, "sample": sample}
    if peer is not None:
        peer.check()
        if world > 1:
            dist.barrier()
        peer.close()
    return line if rank == 0 else None


def run_config3(args, rank, world, local):
    """chunked prefill (BASELINE config 3): per layer, slot-mapped write of the 8 x 512 new rows then causal tile
    attention of those 4096 query rows over their sequences' 4096-token paged contexts"""
    import torch
    import torch.distributed as dist
    from __graft_entry__ import load
    pkg = load()
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W = WORKLOADS["config3"]
    t = TYPES[args.kv_type]
    n_layer, n_seqs, n_head, n_head_kv = args.layers or W["n_layer"], W["n_seqs"], W["n_head"], W["n_head_kv"]
    steps = args.steps or W["steps"]
    CHUNK, P0 = 512, 3584
    lens = ctx_lens("config3", rank)
    bt_h, slot_h, n_blocks = block_tables(lens)
    total = int(lens.sum())
    lay = pkg.KVLayout(t, n_head_kv, D, BS, n_blocks + 8)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    gen = torch.Generator(device="cuda"); gen.manual_seed(4321 + rank)
    slot_d, bt_d, sl_d = dev(slot_h), dev(bt_h), dev(lens)
    starts = np.concatenate([[0], np.cumsum(lens.astype(np.int64))[:-1]])
    rows_seq = np.repeat(np.arange(n_seqs, dtype=np.int32), CHUNK)
    rows_pos = np.tile(np.arange(P0, P0 + CHUNK, dtype=np.int32), n_seqs)
    slot_chunk = dev(slot_h[(starts[rows_seq] + rows_pos).astype(np.int64)])
    qs_d, qp_d = dev(rows_seq), dev(rows_pos)
    n_q = n_seqs * CHUNK
    kcs, vcs = [], []
    for _ in range(n_layer):
        kc, vc = lay.alloc(), lay.alloc()
        for o in range(0, total, 16384):
            n = min(16384, total - o)
            ks = torch.rand((n, n_head_kv * D), generator=gen, device="cuda") * 2 - 1
            vs = torch.rand((n, n_head_kv * D), generator=gen, device="cuda") * 2 - 1
            pkg.kv_write(ks, vs, slot_d[o:o + n], kc, vc, lay)
        kcs.append(kc); vcs.append(vc)
    knew = torch.rand((n_q, n_head_kv * D), generator=gen, device="cuda") * 2 - 1
    vnew = torch.rand((n_q, n_head_kv * D), generator=gen, device="cuda") * 2 - 1
    q_d = torch.rand((n_q, n_head, D), generator=gen, device="cuda") * 2 - 1        # one Q / out buffer shared by the layers
    out_d = torch.empty_like(q_d)
    scale = float(1.0 / np.sqrt(D))
    calls = [pkg.AttnCall(q_d, kcs[l], vcs[l], lay, out_d, scale, block_table=bt_d, seq_lens=sl_d, q_seq=qs_d, q_pos=qp_d,
                          max_seq_len=int(lens.max()), total_tokens=total, kernel=args.kernel) for l in range(n_layer)]
    torch.cuda.synchronize()

    def writes():
        for l in range(n_layer):
            pkg.kv_write(knew, vnew, slot_chunk, kcs[l], vcs[l], lay)

    def attns():
        for c in calls:
            c.run()

    def step():
        for l in range(n_layer):
            pkg.kv_write(knew, vnew, slot_chunk, kcs[l], vcs[l], lay)
            calls[l].run()

    def graphed(fn):
        if args.no_graph:
            return fn
        side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        return g.replay

    l0 = pkg.lib.b200_launch_count(); step(); launches = pkg.lib.b200_launch_count() - l0

    def timed(fn, n):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1) / n], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    sampler = ClockSampler(local); sampler.start()
    ms_step = timed(graphed(step), steps)
    ms_attn = timed(graphed(attns), steps)
    ms_write = timed(graphed(writes), steps)
    clocks = sampler.summary()
    vis = float(sum(min(int(lens[s]), P0 + i + 1) for s in range(n_seqs) for i in range(CHUNK)))
    flops_op = 4.0 * n_head * D * vis                                   # QK^T + PV over the visible (causal) cells
    write_bytes_op = n_q * n_head_kv * D * (4 + row_bytes(t, D) / D) * 2   # F32 rows in, typed rows out, K and V
    if rank == 0:
        tf_attn = flops_op * n_layer / (ms_attn * 1e-3) / 1e12
        tf_step = flops_op * n_layer * world / (ms_step * 1e-3) / 1e12
        peak = 1640.5
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peak = float(json.load(f)["bf16_tflops"])
        except Exception:
            pass
        cfg = workload_config("config3", args.kv_type, world)
        cfg.update({"n_layer": n_layer, "chunk": CHUNK, "chunk_start": P0, "query_rows_per_layer_op": n_q, "cuda_graph": not args.no_graph,
                    "kernel": "tcgen05 tile kernel (plan + main)" if args.kernel in (0, 3) else f"kernel {args.kernel}",
                    "l2_flush": f"per-layer KV caches ({2 * lay.total_bytes * n_layer / 1e9:.1f} GB per step) exceed the 126 MB L2"})
        line = {"metric": "chunked-prefill paged attention TFLOP/s (KV write + attention per layer)", "value": round(tf_step, 2),
                "unit": "TFLOP/s", "n_gpus": world, "steps": steps, "warmup": 3, "ms_per_step": round(ms_step, 4),
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16 (fp32 accumulate)", "data": "synthetic",
                "config": cfg, "prefill_tok_s": round(n_q * world / (ms_step * 1e-3) / n_layer, 1),
                "attention_only": {"ms_per_step": round(ms_attn, 4), "tflops_per_gpu": round(tf_attn, 2)},
                "kv_write_only": {"ms_per_step": round(ms_write, 4), "gbs_per_gpu": round(write_bytes_op * n_layer / (ms_write * 1e-3) / 1e9, 1),
                                  "algorithmic_bytes_per_launch": int(write_bytes_op)},
                "gpu_launches": int(launches * steps),
                "roofline": {"bound": "tensor", "achieved": round(tf_attn, 2), "peak": peak, "unit": "TFLOP/s", "frac": round(tf_attn / peak, 4),
                             "traffic": None, "peak_source": "measured (MEASURED_PEAKS.json bf16_tflops, burst)",
                             "kernel": "pf_main_kernel, one layer", "algorithmic_flops_per_launch": flops_op},
                "clocks": clocks}
        print(json.dumps(line))
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


def _uses_streamk(pkg, call):
    import ctypes as C
    d = pkg.AttnDesc.from_buffer_copy(call.desc)
    d.kernel = pkg.KERNEL_STREAMK
    return bool(pkg.lib.b200_supports_attention(C.byref(d)))


if __name__ == "__main__":
    main()
