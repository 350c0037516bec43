#!/usr/bin/env python
"""profiles/bench_kshift.py — K-shift / context-shift measurement on one B200: Llama-3-8B shape, 32 layers,
one sequence of 8192 tokens drops n_discard tokens after n_keep (the reference's context shift, examples/main), per KV type.
Prints one JSON line per case: the K-shift kernel alone (CUDA events, all layers in one launch; algorithmic bytes =
2 x row_bytes per shifted cell per layer, read + write) and the whole context_shift (host block-table edit + every launch).
The CPU arm is the reference's own K-shift graph (oracle/_ref) on a 1-layer slice of the same cache, all host threads."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g          # noqa: E402
from oracle.loader import Oracle, Ref  # noqa: E402

F16, Q4_0, Q8_0 = 1, 2, 8


def main():
    pkg = g.load()
    torch.cuda.set_device(0)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak = float(peaks.get("hbm_gbs", 6459.0))
    L, hkv, d, bs, n_ctx = 32, 8, 128, 16, 8192
    n_blocks = n_ctx // bs + 8
    orc, ref = Oracle(), Ref.load()
    for t, name in ((F16, "f16"), (Q8_0, "q8_0"), (Q4_0, "q4_0")):
        lay = pkg.KVLayout(t, hkv, d, bs, n_blocks)
        rng = np.random.default_rng(1)
        host = orc.from_float(t, rng.uniform(-1, 1, (lay.n_cells, hkv, d)).astype(np.float32)).reshape(-1)
        kd = [torch.from_numpy(host).cuda() for _ in range(L)]
        vd = [torch.from_numpy(host).cuda() for _ in range(L)]
        for n_keep, n_discard, tag in ((256, 4096 - 128, "block-aligned"), (250, 3961, "ragged")):
            times, ks_times = [], []
            n_shift = 0
            for it in range(6):
                pc = pkg.PagedKVCache(n_blocks, bs)
                pc.append(0, n_ctx)
                plan_cells = n_ctx - n_keep - n_discard
                # kernel alone: the cells that a shift would rotate
                cells = torch.arange(n_keep, n_keep + plan_cells, dtype=torch.int32, device="cuda")
                shift = torch.zeros(lay.n_cells, dtype=torch.int32, device="cuda"); shift[cells.long()] = -n_discard
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); pkg.k_shift(kd, lay, shift, cells); e1.record(); torch.cuda.synchronize()
                ks_times.append(e0.elapsed_time(e1))
                t0 = time.perf_counter()
                ok = pkg.context_shift(pc, 0, n_keep, n_keep + n_discard, kd, vd, lay)
                torch.cuda.synchronize()
                times.append((time.perf_counter() - t0)*1e3)
                assert ok and pc.seq_len(0) == n_ctx - n_discard
                n_shift = plan_cells
            ks_ms, e2e_ms = min(ks_times[1:]), min(times[1:])
            bytes_ = 2*lay.token_bytes*n_shift*L
            line = {"metric": "K-shift GB/s (read + write of the rotated K rows)", "value": round(bytes_/ks_ms/1e6, 1), "unit": "GB/s",
                    "dtype": name, "config": {"workload": f"context shift {tag}: 1 seq x {n_ctx}, keep {n_keep}, discard {n_discard}, {L} layers, 8 kv heads x 128"},
                    "ms_k_shift_all_layers": round(ks_ms, 4), "ms_context_shift_end_to_end": round(e2e_ms, 3), "rows_rotated_per_layer": n_shift,
                    "roofline": {"bound": "hbm", "achieved": round(bytes_/ks_ms/1e6, 1), "peak": peak, "unit": "GB/s", "frac": round(bytes_/ks_ms/1e6/peak, 4)}}
            if ref is not None and tag == "block-aligned":
                c = host.copy(); sh = np.zeros(lay.n_cells, np.int32); sh[n_keep:n_keep + n_shift] = -n_discard
                best = 1e9
                for _ in range(3):
                    t0 = time.perf_counter(); ref.k_shift(c, t, hkv, d, sh, n_threads=ref.n_threads); best = min(best, time.perf_counter() - t0)
                line["cpu_baseline"] = {"value": round(2*lay.token_bytes*lay.n_cells/best/1e9, 3), "unit": "GB/s", "cores": ref.n_threads, "kind": "reference",
                                        "sample": f"the reference's K-shift graph over ALL {lay.n_cells} cells of one layer (what it executes), ms {best*1e3:.2f}; x{L} layers per shift"}
            print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
