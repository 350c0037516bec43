#!/usr/bin/env python
"""tests/golden/make_golden.py — writes tests/golden/ref_vectors.npz from the REFERENCE itself.

The reference's own tests hold no golden vectors for this path (tests/test-backend-ops.cpp seeds its inputs from
std::random_device, :49-52), so the fixtures are outputs of the reference's ggml CPU code compiled where it lies
under /root/reference (oracle/_ref, see oracle/Makefile) and run in the build container:

  * quantize_row_q8_0_ref / quantize_row_q4_0_ref / ggml_fp32_to_fp16_row   (cache bytes of seeded rows)
  * GGML_OP_CPY on the CPU backend                                           (its SIMD quantiser)
  * GGML_OP_FLASH_ATTN_EXT (O2) and the F32 mul_mat/soft_max_ext formulation (O1) on the dense-mask equivalent
    of seeded paged workloads (tests/workloads.py regenerates the inputs from the seed; only outputs are stored)

Run here after `make -C oracle ref`:   python tests/golden/make_golden.py
tests/test_golden.py checks the oracle (CPU) and the CUDA kernels (-m gpu) against the file on machines that
have no reference tree."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.loader import Oracle, Ref          # noqa: E402
from workloads import F16, Q4_0, Q8_0, make_workload   # noqa: E402

# (name, seq_lens, n_head, n_head_kv, d, type_k, type_v, seed)
ATTN_CASES = [
    ("f16_ragged",  [1, 15, 16, 17, 100, 333], 32, 8, 128, F16,  F16,  101),
    ("q8_ragged",   [1, 15, 16, 17, 100, 333], 32, 8, 128, Q8_0, Q8_0, 102),
    ("q4_ragged",   [1, 15, 16, 17, 100, 333], 32, 8, 128, Q4_0, Q4_0, 103),
    ("q8k_f16v",    [200, 31],                 32, 8, 128, Q8_0, F16,  104),
    ("f16_gqa8",    [260, 48, 129],            64, 8, 128, F16,  F16,  105),
    ("q4_gqa8",     [260, 48, 129],            64, 8, 128, Q4_0, Q4_0, 106),
    ("q8_d64",      [77, 300],                 12, 4, 64,  Q8_0, Q8_0, 107),
]


def quant_rows(seed=2024, n=64, k=128):
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    x[0, :32] = 0.0
    x[1, :32] = np.float32(1e-30)
    x[2, 7] = np.float32(6.0e4)
    x[3] = rng.standard_normal(k).astype(np.float32) * 100
    x[4, :32] = np.float32(-0.75)
    x[5, :32] = np.linspace(-1, 1, 32, dtype=np.float32); x[5, 31] = -1.0
    return x


def main():
    orc, ref = Oracle(), Ref.load()
    assert ref is not None, "build oracle/_ref first (make -C oracle ref)"
    out = {}
    x = quant_rows()
    for t, name in ((F16, "f16"), (Q8_0, "q8_0"), (Q4_0, "q4_0")):
        out[f"quant_ref_{name}"] = ref.quantize(t, x)
        out[f"cpy_cpu_{name}"] = ref.cpy_f32(t, x, n_threads=1)
        out[f"dequant_{name}"] = ref.to_float(t, out[f"quant_ref_{name}"], 128)
    for name, lens, nh, nkv, d, tk, tv, seed in ATTN_CASES:
        w = make_workload(orc, lens, nh, nkv, d, tk, tv, seed=seed)
        mask = w.dense_mask()
        f32, _ = ref.attention(1, w.q, w.k_cache, w.v_cache, tk, tv, nkv, w.n_cells, mask, w.scale, n_threads=2)
        fa, _ = ref.attention(0, w.q, w.k_cache, w.v_cache, tk, tv, nkv, w.n_cells, mask, w.scale, n_threads=2)
        out[f"attn_f32_{name}"] = f32
        out[f"attn_fa_{name}"] = fa.astype(np.float16)          # the op's own NMSE bar is 5e-4: half storage is ample
        # a checksum of the regenerated inputs so a drift in workloads.py / numpy is caught, not silently compared
        out[f"attn_in_{name}"] = np.array([w.q.sum(dtype=np.float64), w.k_cache.astype(np.uint64).sum(), w.v_cache.astype(np.uint64).sum()])
    path = os.path.join(HERE, "ref_vectors.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {os.path.getsize(path)/1024:.0f} KiB, {len(out)} arrays, reference isa {ref.isa}")


if __name__ == "__main__":
    main()
