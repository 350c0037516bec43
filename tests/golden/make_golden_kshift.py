#!/usr/bin/env python
"""tests/golden/make_golden_kshift.py — writes tests/golden/k_shift_vectors.npz from the REFERENCE itself.

The reference holds no golden vectors for the K-shift (its rope test, tests/test-backend-ops.cpp test_rope, compares
backends on random inputs).  The fixtures are the cache bytes the reference's own K-shift graph
(llama_context::build_kv_self_shift, src/llama-context.cpp:524-571, rebuilt by oracle/ref_harness.cpp ref_k_shift and
executed by the reference's ggml CPU code compiled under oracle/_ref) leaves behind on seeded caches.

Run here after `make -C oracle ref`:   python tests/golden/make_golden_kshift.py
tests/test_k_shift.py checks the oracle (CPU) and the CUDA kernel (-m gpu) against the file."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.loader import Oracle, Ref          # noqa: E402

F16, Q4_0, Q8_0 = 1, 2, 8
N_CELLS, N_HEAD_KV, D = 48, 2, 128

# (name, rope keyword arguments) — Llama-3 defaults live in the loaders' signatures
ROPE_CASES = [
    ("llama3",   {}),
    ("neox",     dict(mode=2, freq_base=10000.0)),
    ("partial",  dict(n_rot=64, freq_base=10000.0)),
    ("yarn",     dict(freq_scale=0.25, ext_factor=1.0, attn_factor=1.2, n_ctx_orig=4096, freq_base=10000.0)),
    ("factors",  dict(freq_factors=np.linspace(1.0, 8.0, 64).astype(np.float32))),
]
TYPES = [(F16, "f16"), (Q8_0, "q8_0"), (Q4_0, "q4_0")]


def inputs(seed=4242):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((N_CELLS, N_HEAD_KV, D)).astype(np.float32)
    shift = rng.integers(-2048, 2048, N_CELLS).astype(np.int32)
    shift[::6] = 0                      # untouched cells
    shift[1] = -512; shift[2] = 1; shift[3] = 32768
    return x, shift


def main():
    orc, ref = Oracle(), Ref.load()
    assert ref is not None, "build oracle/_ref first (make -C oracle ref)"
    x, shift = inputs()
    out = {}
    for t, tn in TYPES:
        base = orc.from_float(t, x).reshape(-1)
        for name, kw in ROPE_CASES:
            c = base.copy()
            ref.k_shift(c, t, N_HEAD_KV, D, shift, **kw)
            out[f"{tn}_{name}"] = c
    path = os.path.join(HERE, "k_shift_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", ref.features())


if __name__ == "__main__":
    main()
