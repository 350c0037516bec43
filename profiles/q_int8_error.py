#!/usr/bin/env python
"""profiles/q_int8_error.py — would `K.Q^T` on int8 tensor-core MMAs (Q quantised to int8 like the reference's CPU op does for
quantised K, ggml-cpu.c:12300-12343, raw Q8_0 cache bytes as the other operand: no expansion instructions) stay inside
north_star's bar?  CPU experiment, numpy in float64 over the SAME dequantised Q8_0 cache: attention with Q rounded to half (what
sk_main_kernel does), to int8 per 32-element block (the reference's Q8_0 Q) and to int8 per head row, against exact Q; the
metric is the parity tests' (max |a-b| / max(|b|, rms(row))).  Result (profiles/r2_q_int8_error.txt): half 3e-4 .. 1e-3; int8
4.8e-3 on U(-1,1) inputs but 2.2e-2 .. 2.8e-2 on N(0,1) inputs — twice the 1e-2 bar.  The formulation was therefore not built."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.loader import Oracle  # noqa: E402

orc = Oracle()
rng = np.random.default_rng(0)


def run(dist, n_ctx, n_q=8, H=8, Hkv=2, d=128):
    draw = (lambda s: rng.standard_normal(s).astype(np.float32)) if dist == "normal" else (lambda s: rng.uniform(-1, 1, s).astype(np.float32))
    q, k, v = draw((n_q, H, d)), draw((n_ctx, Hkv, d)), draw((n_ctx, Hkv, d))
    kq = orc.to_float(8, orc.from_float(8, k.reshape(-1, d)), d).reshape(k.shape)
    vq = orc.to_float(8, orc.from_float(8, v.reshape(-1, d)), d).reshape(v.shape)
    scale = 1 / np.sqrt(d)

    def attn(qq):
        out = np.zeros((n_q, H, d), np.float64)
        for h in range(H):
            kh, vh = kq[:, h // (H // Hkv)].astype(np.float64), vq[:, h // (H // Hkv)].astype(np.float64)
            s = qq[:, h].astype(np.float64) @ kh.T * scale
            p = np.exp(s - s.max(1, keepdims=True)); p /= p.sum(1, keepdims=True)
            out[:, h] = p @ vh
        return out

    def q_int8(x, blk):
        xs = x.reshape(-1, blk); dq = np.abs(xs).max(1, keepdims=True) / 127; dq[dq == 0] = 1
        return (np.round(xs / dq) * dq).reshape(x.shape)

    ref = attn(q)
    rms = np.sqrt((ref ** 2).mean(-1, keepdims=True))
    res = {}
    for name, qq in (("half", (q * scale).astype(np.float16).astype(np.float32) / scale), ("int8 per 32", q_int8(q, 32)), ("int8 per row", q_int8(q, 128))):
        res[name] = float((np.abs(attn(qq) - ref) / np.maximum(np.abs(ref), rms)).max())
    return res


if __name__ == "__main__":
    for dist in ("uniform", "normal"):
        for n in (512, 4096):
            print(dist, n, {k: f"{v:.2e}" for k, v in run(dist, n).items()})
