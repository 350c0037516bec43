#!/usr/bin/env python
"""profiles/make_random_gguf.py OUT.gguf [--layers 32] — a random-init Llama-3-8B in GGUF with every matrix in Q4_0
(BASELINE config 5: "llama-batched-bench end-to-end, random-init Llama-3-8B Q4_0 weights"; there is no network for a
checkpoint).  Shapes are Llama-3-8B's (n_embd 4096, n_ff 14336, 32 heads / 8 kv heads, vocab 128256, rope theta 5e5);
the tokenizer is "no_vocab" (llama-batched-bench feeds random token ids and never detokenises,
examples/batched-bench/batched-bench.cpp:60-75).  Q4_0 blocks are written directly: a fixed half scale and random
nibbles, i.e. weights uniform in scale*[-8, 7]; the values only have to be finite — the benchmark measures time.
Uses the `gguf` package of the image (the reference's own gguf-py is the same writer)."""
import argparse
import sys

import numpy as np
import gguf


def q4_0_random(rng, rows, cols, scale):
    nb = cols // 32
    blk = np.empty((rows, nb, 18), dtype=np.uint8)
    blk[:, :, :2] = np.frombuffer(np.float16(scale).tobytes(), dtype=np.uint8)
    blk[:, :, 2:] = rng.integers(0, 256, size=(rows, nb, 16), dtype=np.uint8)
    return blk.reshape(rows, nb * 18)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--vocab", type=int, default=128256)
    a = ap.parse_args()
    n_embd, n_ff, n_head, n_head_kv, n_layer = 4096, 14336, 32, 8, a.layers
    w = gguf.GGUFWriter(a.out, "llama")
    w.add_name("random-init Llama-3-8B shape, Q4_0")
    w.add_context_length(131072)
    w.add_embedding_length(n_embd)
    w.add_block_count(n_layer)
    w.add_feed_forward_length(n_ff)
    w.add_head_count(n_head)
    w.add_head_count_kv(n_head_kv)
    w.add_rope_dimension_count(n_embd // n_head)
    w.add_rope_freq_base(500000.0)
    w.add_layer_norm_rms_eps(1e-5)
    w.add_vocab_size(a.vocab)
    w.add_file_type(2)                                  # LLAMA_FTYPE_MOSTLY_Q4_0
    w.add_tokenizer_model("no_vocab")
    rng = np.random.default_rng(5)
    Q4 = gguf.GGMLQuantizationType.Q4_0
    s_in = 1.0 / np.sqrt(n_embd) / 4.6                  # rms of uniform{-8..7} is ~4.6
    s_ff = 1.0 / np.sqrt(n_ff) / 4.6

    def mat(name, rows, cols, scale):
        w.add_tensor(name, q4_0_random(rng, rows, cols, scale), raw_dtype=Q4)

    mat("token_embd.weight", a.vocab, n_embd, 0.02)
    for il in range(n_layer):
        p = f"blk.{il}."
        w.add_tensor(p + "attn_norm.weight", np.ones(n_embd, dtype=np.float32))
        mat(p + "attn_q.weight", n_embd, n_embd, s_in)
        mat(p + "attn_k.weight", n_embd // n_head * n_head_kv, n_embd, s_in)
        mat(p + "attn_v.weight", n_embd // n_head * n_head_kv, n_embd, s_in)
        mat(p + "attn_output.weight", n_embd, n_embd, s_in)
        w.add_tensor(p + "ffn_norm.weight", np.ones(n_embd, dtype=np.float32))
        mat(p + "ffn_gate.weight", n_ff, n_embd, s_in)
        mat(p + "ffn_up.weight", n_ff, n_embd, s_in)
        mat(p + "ffn_down.weight", n_embd, n_ff, s_ff)
    w.add_tensor("output_norm.weight", np.ones(n_embd, dtype=np.float32))
    mat("output.weight", a.vocab, n_embd, s_in)
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    print("wrote", a.out, file=sys.stderr)


if __name__ == "__main__":
    main()
