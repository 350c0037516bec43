// oracle/llama_logits_dump.cpp — a few decode steps through libllama's public API (include/llama.h) with the logits of
// every step written to a file, so that the three ways of running the attention path (reference kernels; libb200attn on
// the unmodified dense-mask graphs; libllama with the paged KV section, llama.cpp_test_b200/host/libllama_paged.patch)
// can be compared on the SAME model, tokens and call sequence.  TEST INFRASTRUCTURE ONLY; no arithmetic of its own.
//
// Call sequence (what llama-batched-bench and llama-server do): one batch with the prompts of 5 sequences of unequal
// length, interleaved position by position like examples/batched-bench/batched-bench.cpp:118-124; llama_kv_self_seq_cp of
// sequence 0 into sequence 5 (a fork sharing a partly filled block -> copy-on-write in the paged form); 4 decode steps of
// one token per sequence; llama_kv_self_seq_rm of sequence 2 (whole) and of the tail of sequence 3; 2 more steps.
//   llama_logits_dump MODEL.gguf OUT.bin [-ctk q4_0|q8_0|f16]
#include "llama.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

int main(int argc, char ** argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s MODEL.gguf OUT.bin [-ctk type]\n", argv[0]); return 2; }
    ggml_type kvt = GGML_TYPE_Q4_0;
    for (int i = 3; i + 1 < argc; i += 2) {
        if (!strcmp(argv[i], "-ctk")) kvt = !strcmp(argv[i + 1], "f16") ? GGML_TYPE_F16 : !strcmp(argv[i + 1], "q8_0") ? GGML_TYPE_Q8_0 : GGML_TYPE_Q4_0;
    }
    llama_backend_init();
    llama_model_params mp = llama_model_default_params();
    mp.n_gpu_layers = 99;
    llama_model * model = llama_model_load_from_file(argv[1], mp);
    if (!model) return 1;
    llama_context_params cp = llama_context_default_params();
    cp.n_ctx = 2048; cp.n_batch = 512; cp.n_ubatch = 512; cp.n_seq_max = 8;
    cp.flash_attn = true; cp.type_k = kvt; cp.type_v = kvt;
    llama_context * ctx = llama_init_from_model(model, cp);
    if (!ctx) return 1;
    const int n_vocab = llama_vocab_n_tokens(llama_model_get_vocab(model));
    FILE * out = fopen(argv[2], "wb");
    if (!out) return 1;

    const int lens[5] = {37, 5, 64, 100, 16};
    std::vector<int> n_past(8, 0);
    unsigned rs = 12345;
    auto tok = [&]() { rs = rs*1664525u + 1013904223u; return (llama_token) ((rs >> 8) % (unsigned) n_vocab); };
    llama_batch batch = llama_batch_init(512, 0, 1);
    auto add = [&](llama_token t, int pos, int seq, bool logits) {
        const int i = batch.n_tokens++;
        batch.token[i] = t; batch.pos[i] = pos; batch.n_seq_id[i] = 1; batch.seq_id[i][0] = seq; batch.logits[i] = logits;
    };
    auto dump = [&]() {                                       // logits of every row that asked for them, in batch order
        for (int i = 0; i < batch.n_tokens; ++i) {
            if (!batch.logits[i]) continue;
            const float * l = llama_get_logits_ith(ctx, i);
            fwrite(l, sizeof(float), (size_t) n_vocab, out);
        }
    };
    // prompts, interleaved position by position
    batch.n_tokens = 0;
    for (int p = 0; p < 100; ++p) for (int s = 0; s < 5; ++s) if (p < lens[s]) add(tok(), p, s, p == lens[s] - 1);
    if (llama_decode(ctx, batch) != 0) { fprintf(stderr, "prompt decode failed\n"); return 1; }
    dump();
    for (int s = 0; s < 5; ++s) n_past[s] = lens[s];
    llama_kv_self_seq_cp(ctx, 0, 5, -1, -1);
    n_past[5] = n_past[0];
    std::vector<int> live = {0, 1, 2, 3, 4, 5};
    for (int step = 0; step < 6; ++step) {
        if (step == 4) {
            llama_kv_self_seq_rm(ctx, 2, -1, -1);             // sequence 2 leaves
            llama_kv_self_seq_rm(ctx, 3, 90, -1);             // sequence 3 is cut back to 90 tokens
            n_past[3] = 90;
            live = {0, 1, 3, 4, 5};
        }
        batch.n_tokens = 0;
        for (int s : live) add(tok(), n_past[s]++, s, true);
        if (llama_decode(ctx, batch) != 0) { fprintf(stderr, "decode step %d failed\n", step); return 1; }
        dump();
    }
    fclose(out);
    llama_batch_free(batch);
    llama_free(ctx);
    llama_model_free(model);
    return 0;
}
