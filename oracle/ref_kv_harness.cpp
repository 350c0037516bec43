// oracle/ref_kv_harness.cpp — TEST INFRASTRUCTURE ONLY.  Drives the REFERENCE's own llama_kv_cache_unified
// (src/llama-kv-cache.{h,cpp}, compiled where it lies together with the rest of libllama's sources by `make refkv`;
// nothing is copied) with a scripted call stream, so that the restatements the tests lean on — the cell-per-token
// model in tests/test_paged_cache.py and the session-stream writer in oracle/kv_state_oracle.py — are pinned to what
// the reference really does.  Only object construction is ours: a llama_model shell with hand-set hparams (no file is
// loaded) and the cache's own init(); every bookkeeping call below is the reference's.
//
// Protocol (stdin, one command per line; every command answers one line on stdout):
//   I n_layer n_embd_head n_head_kv kv_size type_k type_v      init                            -> "ok"
//   A seq n            ubatch of n tokens for seq at the positions that continue it, find_slot, the K row of layer 0 of each
//                      placed cell is stamped with a running token id, head += n (llama-context.cpp:1291-1353)
//                                                                                       -> "slot h" | "noslot"
//   B n s0 s1 ...      one ubatch of n tokens, token i for sequence s_i (interleaved), as above  -> "slot h" | "noslot"
//   M n s0 s1 ...      as B, then kv_self->n and the reference's mask fill for the ubatch: per token row, the token ids its
//                      mask row leaves visible, by position                         -> "m n_kv ids ids ..." | "noslot"
//   C src dst p0 p1 / R seq p0 p1 / K seq / X                  seq_cp / seq_rm / seq_keep / clear -> "ok" | "0"/"1" for R
//   T seq p0 p1 delta  seq_add                                                                 -> "ok"
//   E seq              has_shift and, per cell of seq with a pending K rotation, pos:delta      -> "e 0|1 pos:delta ..."
//   P seq              seq_pos_max                                                             -> number
//   N                  get_n_tokens get_used_cells                                             -> "n used"
//   V seq              token ids seq sees, in position order (cells whose set holds seq)       -> "v id id ..."
//   S seq              state_write(io, seq) as hex (seq -1: the whole cache)                   -> "s hex"
//   L seq hex          state_read(io, seq) of that stream                                      -> "ok" | "fail"
#include "llama-kv-cache.h"
#include "llama-model.h"
#include "llama-cparams.h"
#include "llama-batch.h"
#include "llama-io.h"
#include "llama-graph.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

struct mem_writer : llama_io_write_i {
    std::vector<uint8_t> buf;
    void write(const void * src, size_t size) override { const uint8_t * p = (const uint8_t *) src; buf.insert(buf.end(), p, p + size); }
    void write_tensor(const ggml_tensor * t, size_t offset, size_t size) override { write((const char *) t->data + offset, size); }
    size_t n_bytes() override { return buf.size(); }
};
struct mem_reader : llama_io_read_i {
    std::vector<uint8_t> buf; size_t at = 0;
    const uint8_t * read(size_t size) override { if (at + size > buf.size()) throw std::runtime_error("unexpectedly reached end of buffer"); const uint8_t * p = buf.data() + at; at += size; return p; }
    void read_to(void * dst, size_t size) override { memcpy(dst, read(size), size); }
    size_t n_bytes() override { return at; }
};

int main() {
    llama_log_set([](ggml_log_level, const char *, void *) {}, nullptr);
    llama_model model(llama_model_default_params());
    model.arch = LLM_ARCH_LLAMA;
    llama_cparams cparams = {};
    cparams.flash_attn = true;
    cparams.causal_attn = true;                                   // v_trans = 0, like -fa runs (the paged path's layout)
    std::unique_ptr<llama_kv_cache_unified> kv;
    int32_t next_tok = 0;
    size_t k_row = 0;
    std::string line;
    while (std::getline(std::cin, line)) {
        std::istringstream in(line);
        char op = 0; in >> op;
        if (op == 'I') {
            int n_layer, hd, nkv, size, tk, tv; in >> n_layer >> hd >> nkv >> size >> tk >> tv;
            model.hparams.n_layer = (uint32_t) n_layer;
            model.hparams.n_embd_head_k = model.hparams.n_embd_head_v = (uint32_t) hd;
            std::fill(model.hparams.n_head_arr.begin(), model.hparams.n_head_arr.end(), (uint32_t) nkv);
            std::fill(model.hparams.n_head_kv_arr.begin(), model.hparams.n_head_kv_arr.end(), (uint32_t) nkv);
            kv.reset(new llama_kv_cache_unified(model.hparams, { [](uint32_t, int) -> ggml_tensor * { return nullptr; } }));
            const bool ok = kv->init(model, cparams, (ggml_type) tk, (ggml_type) tv, (uint32_t) size, false);
            k_row = ggml_row_size((ggml_type) tk, (int64_t) hd*nkv);
            next_tok = 0;
            std::printf(ok ? "ok\n" : "fail\n");
        } else if (op == 'A' || op == 'B') {
            std::vector<llama_seq_id> seq;
            if (op == 'A') { int s, n; in >> s >> n; seq.assign((size_t) n, s); }
            else { int n; in >> n; seq.resize((size_t) n); for (auto & s : seq) in >> s; }
            const uint32_t n = (uint32_t) seq.size();
            std::vector<llama_pos> pos(n); std::vector<int32_t> n_seq_id(n, 1); std::vector<llama_seq_id *> seq_ptr(n);
            std::vector<llama_pos> nxt(64, -2);
            for (uint32_t i = 0; i < n; ++i) {
                llama_pos & p = nxt[(size_t) seq[i]];
                if (p == -2) {                                    // next position: seq_pos_max answers 0 for an empty sequence too (:431-441)
                    p = 0;
                    for (uint32_t c = 0; c < kv->size; ++c) if (kv->cells[c].has_seq_id(seq[i])) p = std::max(p, kv->cells[c].pos + 1);
                }
                pos[i] = p++; seq_ptr[i] = &seq[i];
            }
            llama_ubatch ub = {};
            ub.equal_seqs = false; ub.n_tokens = n; ub.n_seq_tokens = 1; ub.n_seqs = n;      // one token per "sequence slot": interleaving allowed
            ub.pos = pos.data(); ub.n_seq_id = n_seq_id.data(); ub.seq_id = seq_ptr.data();
            const auto slot = kv->find_slot(ub);
            if (!slot) { std::printf("noslot\n"); }
            else {
                for (uint32_t i = 0; i < n; ++i) {                // what the graph's ggml_cpy of k_cur does to layer 0
                    int32_t tok = next_tok++;
                    memcpy((char *) kv->k_l[0]->data + (size_t) (kv->head + i)*k_row, &tok, 4);
                }
                std::printf("slot %u\n", kv->head);
                kv->head += n;
                if (kv->head >= kv->size) kv->head = 0;
            }
        } else if (op == 'M') {
            // one decode step's KV section with the reference's own mask fill: find_slot, kv_self->n as llama_context::decode
            // sets it (src/llama-context.cpp:1304-1312), llm_graph_input_attn_kv_unified::set_input (src/llama-graph.cpp:403-477)
            int n0; in >> n0;
            std::vector<llama_seq_id> seq((size_t) n0); for (auto & q : seq) in >> q;
            const uint32_t n = (uint32_t) n0;
            std::vector<llama_pos> pos(n); std::vector<int32_t> n_seq_id(n, 1); std::vector<llama_seq_id *> seq_ptr(n);
            std::vector<llama_pos> nxt(64, -2);
            for (uint32_t i = 0; i < n; ++i) {
                llama_pos & p = nxt[(size_t) seq[i]];
                if (p == -2) { p = 0; for (uint32_t c = 0; c < kv->size; ++c) if (kv->cells[c].has_seq_id(seq[i])) p = std::max(p, kv->cells[c].pos + 1); }
                pos[i] = p++; seq_ptr[i] = &seq[i];
            }
            llama_ubatch ub = {};
            ub.equal_seqs = false; ub.n_tokens = n; ub.n_seq_tokens = 1; ub.n_seqs = n;
            ub.pos = pos.data(); ub.n_seq_id = n_seq_id.data(); ub.seq_id = seq_ptr.data();
            if (!kv->find_slot(ub)) { std::printf("noslot\n"); std::fflush(stdout); continue; }
            for (uint32_t i = 0; i < n; ++i) { int32_t tok = next_tok++; memcpy((char *) kv->k_l[0]->data + (size_t) (kv->head + i)*k_row, &tok, 4); }
            const uint32_t pad = kv->get_padding(cparams);
            kv->n = std::min(kv->size, std::max(pad, GGML_PAD(kv->cell_max(), pad)));
            ggml_init_params ip = { ggml_tensor_overhead()*4, nullptr, true };
            ggml_context * ctx = ggml_init(ip);
            llm_graph_input_attn_kv_unified inp(model.hparams, cparams, kv.get());
            inp.self_kq_mask = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, kv->n, GGML_PAD(n, GGML_KQ_MASK_PAD));
            ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors_from_buft(ctx, ggml_backend_cpu_buffer_type());
            inp.set_input(&ub);
            const float * m = (const float *) inp.self_kq_mask->data;
            std::printf("m %u", kv->n);
            for (uint32_t i = 0; i < n; ++i) {                     // row i: token ids of the cells its mask row leaves visible, by position
                std::vector<std::pair<llama_pos, int32_t>> v;
                for (uint32_t c = 0; c < kv->n; ++c) if (m[(size_t) i*kv->n + c] == 0.0f) {
                    int32_t tok; memcpy(&tok, (const char *) kv->k_l[0]->data + (size_t) c*k_row, 4);
                    v.emplace_back(kv->cells[c].pos, tok);
                }
                std::sort(v.begin(), v.end());
                std::printf(" ");
                for (size_t k = 0; k < v.size(); ++k) std::printf(k ? ",%d" : "%d", v[k].second);
            }
            std::printf("\n");
            ggml_backend_buffer_free(buf); ggml_free(ctx);
            kv->head += n;
            if (kv->head >= kv->size) kv->head = 0;
        } else if (op == 'C') { int a, b, p0, p1; in >> a >> b >> p0 >> p1; kv->seq_cp(a, b, p0, p1); std::printf("ok\n"); }
        else if (op == 'R') { int s, p0, p1; in >> s >> p0 >> p1; std::printf("%d\n", kv->seq_rm(s, p0, p1) ? 1 : 0); }
        else if (op == 'T') { int s, p0, p1, d; in >> s >> p0 >> p1 >> d; kv->seq_add(s, p0, p1, d); std::printf("ok\n"); }
        else if (op == 'E') {                                      // the cells with a pending K rotation and their delta (has_shift)
            int s; in >> s;
            std::vector<std::pair<llama_pos, llama_pos>> v;
            for (uint32_t i = 0; i < kv->size; ++i) if (kv->cells[i].has_seq_id(s) && kv->cells[i].delta != 0) v.emplace_back(kv->cells[i].pos, kv->cells[i].delta);
            std::sort(v.begin(), v.end());
            std::printf("e %d", kv->has_shift ? 1 : 0);
            for (auto & e : v) std::printf(" %d:%d", e.first, e.second);
            std::printf("\n");
        }
        else if (op == 'K') { int s; in >> s; kv->seq_keep(s); std::printf("ok\n"); }
        else if (op == 'X') { kv->clear(); std::printf("ok\n"); }
        else if (op == 'P') { int s; in >> s; std::printf("%d\n", kv->seq_pos_max(s)); }
        else if (op == 'N') { std::printf("%d %u\n", kv->get_n_tokens(), kv->get_used_cells()); }
        else if (op == 'V') {
            int s; in >> s;
            std::vector<std::pair<llama_pos, int32_t>> v;
            for (uint32_t i = 0; i < kv->size; ++i) if (kv->cells[i].has_seq_id(s)) {
                int32_t tok; memcpy(&tok, (const char *) kv->k_l[0]->data + (size_t) i*k_row, 4);
                v.emplace_back(kv->cells[i].pos, tok);
            }
            std::sort(v.begin(), v.end());
            std::printf("v");
            for (auto & e : v) std::printf(" %d:%d", e.first, e.second);
            std::printf("\n");
        } else if (op == 'S') {
            int s; in >> s;
            mem_writer w; kv->state_write(w, s);
            std::printf("s ");
            for (uint8_t b : w.buf) std::printf("%02x", b);
            std::printf("\n");
        } else if (op == 'L') {
            int s; std::string hex; in >> s >> hex;
            mem_reader r;
            for (size_t i = 0; i + 1 < hex.size(); i += 2) r.buf.push_back((uint8_t) std::stoi(hex.substr(i, 2), nullptr, 16));
            try { kv->state_read(r, s); std::printf("ok\n"); } catch (const std::exception & e) { std::printf("fail %s\n", e.what()); }
        } else { std::printf("?\n"); }
        std::fflush(stdout);
    }
    return 0;
}
