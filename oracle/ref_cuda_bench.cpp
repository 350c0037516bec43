// oracle/ref_cuda_bench.cpp — times GGML_OP_FLASH_ATTN_EXT / GGML_OP_CPY on the REFERENCE's CUDA backend built for the
// B200 (oracle/Makefile target `refcuda`): SURVEY 8(d)'s secondary bar, "the reference's own kernels run on B200".
//
// TEST / MEASUREMENT INFRASTRUCTURE ONLY.  No arithmetic of its own: it builds the graph the reference builds for
// batched decode (src/llama-graph.cpp:1160-1250 build_attn_mha: q permuted [d, n_q, n_head], K/V strided views
// [d, n_kv, n_head_kv] into the flat cell-major cache, F16 mask [n_kv, PAD(n_q, 64)] filled like
// llm_graph_input_attn_kv_unified::set_input :403-477) and runs it through ggml_backend_graph_compute, in the style of
// tests/test-backend-ops.cpp's perf mode (:270-340).  Linked twice by the Makefile: against libggml-ref.so (the
// reference's kernels) and libggml-b200.so (the same backend with the two hot-path calls routed to libb200attn.so) —
// an unmodified ggml graph either way.  The same graph is run on the reference's CPU backend and the NMSE is printed.
//
//   ref_cuda_bench --ctx 1024,1024,... --type f16|q8_0|q4_0 [--heads 32 --kv-heads 8] [--iters 50] [--no-check]
//   ref_cuda_bench --cpy 512 --type q8_0          KV store of 512 tokens x 1024 elements (src/llama-graph.cpp:1375-1397)
#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"
#include "ggml-cpu.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <thread>
#include <vector>

static ggml_type parse_type(const std::string & s) {
    if (s == "f16")  return GGML_TYPE_F16;
    if (s == "q8_0") return GGML_TYPE_Q8_0;
    if (s == "q4_0") return GGML_TYPE_Q4_0;
    fprintf(stderr, "unknown type %s\n", s.c_str()); exit(2);
}

struct fa_graph {
    ggml_context * ctx = nullptr;
    ggml_backend_buffer_t buf = nullptr;
    ggml_tensor * q = nullptr, * k = nullptr, * v = nullptr, * mask = nullptr, * out = nullptr;
    ggml_cgraph * gf = nullptr;
};

static fa_graph build_fa(ggml_backend_t be, int d, int n_q, int n_head, int n_head_kv, int64_t n_kv, ggml_type tk, ggml_type tv) {
    fa_graph g;
    ggml_init_params ip = { ggml_tensor_overhead()*64 + ggml_graph_overhead(), nullptr, /*no_alloc=*/true };
    g.ctx = ggml_init(ip);
    g.q    = ggml_new_tensor_3d(g.ctx, GGML_TYPE_F32, d, n_head, n_q);                  // as the model produces it
    g.k    = ggml_new_tensor_2d(g.ctx, tk, (int64_t) d*n_head_kv, n_kv);                // flat cell-major cache
    g.v    = ggml_new_tensor_2d(g.ctx, tv, (int64_t) d*n_head_kv, n_kv);
    g.mask = ggml_new_tensor_2d(g.ctx, GGML_TYPE_F16, n_kv, GGML_PAD(n_q, GGML_KQ_MASK_PAD));
    ggml_tensor * qp = ggml_permute(g.ctx, g.q, 0, 2, 1, 3);
    ggml_tensor * k3 = ggml_view_3d(g.ctx, g.k, d, n_kv, n_head_kv, g.k->nb[1], ggml_row_size(tk, d), 0);
    ggml_tensor * v3 = ggml_view_3d(g.ctx, g.v, d, n_kv, n_head_kv, g.v->nb[1], ggml_row_size(tv, d), 0);
    g.out = ggml_flash_attn_ext(g.ctx, qp, k3, v3, g.mask, 1.0f/sqrtf((float) d), 0.0f, 0.0f);
    ggml_flash_attn_ext_set_prec(g.out, GGML_PREC_F32);
    g.gf = ggml_new_graph(g.ctx);
    ggml_build_forward_expand(g.gf, g.out);
    g.buf = ggml_backend_alloc_ctx_tensors(g.ctx, be);
    if (!g.buf) { fprintf(stderr, "buffer allocation failed\n"); exit(1); }
    return g;
}

static double nmse(const std::vector<float> & a, const std::vector<float> & b) {
    double num = 0, den = 0;
    for (size_t i = 0; i < a.size(); ++i) { const double e = (double) a[i] - b[i]; num += e*e; den += (double) a[i]*a[i]; }
    return den > 0 ? num/den : num;
}

int main(int argc, char ** argv) {
    std::vector<int> ctx;
    std::string type = "f16";
    int n_head = 32, n_head_kv = 8, d = 128, iters = 50, cpy_tokens = 0;
    bool check = true;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto next = [&]() -> std::string { if (i + 1 >= argc) { fprintf(stderr, "missing value for %s\n", a.c_str()); exit(2); } return argv[++i]; };
        if      (a == "--ctx")      { const std::string s = next(); size_t p = 0; while (p < s.size()) { size_t e = s.find(',', p); if (e == std::string::npos) e = s.size(); ctx.push_back(atoi(s.substr(p, e - p).c_str())); p = e + 1; } }
        else if (a == "--type")     type = next();
        else if (a == "--heads")    n_head = atoi(next().c_str());
        else if (a == "--kv-heads") n_head_kv = atoi(next().c_str());
        else if (a == "--iters")    iters = atoi(next().c_str());
        else if (a == "--cpy")      cpy_tokens = atoi(next().c_str());
        else if (a == "--no-check") check = false;
        else { fprintf(stderr, "unknown argument %s\n", a.c_str()); return 2; }
    }
    const ggml_type t = parse_type(type);
    ggml_time_init();
    ggml_backend_load_all();
    ggml_backend_t gpu = ggml_backend_init_by_type(GGML_BACKEND_DEVICE_TYPE_GPU, nullptr);
    if (!gpu) { printf("{\"impl\": \"reference-cuda\", \"unavailable\": \"no GPU backend device\"}\n"); return 0; }
    ggml_backend_t cpu = ggml_backend_init_by_type(GGML_BACKEND_DEVICE_TYPE_CPU, nullptr);
    std::mt19937 rng(1234);
    std::uniform_real_distribution<float> uni(-1.0f, 1.0f);

    if (cpy_tokens > 0) {                                     // KV store: F32 [1024, n_tok] -> type view of n_tok*1024 elements
        const int64_t row = (int64_t) d*n_head_kv, n = (int64_t) cpy_tokens*row;
        ggml_init_params ip = { ggml_tensor_overhead()*16 + ggml_graph_overhead(), nullptr, true };
        ggml_context * c = ggml_init(ip);
        ggml_tensor * src = ggml_new_tensor_2d(c, GGML_TYPE_F32, row, cpy_tokens);
        ggml_tensor * dst = ggml_new_tensor_1d(c, t, n);
        ggml_tensor * op = ggml_cpy(c, src, ggml_view_1d(c, dst, n, 0));
        ggml_cgraph * gf = ggml_new_graph(c);
        ggml_build_forward_expand(gf, op);
        ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(c, gpu);
        std::vector<float> h(n); for (auto & x : h) x = uni(rng);
        ggml_backend_tensor_set(src, h.data(), 0, n*4);
        for (int i = 0; i < 5; ++i) ggml_backend_graph_compute(gpu, gf);
        std::vector<double> us;
        for (int i = 0; i < iters; ++i) { const int64_t t0 = ggml_time_us(); ggml_backend_graph_compute(gpu, gf); us.push_back((double) (ggml_time_us() - t0)); }
        std::sort(us.begin(), us.end());
        const double med = us[us.size()/2], bytes = (double) n*4 + (double) ggml_row_size(t, n);
        printf("{\"op\": \"CPY\", \"backend\": \"%s\", \"type\": \"%s\", \"n_tokens\": %d, \"us_median\": %.2f, \"us_min\": %.2f, \"GBps\": %.1f}\n",
               ggml_backend_name(gpu), type.c_str(), cpy_tokens, med, us[0], bytes/med*1e-3);
        ggml_backend_buffer_free(buf); ggml_free(c);
        return 0;
    }

    if (ctx.empty()) ctx.assign(8, 1024);
    const int n_q = (int) ctx.size();
    int64_t total = 0; for (int c : ctx) total += c;
    const int64_t n_kv = GGML_PAD(total, 256);                // get_padding() with flash attention (src/llama-kv-cache.cpp:674-677)
    const int64_t row = (int64_t) d*n_head_kv;

    // host data: Q, K, V ~ U(-1, 1) (init_tensor_uniform, tests/test-backend-ops.cpp:39-110), K/V quantised by the reference
    std::vector<float> hq((size_t) n_q*n_head*d); for (auto & x : hq) x = uni(rng);
    std::vector<uint8_t> hk(ggml_row_size(t, row)*n_kv), hv(hk.size());
    {
        std::vector<float> tmp((size_t) row*1024);
        for (int64_t c0 = 0; c0 < n_kv; c0 += 1024) {
            const int64_t nr = std::min<int64_t>(1024, n_kv - c0);
            for (int pass = 0; pass < 2; ++pass) {
                for (int64_t i = 0; i < nr*row; ++i) tmp[i] = uni(rng);
                uint8_t * dstp = (pass ? hv : hk).data() + ggml_row_size(t, row)*c0;
                if (t == GGML_TYPE_F16) ggml_fp32_to_fp16_row(tmp.data(), (ggml_fp16_t *) dstp, nr*row);
                else                    ggml_quantize_chunk(t, tmp.data(), dstp, 0, nr, row, nullptr);
            }
        }
    }
    // mask: row s sees its own contiguous cells (a batch of prompts placed one after the other, then decoded)
    const int64_t n_q_pad = GGML_PAD(n_q, GGML_KQ_MASK_PAD);
    std::vector<ggml_fp16_t> hm((size_t) n_kv*n_q_pad, ggml_fp32_to_fp16(-INFINITY));
    { int64_t off = 0; for (int s = 0; s < n_q; ++s) { for (int64_t c = off; c < off + ctx[s]; ++c) hm[(size_t) s*n_kv + c] = ggml_fp32_to_fp16(0.0f); off += ctx[s]; } }

    fa_graph g = build_fa(gpu, d, n_q, n_head, n_head_kv, n_kv, t, t);
    ggml_backend_tensor_set(g.q, hq.data(), 0, hq.size()*4);
    ggml_backend_tensor_set(g.k, hk.data(), 0, hk.size());
    ggml_backend_tensor_set(g.v, hv.data(), 0, hv.size());
    ggml_backend_tensor_set(g.mask, hm.data(), 0, hm.size()*2);
    if (!ggml_backend_supports_op(gpu, g.out)) { printf("{\"impl\": \"reference-cuda\", \"unavailable\": \"op not supported by the backend\"}\n"); return 0; }

    for (int i = 0; i < 5; ++i) ggml_backend_graph_compute(gpu, g.gf);
    std::vector<double> us;
    for (int i = 0; i < iters; ++i) { const int64_t t0 = ggml_time_us(); ggml_backend_graph_compute(gpu, g.gf); us.push_back((double) (ggml_time_us() - t0)); }
    std::sort(us.begin(), us.end());
    const double med = us[us.size()/2];
    std::vector<float> got((size_t) n_q*n_head*d);
    ggml_backend_tensor_get(g.out, got.data(), 0, got.size()*4);

    double err = -1.0, cpu_us = -1.0;
    if (check && cpu) {
        ggml_backend_cpu_set_n_threads(cpu, (int) std::max(1u, std::min(64u, std::thread::hardware_concurrency())));
        fa_graph c = build_fa(cpu, d, n_q, n_head, n_head_kv, n_kv, t, t);
        ggml_backend_tensor_set(c.q, hq.data(), 0, hq.size()*4);
        ggml_backend_tensor_set(c.k, hk.data(), 0, hk.size());
        ggml_backend_tensor_set(c.v, hv.data(), 0, hv.size());
        ggml_backend_tensor_set(c.mask, hm.data(), 0, hm.size()*2);
        const int64_t t0 = ggml_time_us();
        ggml_backend_graph_compute(cpu, c.gf);
        cpu_us = (double) (ggml_time_us() - t0);
        std::vector<float> want(got.size());
        ggml_backend_tensor_get(c.out, want.data(), 0, want.size()*4);
        err = nmse(want, got);
    }
    // algorithmic bytes: every visible cached element once; the dense form also reads the mask and the
    // other sequences' cells, which is what this number exposes
    const double b_kv = (double) total*2.0*(double) ggml_row_size(t, row);
    printf("{\"op\": \"FLASH_ATTN_EXT\", \"backend\": \"%s\", \"form\": \"dense mask, unmodified ggml graph\", \"type\": \"%s\", \"n_q\": %d, \"n_head\": %d, "
           "\"n_head_kv\": %d, \"n_kv\": %lld, \"visible_tokens\": %lld, \"us_median\": %.2f, \"us_min\": %.2f, \"algorithmic_GBps\": %.1f, "
           "\"nmse_vs_cpu_backend\": %.3e, \"cpu_backend_us\": %.0f}\n",
           ggml_backend_name(gpu), type.c_str(), n_q, n_head, n_head_kv, (long long) n_kv, (long long) total, med, us[0], b_kv/med*1e-3, err, cpu_us);
    return 0;
}
