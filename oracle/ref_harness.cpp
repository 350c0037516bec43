// oracle/ref_harness.cpp — thin C entry points over the REFERENCE's own ggml CPU code.
//
// TEST INFRASTRUCTURE ONLY (see oracle/paged_attn_oracle.h).  This file holds no arithmetic of its
// own: it builds the same ggml graphs the reference builds (tests/test-backend-ops.cpp:3232-3267
// test_flash_attn_ext; src/llama-graph.cpp:1160-1250 build_attn_mha; :1375-1397 KV cpy) and runs them
// with the reference's ggml_graph_compute (ggml/src/ggml-cpu/ggml-cpu.c:15690).  It is compiled
// together with the reference sources where they lie under /root/reference by oracle/Makefile
// (target `ref`), output only into oracle/_ref/.  No reference source is copied into this repo.
#include "ggml.h"
#include "ggml-cpu.h"

#include <algorithm>
#include <chrono>
#include <thread>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

namespace {

struct ctx_guard {
    ggml_context * ctx;
    explicit ctx_guard(size_t mem) {
        ggml_init_params ip = { mem, nullptr, /*no_alloc=*/false };
        ctx = ggml_init(ip);
    }
    ~ctx_guard() { if (ctx) ggml_free(ctx); }
};

// tensor header only; data points at caller memory
ggml_tensor * wrap(ggml_context * ctx, ggml_type type, int64_t ne0, int64_t ne1, int64_t ne2, void * data) {
    ggml_set_no_alloc(ctx, true);
    ggml_tensor * t = ggml_new_tensor_3d(ctx, type, ne0, ne1, ne2);
    ggml_set_no_alloc(ctx, false);
    t->data = data;
    return t;
}

// n_iter >= 3: one warm-up run, then the MEDIAN of the other n_iter - 1 (SURVEY.md 8d "CPU baseline timing": median of
// >= 5 runs after 1 warm-up); fewer: the best of them
double run_graph(ggml_cgraph * gf, int n_threads, int n_iter) {
    ggml_cplan plan = ggml_graph_plan(gf, n_threads, nullptr);
    std::vector<uint8_t> work(plan.work_size + 64);
    plan.work_data = work.data();
    std::vector<double> us_all;
    for (int it = 0; it < (n_iter < 1 ? 1 : n_iter); ++it) {
        auto t0 = std::chrono::steady_clock::now();
        ggml_graph_compute(gf, &plan);
        auto t1 = std::chrono::steady_clock::now();
        us_all.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
    }
    if (us_all.size() >= 3) {
        us_all.erase(us_all.begin());
        std::sort(us_all.begin(), us_all.end());
        return us_all[us_all.size()/2];
    }
    return *std::min_element(us_all.begin(), us_all.end());
}

} // namespace

// ggml fills its fp16 lookup table inside the first ggml_init (ggml/src/ggml.c ggml_init, is_first_call)
static void ref_init_once() {
    static bool done = false;
    if (done) return;
    ggml_init_params ip = { 1u << 20, nullptr, false };
    ggml_context * c = ggml_init(ip);
    if (c) ggml_free(c);
    ggml_cpu_init();
    done = true;
}

extern "C" {

int ref_n_threads_default(void) {
    return (int) std::thread::hardware_concurrency();
}

const char * ref_cpu_features(void) {
    static char buf[256];
    snprintf(buf, sizeof(buf), "avx2=%d fma=%d f16c=%d avx512=%d avx512_vnni=%d avx_vnni=%d",
             ggml_cpu_has_avx2(), ggml_cpu_has_fma(), ggml_cpu_has_f16c(), ggml_cpu_has_avx512(),
             ggml_cpu_has_avx512_vnni(), ggml_cpu_has_avx_vnni());
    return buf;
}

size_t ref_row_size(int type, int64_t n) { return ggml_row_size((ggml_type) type, n); }

// type_traits from_float_ref / to_float (ggml/src/ggml.c:600-720)
int ref_quantize_row_ref(int type, const float * x, void * y, int64_t k) {
    ref_init_once();
    if (type == GGML_TYPE_F16) { ggml_fp32_to_fp16_row(x, (ggml_fp16_t *) y, k); return 0; }
    const ggml_type_traits * tt = ggml_get_type_traits((ggml_type) type);
    if (!tt || !tt->from_float_ref) return -1;
    tt->from_float_ref(x, y, k);
    return 0;
}

// the CPU backend's own (SIMD) quantiser: type_traits_cpu[type].from_float (ggml-cpu.c:253-310)
int ref_quantize_row_cpu(int type, const float * x, void * y, int64_t k) {
    ref_init_once();
    const ggml_type_traits_cpu * tt = ggml_get_type_traits_cpu((ggml_type) type);
    if (!tt || !tt->from_float) return -1;
    tt->from_float(x, y, k);
    return 0;
}

int ref_to_float(int type, const void * x, float * y, int64_t k) {
    ref_init_once();
    const ggml_type_traits * tt = ggml_get_type_traits((ggml_type) type);
    if (!tt || !tt->to_float) return -1;
    tt->to_float(x, y, k);
    return 0;
}

// GGML_OP_CPY F32 [row_elems, n_rows] -> `type`, contiguous dst: the KV write as the graph
// schedules it for n_rows consecutive cells (src/llama-graph.cpp:1375-1379).
int ref_cpy_f32(int type, const float * src, int64_t row_elems, int64_t n_rows, void * dst, int n_threads) {
    ref_init_once();
    ctx_guard g(16u << 20);
    if (!g.ctx) return -1;
    ggml_tensor * a = wrap(g.ctx, GGML_TYPE_F32, row_elems, n_rows, 1, (void *) src);
    ggml_tensor * b = wrap(g.ctx, (ggml_type) type, row_elems, n_rows, 1, dst);
    ggml_tensor * c = ggml_cpy(g.ctx, a, ggml_view_1d(g.ctx, b, row_elems*n_rows, 0));
    ggml_cgraph * gf = ggml_new_graph(g.ctx);
    ggml_build_forward_expand(gf, c);
    run_graph(gf, n_threads, 1);
    return 0;
}

// Attention on the reference CPU backend over a flat token-major cache
//   q    F32  [n_q][n_head][d]                         (permuted to [d, n_q, n_head] as libllama does)
//   k, v type [n_kv][n_head_kv][row(d)]                (viewed [d, n_kv, n_head_kv], nb1 = row(n_head_kv*d))
//   mask F16  [pad64(n_q)][n_kv] or NULL
//   dst  F32  [n_q][n_head][d]
// mode 0: GGML_OP_FLASH_ATTN_EXT, prec F32           (src/llama-graph.cpp:1187-1199)
// mode 1: "F32 CPU result": K/V -> F32 by to_float, mul_mat -> soft_max_ext -> mul_mat (:1201-1239)
// returns through *us the wall time of one graph compute in us: the median of n_iter - 1 runs after one warm-up run
// when n_iter >= 3, else the best of n_iter
int ref_attention(int mode, int n_q, int n_head, int n_head_kv, int d, int64_t n_kv,
                  int type_k, int type_v,
                  const float * q, const void * k, const void * v, const uint16_t * mask,
                  float scale, float max_bias, float softcap,
                  float * dst, int n_threads, int n_iter, double * us) {
    ref_init_once();
    const size_t inter = mode == 1 ? (size_t) n_kv*n_q*n_head*4*3 + (size_t) n_kv*n_head_kv*d*4*2 : 0;
    ctx_guard g((64u << 20) + inter + (size_t) n_q*n_head*d*8);
    if (!g.ctx) return -1;
    ggml_context * ctx = g.ctx;
    const int64_t n_q_pad = GGML_PAD(n_q, GGML_KQ_MASK_PAD);

    ggml_tensor * q0 = wrap(ctx, GGML_TYPE_F32, d, n_head, n_q, (void *) q);
    ggml_tensor * qp = ggml_permute(ctx, q0, 0, 2, 1, 3);                       // [d, n_q, n_head]
    ggml_tensor * m  = mask ? wrap(ctx, GGML_TYPE_F16, n_kv, n_q_pad, 1, (void *) mask) : nullptr;
    ggml_tensor * out = nullptr;
    std::vector<float> k32, v32;

    if (mode == 0) {
        ggml_tensor * kc = wrap(ctx, (ggml_type) type_k, (int64_t) d*n_head_kv, n_kv, 1, (void *) k);
        ggml_tensor * vc = wrap(ctx, (ggml_type) type_v, (int64_t) d*n_head_kv, n_kv, 1, (void *) v);
        ggml_tensor * kv3 = ggml_view_3d(ctx, kc, d, n_kv, n_head_kv, kc->nb[1], ggml_row_size(kc->type, d), 0);
        ggml_tensor * vv3 = ggml_view_3d(ctx, vc, d, n_kv, n_head_kv, vc->nb[1], ggml_row_size(vc->type, d), 0);
        out = ggml_flash_attn_ext(ctx, qp, kv3, vv3, m, scale, max_bias, softcap);
        ggml_flash_attn_ext_set_prec(out, GGML_PREC_F32);
    } else {
        const int64_t row = (int64_t) d*n_head_kv;
        k32.resize((size_t) n_kv*row); v32.resize((size_t) n_kv*row);
        const size_t rbk = ggml_row_size((ggml_type) type_k, row), rbv = ggml_row_size((ggml_type) type_v, row);
        for (int64_t c = 0; c < n_kv; ++c) {
            ref_to_float(type_k, (const char *) k + c*rbk, k32.data() + c*row, row);
            ref_to_float(type_v, (const char *) v + c*rbv, v32.data() + c*row, row);
        }
        ggml_tensor * kc = wrap(ctx, GGML_TYPE_F32, row, n_kv, 1, k32.data());
        ggml_tensor * vc = wrap(ctx, GGML_TYPE_F32, row, n_kv, 1, v32.data());
        ggml_tensor * kv3 = ggml_view_3d(ctx, kc, d, n_kv, n_head_kv, kc->nb[1], (size_t) d*4, 0);
        ggml_tensor * vv3 = ggml_view_3d(ctx, vc, d, n_kv, n_head_kv, vc->nb[1], (size_t) d*4, 0);
        ggml_tensor * kq = ggml_mul_mat(ctx, kv3, ggml_cont(ctx, qp));          // [n_kv, n_q, n_head]
        float sm_scale = scale;
        if (softcap != 0.0f) {
            kq = ggml_scale(ctx, kq, scale/softcap);
            kq = ggml_tanh(ctx, kq);
            kq = ggml_scale(ctx, kq, softcap);
            sm_scale = 1.0f;
        }
        kq = ggml_soft_max_ext(ctx, kq, m, sm_scale, max_bias);
        ggml_tensor * vt = ggml_cont(ctx, ggml_transpose(ctx, vv3));            // [n_kv, d, n_head_kv]
        ggml_tensor * kqv = ggml_mul_mat(ctx, vt, kq);                          // [d, n_q, n_head]
        out = ggml_cont(ctx, ggml_permute(ctx, kqv, 0, 2, 1, 3));               // [d, n_head, n_q]
    }
    ggml_cgraph * gf = ggml_new_graph(ctx);
    ggml_build_forward_expand(gf, out);
    const double t = run_graph(gf, n_threads, n_iter);
    if (us) *us = t;
    memcpy(dst, out->data, (size_t) n_q*n_head*d*sizeof(float));
    return 0;
}

// K-shift exactly as llama_context::build_kv_self_shift / build_rope_shift emit it (src/llama-context.cpp:524-571,
// :447-495): view [head_dim, n_head_kv, n_cells] of the flat K cache; F16/F32: ggml_rope_ext_inplace on the view;
// quantised: ggml_cast -> F32, ggml_rope_ext_inplace, ggml_cpy back into the view.  shift = I32 [n_cells].  In place.
int ref_k_shift(int type, void * k_cache, int64_t n_cells, int n_head_kv, int d, const int32_t * shift,
                int n_rot, int mode, int n_ctx_orig, float freq_base, float freq_scale, float ext_factor,
                float attn_factor, float beta_fast, float beta_slow, const float * freq_factors, int n_threads) {
    ref_init_once();
    const int64_t row = (int64_t) d*n_head_kv;
    ctx_guard g((size_t) 64*1024*1024 + (size_t) n_cells*row*8);
    ggml_context * ctx = g.ctx;
    if (!ctx) return -1;
    ggml_tensor * kc = wrap(ctx, (ggml_type) type, row*n_cells, 1, 1, k_cache);          // k_l[il]: flat 1-D (llama-kv-cache.cpp:100)
    ggml_tensor * sh = wrap(ctx, GGML_TYPE_I32, n_cells, 1, 1, (void *) shift);
    ggml_tensor * ff = freq_factors ? wrap(ctx, GGML_TYPE_F32, n_rot/2, 1, 1, (void *) freq_factors) : nullptr;
    ggml_tensor * cur = ggml_view_3d(ctx, kc, d, n_head_kv, n_cells, ggml_row_size(kc->type, d), ggml_row_size(kc->type, row), 0);
    ggml_tensor * tmp;
    if (ggml_is_quantized(cur->type)) {
        tmp = ggml_cast(ctx, cur, GGML_TYPE_F32);
        tmp = ggml_rope_ext_inplace(ctx, tmp, sh, ff, n_rot, mode, n_ctx_orig, freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow);
        tmp = ggml_cpy(ctx, tmp, cur);
    } else {
        tmp = ggml_rope_ext_inplace(ctx, cur, sh, ff, n_rot, mode, n_ctx_orig, freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow);
    }
    ggml_cgraph * gf = ggml_new_graph(ctx);
    ggml_build_forward_expand(gf, tmp);
    run_graph(gf, n_threads, 1);
    return 0;
}

} // extern "C"
