/*
 * oracle/paged_attn_oracle.c — see paged_attn_oracle.h.  TEST INFRASTRUCTURE ONLY.
 * Plain C restatement of the reference's algorithm for the paged-attention hot path; parity
 * pinned against oracle/_ref (the reference's own ggml CPU code) and tests/golden/.
 */
#include "paged_attn_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ---------------------------------------------------------------- half conversion */

/* ggml/src/ggml-impl.h:317-440 (GGML_FP32_TO_FP16 / GGML_FP16_TO_FP32): IEEE RNE both ways */
uint16_t orc_fp32_to_fp16(float f) {
    uint32_t x; memcpy(&x, &f, 4);
    const uint32_t sign = (x >> 16) & 0x8000u;
    const uint32_t absx = x & 0x7fffffffu;
    if (absx >= 0x7f800000u) {                       /* inf / nan */
        return (uint16_t)(sign | 0x7c00u | ((absx > 0x7f800000u) ? (0x0200u | ((absx & 0x007fffffu) >> 13)) : 0u));  /* NaN quieted, payload kept (vcvtps2ph) */
    }
    if (absx >= 0x477ff000u) {                       /* >= 65520 rounds to inf */
        return (uint16_t)(sign | 0x7c00u);
    }
    if (absx < 0x33000001u) {                        /* <= 2^-25 rounds to zero */
        return (uint16_t)sign;
    }
    int32_t  e = (int32_t)(absx >> 23) - 127;        /* unbiased exponent */
    uint32_t m = (absx & 0x007fffffu) | 0x00800000u; /* 24-bit significand */
    uint32_t h;
    if (e < -14) {                                   /* subnormal half */
        const int shift = (-14 - e) + 13;            /* bits dropped */
        const uint32_t q = m >> shift;
        const uint32_t r = m & ((1u << shift) - 1u);
        const uint32_t half = 1u << (shift - 1);
        h = q + ((r > half) || (r == half && (q & 1u)));
    } else {
        const uint32_t q = m >> 13;                  /* 11 bits incl. hidden */
        const uint32_t r = m & 0x1fffu;
        h = ((uint32_t)(e + 15) << 10) + (q - 0x400u);
        h += ((r > 0x1000u) || (r == 0x1000u && (q & 1u)));  /* carry may bump exponent: correct */
    }
    return (uint16_t)(sign | h);
}

float orc_fp16_to_fp32(uint16_t h) {
    const uint32_t sign = ((uint32_t)h & 0x8000u) << 16;
    const uint32_t e = (h >> 10) & 0x1fu;
    const uint32_t m = h & 0x3ffu;
    uint32_t x;
    if (e == 0) {
        if (m == 0) { x = sign; }
        else {                                        /* subnormal: normalise */
            int sh = 0; uint32_t mm = m;
            while (!(mm & 0x400u)) { mm <<= 1; ++sh; }
            x = sign | ((uint32_t)(127 - 15 - sh + 1) << 23) | ((mm & 0x3ffu) << 13);
        }
    } else if (e == 31) {
        x = sign | 0x7f800000u | (m << 13) | (m ? 0x00400000u : 0u);   /* NaN quieted (vcvtph2ps) */
    } else {
        x = sign | ((e + 112u) << 23) | (m << 13);
    }
    float f; memcpy(&f, &x, 4);
    return f;
}

/* ggml/src/ggml.c:387 ggml_fp32_to_fp16_row */
void orc_fp32_to_fp16_row(const float *x, uint16_t *y, int64_t n) {
    for (int64_t i = 0; i < n; ++i) y[i] = orc_fp32_to_fp16(x[i]);
}

/* ---------------------------------------------------------------- block formats */

/* ggml/src/ggml-common.h:161-166 block_q4_0 = {half d; uint8 qs[16]} (18 B / 32 elems)
 * ggml/src/ggml-common.h:203-208 block_q8_0 = {half d; int8  qs[32]} (34 B / 32 elems) */
#define QK 32
#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif
#define Q4_BYTES 18
#define Q8_BYTES 34

size_t orc_row_size(int type, int64_t n) {
    switch (type) {
        case ORC_TYPE_F32:  return (size_t)n * 4;
        case ORC_TYPE_F16:  return (size_t)n * 2;
        case ORC_TYPE_Q8_0: return (size_t)(n / QK) * Q8_BYTES;
        case ORC_TYPE_Q4_0: return (size_t)(n / QK) * Q4_BYTES;
        default: return 0;
    }
}

/* ggml/src/ggml-quants.c:194-216 quantize_row_q8_0_ref */
void orc_quantize_row_q8_0(const float *x, void *vy, int64_t k) {
    uint8_t *y = (uint8_t *)vy;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; ++i) {
        float amax = 0.0f;
        for (int j = 0; j < QK; ++j) {
            const float v = fabsf(x[i*QK + j]);
            amax = amax > v ? amax : v;              /* MAX(amax, fabsf(v)) */
        }
        const float d  = amax / 127;                 /* ((1 << 7) - 1) */
        const float id = d ? 1.0f/d : 0.0f;
        const uint16_t hd = orc_fp32_to_fp16(d);
        memcpy(y + i*Q8_BYTES, &hd, 2);
        int8_t *qs = (int8_t *)(y + i*Q8_BYTES + 2);
        for (int j = 0; j < QK; ++j) {
            const float x0 = x[i*QK + j]*id;
            qs[j] = (int8_t)roundf(x0);
        }
    }
}

/* ggml/src/ggml-quants.c:31-66 quantize_row_q4_0_ref */
void orc_quantize_row_q4_0(const float *x, void *vy, int64_t k) {
    uint8_t *y = (uint8_t *)vy;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; ++i) {
        float amax = 0.0f, max = 0.0f;
        for (int j = 0; j < QK; ++j) {
            const float v = x[i*QK + j];
            if (amax < fabsf(v)) { amax = fabsf(v); max = v; }
        }
        const float d  = max / -8;
        const float id = d ? 1.0f/d : 0.0f;
        const uint16_t hd = orc_fp32_to_fp16(d);
        memcpy(y + i*Q4_BYTES, &hd, 2);
        uint8_t *qs = y + i*Q4_BYTES + 2;
        for (int j = 0; j < QK/2; ++j) {
            const float x0 = x[i*QK + 0    + j]*id;
            const float x1 = x[i*QK + QK/2 + j]*id;
            const int8_t  t0 = (int8_t)(x0 + 8.5f);
            const int8_t  t1 = (int8_t)(x1 + 8.5f);
            const uint8_t xi0 = (uint8_t)(t0 < 15 ? t0 : 15);   /* MIN(15, (int8_t)(x0 + 8.5f)) */
            const uint8_t xi1 = (uint8_t)(t1 < 15 ? t1 : 15);
            qs[j] = (uint8_t)(xi0 | (xi1 << 4));
        }
    }
}

/* ggml/src/ggml-quants.c:349-362 dequantize_row_q8_0 */
void orc_dequantize_row_q8_0(const void *vx, float *y, int64_t k) {
    const uint8_t *x = (const uint8_t *)vx;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; ++i) {
        uint16_t hd; memcpy(&hd, x + i*Q8_BYTES, 2);
        const float d = orc_fp16_to_fp32(hd);
        const int8_t *qs = (const int8_t *)(x + i*Q8_BYTES + 2);
        for (int j = 0; j < QK; ++j) y[i*QK + j] = qs[j]*d;
    }
}

/* ggml/src/ggml-quants.c:255-273 dequantize_row_q4_0 */
void orc_dequantize_row_q4_0(const void *vx, float *y, int64_t k) {
    const uint8_t *x = (const uint8_t *)vx;
    const int64_t nb = k / QK;
    for (int64_t i = 0; i < nb; ++i) {
        uint16_t hd; memcpy(&hd, x + i*Q4_BYTES, 2);
        const float d = orc_fp16_to_fp32(hd);
        const uint8_t *qs = x + i*Q4_BYTES + 2;
        for (int j = 0; j < QK/2; ++j) {
            const int x0 = (qs[j] & 0x0F) - 8;
            const int x1 = (qs[j] >>   4) - 8;
            y[i*QK + j       ] = x0*d;
            y[i*QK + j + QK/2] = x1*d;
        }
    }
}

int orc_from_float(int type, const float *x, void *y, int64_t k) {
    switch (type) {
        case ORC_TYPE_F32:  memcpy(y, x, (size_t)k*4); return 0;
        case ORC_TYPE_F16:  orc_fp32_to_fp16_row(x, (uint16_t *)y, k); return 0;
        case ORC_TYPE_Q8_0: if (k % QK) return -1; orc_quantize_row_q8_0(x, y, k); return 0;
        case ORC_TYPE_Q4_0: if (k % QK) return -1; orc_quantize_row_q4_0(x, y, k); return 0;
        default: return -1;
    }
}

int orc_to_float(int type, const void *x, float *y, int64_t k) {
    switch (type) {
        case ORC_TYPE_F32:  memcpy(y, x, (size_t)k*4); return 0;
        case ORC_TYPE_F16:  for (int64_t i = 0; i < k; ++i) y[i] = orc_fp16_to_fp32(((const uint16_t *)x)[i]); return 0;
        case ORC_TYPE_Q8_0: if (k % QK) return -1; orc_dequantize_row_q8_0(x, y, k); return 0;
        case ORC_TYPE_Q4_0: if (k % QK) return -1; orc_dequantize_row_q4_0(x, y, k); return 0;
        default: return -1;
    }
}

/* ---------------------------------------------------------------- KV write / block copy */

/* src/llama-graph.cpp:1375-1397 + ggml-cpu.c:3742-3817 (dup_f32, quantise branch: from_float per row).
 * NOTE: the reference CPU backend quantises Q8_0 with its SIMD quantize_row_q8_0
 * (ggml-cpu-quants.c:738-), whose AVX path multiplies by 127/amax and rounds half-to-even; the
 * deterministic definition ("reference implementation for deterministic creation of model files",
 * ggml-quants.c:193) is the _ref one restated here.  tests/test_oracle_vs_ref.py counts the codes
 * on which the two disagree (ties / 1-ulp cases, SURVEY.md §4 test_cpy row). */
int orc_kv_write(const float *src, size_t src_nb1, int64_t n_tokens,
                 const int32_t *slot_mapping,
                 void *cache, int type, int n_head_kv, int head_dim,
                 size_t nb_token, size_t nb_head, int64_t n_cells) {
    if (orc_row_size(type, head_dim) == 0) return -1;
    if ((type == ORC_TYPE_Q8_0 || type == ORC_TYPE_Q4_0) && head_dim % QK) return -1;
    for (int64_t t = 0; t < n_tokens; ++t) {
        const int64_t c = slot_mapping ? slot_mapping[t] : t;
        if (c < 0 || c >= n_cells) continue;
        const float *row = (const float *)((const char *)src + t*src_nb1);
        for (int h = 0; h < n_head_kv; ++h) {
            void *dst = (char *)cache + (size_t)c*nb_token + (size_t)h*nb_head;
            if (orc_from_float(type, row + (size_t)h*head_dim, dst, head_dim)) return -1;
        }
    }
    return 0;
}

/* src/llama-context.cpp:653-711 + ggml-cpu.c:4055-4200 (dup_bytes); pairs applied in order */
int orc_block_copy(void *cache, size_t block_bytes, int64_t n_blocks,
                   const int32_t *pairs, int64_t n_pairs) {
    for (int64_t i = 0; i < n_pairs; ++i) {
        const int64_t s = pairs[2*i], d = pairs[2*i + 1];
        if (s < 0 || d < 0 || s >= n_blocks || d >= n_blocks) return -1;
        if (s == d) continue;
        memcpy((char *)cache + (size_t)d*block_bytes, (const char *)cache + (size_t)s*block_bytes, block_bytes);
    }
    return 0;
}

/* ---------------------------------------------------------------- K-shift (rope on cached K rows) */

/* ggml/src/ggml.c:3735-3750 */
static float rope_yarn_corr_dim(int n_dims, int n_ctx_orig, float n_rot, float base) {
    return n_dims * logf(n_ctx_orig / (n_rot * 2 * (float)M_PI)) / (2 * logf(base));
}
static void rope_yarn_corr_dims(int n_dims, int n_ctx_orig, float freq_base, float beta_fast, float beta_slow, float dims[2]) {
    const float start = floorf(rope_yarn_corr_dim(n_dims, n_ctx_orig, beta_fast, freq_base));
    const float end   =  ceilf(rope_yarn_corr_dim(n_dims, n_ctx_orig, beta_slow, freq_base));
    dims[0] = start > 0 ? start : 0;
    dims[1] = end < n_dims - 1 ? end : n_dims - 1;
}

/* ggml-cpu.c:10571-10611: cache[i0] = cos, cache[i0+1] = sin for the pair starting at i0 (forward: sin_sign = 1) */
static void rope_cache_init(float theta_base, float freq_scale, const float *freq_factors, const float corr_dims[2],
                            int64_t ne0, float ext_factor, float mscale0, float *cache, float theta_scale) {
    float theta = theta_base;
    for (int64_t i0 = 0; i0 < ne0; i0 += 2) {
        const float ff = freq_factors ? freq_factors[i0/2] : 1.0f;
        const float theta_extrap = theta/ff;
        const float theta_interp = freq_scale * theta_extrap;
        float th = theta_interp, mscale = mscale0;
        if (ext_factor != 0.0f) {
            const float hl = corr_dims[1] - corr_dims[0];
            const float y = (i0 / 2 - corr_dims[0]) / (0.001f > hl ? 0.001f : hl);
            const float cl = y < 0 ? 0 : y;
            const float ramp_mix = (1 - (1 < cl ? 1 : cl)) * ext_factor;
            th = theta_interp * (1 - ramp_mix) + theta_extrap * ramp_mix;
            mscale *= 1.0f + 0.1f * logf(1.0f / freq_scale);
        }
        cache[i0 + 0] = cosf(th) * mscale;
        cache[i0 + 1] = sinf(th) * mscale;
        theta *= theta_scale;
    }
}

int orc_k_shift(void *cache, int type, int n_head_kv, int head_dim, size_t nb_token, size_t nb_head, int64_t n_cells,
                const int32_t *shift, const int32_t *cells, int64_t n_rows,
                int n_rot, int mode, int n_ctx_orig, float freq_base, float freq_scale, float ext_factor,
                float attn_factor, float beta_fast, float beta_slow, const float *freq_factors) {
    if (orc_row_size(type, head_dim) == 0 || (mode & ~2) || n_rot <= 0 || (n_rot & 1) || n_rot > head_dim || (head_dim & 1)) return -1;
    if ((type == ORC_TYPE_Q8_0 || type == ORC_TYPE_Q4_0) && head_dim % QK) return -1;
    const float theta_scale = powf(freq_base, -2.0f/n_rot);                       /* ggml-cpu.c:10720 */
    float corr_dims[2];
    rope_yarn_corr_dims(n_rot, n_ctx_orig, freq_base, beta_fast, beta_slow, corr_dims);
    const int neox = (mode & 2) != 0;
    float *x = (float *)malloc(sizeof(float)*(size_t)head_dim), *cs = (float *)malloc(sizeof(float)*(size_t)head_dim);
    if (!x || !cs) { free(x); free(cs); return -1; }
    const int64_t n = cells ? n_rows : n_cells;
    for (int64_t r = 0; r < n; ++r) {
        const int64_t c = cells ? cells[r] : r;
        if (c < 0 || c >= n_cells) continue;
        rope_cache_init((float)(int64_t)shift[c], freq_scale, freq_factors, corr_dims, head_dim, ext_factor, attn_factor, cs, theta_scale);
        for (int h = 0; h < n_head_kv; ++h) {
            void *row = (char *)cache + (size_t)c*nb_token + (size_t)h*nb_head;
            if (orc_to_float(type, row, x, head_dim)) { free(x); free(cs); return -1; }   /* ggml_cast / FP16_TO_FP32 */
            for (int i0 = 0; i0 < n_rot; i0 += 2) {                                        /* ggml-cpu.c:10800-10830 */
                const float cos_theta = cs[i0], sin_theta = cs[i0 + 1];
                const int a = neox ? i0/2 : i0, b = neox ? i0/2 + n_rot/2 : i0 + 1;
                const float x0 = x[a], x1 = x[b];
                x[a] = x0*cos_theta - x1*sin_theta;
                x[b] = x0*sin_theta + x1*cos_theta;
            }
            if (orc_from_float(type, x, row, head_dim)) { free(x); free(cs); return -1; }   /* ggml_cpy back / FP32_TO_FP16 */
        }
    }
    free(x); free(cs);
    return 0;
}

/* ---------------------------------------------------------------- attention */

/* ALiBi slope: ggml-cpu.c:12294-12298, :12316 */
static float alibi_slope(float max_bias, uint32_t n_head, uint32_t h) {
    if (!(max_bias > 0.0f)) return 1.0f;
    const uint32_t n_head_log2 = 1u << (uint32_t)floor(log2((double)n_head));
    const float m0 = powf(2.0f, -(max_bias       ) / n_head_log2);
    const float m1 = powf(2.0f, -(max_bias / 2.0f) / n_head_log2);
    return h < n_head_log2 ? powf(m0, (float)(h + 1)) : powf(m1, (float)(2*(h - n_head_log2) + 1));
}

/* one (query row, head): ggml-cpu.c:12309-12431, online softmax replaced by the equivalent
 * two-pass form in double (exact same quantity, fewer roundings). */
int orc_attention(const orc_attn_params *p) {
    const int D = p->head_dim;
    if (p->n_head % p->n_head_kv) return -1;
    if (orc_row_size(p->type_k, D) == 0 || orc_row_size(p->type_v, D) == 0) return -1;
    const int rk2 = p->n_head / p->n_head_kv;       /* broadcast factor, ggml-cpu.c:12264 */
    float scale = p->scale;
    if (p->logit_softcap != 0.0f) scale /= p->logit_softcap;   /* :12290-12292 */

    float  *krow = (float  *)malloc(sizeof(float)*D);
    float  *vrow = (float  *)malloc(sizeof(float)*D);
    double *acc  = (double *)malloc(sizeof(double)*D);
    int64_t cap = 0; double *sc = NULL; int64_t *cell = NULL;
    int rc = 0;

    for (int iq = 0; iq < p->n_q && !rc; ++iq) {
        /* visible cells of this query row */
        int64_t n_vis;
        int seq = 0;
        if (p->block_table) {
            seq = p->q_seq ? p->q_seq[iq] : iq;
            if (seq < 0 || seq >= p->n_seqs) { rc = -2; break; }
            n_vis = p->seq_lens[seq];
            if (p->q_pos && (int64_t)p->q_pos[iq] + 1 < n_vis) n_vis = (int64_t)p->q_pos[iq] + 1;
            if (n_vis < 0) n_vis = 0;
        } else {
            n_vis = p->n_kv;
        }
        if (n_vis > cap) {
            cap = n_vis;
            sc   = (double  *)realloc(sc,   sizeof(double)*(size_t)cap);
            cell = (int64_t *)realloc(cell, sizeof(int64_t)*(size_t)cap);
        }
        for (int h = 0; h < p->n_head; ++h) {
            const int g = h / rk2;                   /* ik2 = iq2 / rk2, :12336 */
            const float slope = alibi_slope(p->max_bias, (uint32_t)p->n_head, (uint32_t)h);
            const float *q = (const float *)((const char *)p->q + (size_t)iq*p->nbq1 + (size_t)h*p->nbq2);
            int64_t n = 0; double M = -INFINITY;
            for (int64_t ic = 0; ic < n_vis; ++ic) {
                int64_t c; float mv = 0.0f;
                if (p->block_table) {
                    const int32_t b = p->block_table[(size_t)seq*p->bt_stride + ic/p->block_size];
                    c = (int64_t)b*p->block_size + ic % p->block_size;
                } else {
                    c = ic;
                    if (p->mask) {
                        const uint16_t *mp = (const uint16_t *)((const char *)p->mask + (size_t)iq*p->nb_mask1);
                        mv = slope*orc_fp16_to_fp32(mp[ic]);     /* :12349 */
                        if (mv == -INFINITY) continue;           /* :12350 */
                    }
                }
                orc_to_float(p->type_k, (const char *)p->k + (size_t)c*p->nbk1 + (size_t)g*p->nbk2, krow, D);
                double s = 0.0;
                for (int d = 0; d < D; ++d) s += (double)q[d]*(double)krow[d];
                s *= scale;                                       /* :12359 */
                if (p->logit_softcap != 0.0f) s = p->logit_softcap*tanh(s);   /* :12361-12363 */
                s += mv;                                          /* :12365 */
                sc[n] = s; cell[n] = c; ++n;
                if (s > M) M = s;
            }
            for (int d = 0; d < D; ++d) acc[d] = 0.0;
            double S = 0.0;
            for (int64_t i = 0; i < n; ++i) {
                const double w = exp(sc[i] - M);                  /* vs = expf(s - M), :12384 */
                S += w;
                orc_to_float(p->type_v, (const char *)p->v + (size_t)cell[i]*p->nbv1 + (size_t)g*p->nbv2, vrow, D);
                for (int d = 0; d < D; ++d) acc[d] += w*(double)vrow[d];
            }
            /* dst [d, n_head, n_q] contiguous: permute(0,2,1,3) store, :12429-12430.
             * S == 0 (no visible cell) gives 0/0 = NaN in the reference (1.0f/S, :12418); kept. */
            float *o = p->dst + ((size_t)iq*p->n_head + h)*D;
            for (int d = 0; d < D; ++d) o[d] = (float)(acc[d]*(1.0/S));
        }
    }
    free(krow); free(vrow); free(acc); free(sc); free(cell);
    return rc;
}

/* ---------------------------------------------------------------- metrics */

/* tests/test-backend-ops.cpp:177-190 (a = reference) */
double orc_nmse(const float *ref, const float *out, size_t n) {
    double mse_a_b = 0.0, mse_a_0 = 0.0;
    for (size_t i = 0; i < n; ++i) {
        const double a = ref[i], b = out[i];
        mse_a_b += (a - b)*(a - b);
        mse_a_0 += a*a;
    }
    return mse_a_b / mse_a_0;
}

double orc_max_rel_err(const float *ref, const float *out, size_t n, size_t row) {
    double worst = 0.0;
    for (size_t r0 = 0; r0 < n; r0 += row) {
        const size_t m = (n - r0 < row) ? n - r0 : row;
        double ss = 0.0;
        for (size_t i = 0; i < m; ++i) ss += (double)ref[r0+i]*(double)ref[r0+i];
        const double rms = sqrt(ss/(double)m);
        for (size_t i = 0; i < m; ++i) {
            const double a = ref[r0+i], b = out[r0+i];
            if (isnan(a) && isnan(b)) continue;
            if (isnan(a) != isnan(b)) return INFINITY;
            double den = fabs(a) > rms ? fabs(a) : rms;
            if (den == 0.0) den = 1.0;
            const double e = fabs(a - b)/den;
            if (e > worst) worst = e;
        }
    }
    return worst;
}
